#!/bin/bash
# round 2 GPU call: specialised embedding + LayerNorm kernel (table type / full-row templates), ncu of it and of the GEMMs at the
# 262 144-token micro-batch (dram traffic per launch for bench.py's roofline.traffic)
mkdir -p gpurun_out
echo "== parity"; timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "zero_layer or golden or full_size or staged" > gpurun_out/r2_embed2_pytest.log 2>&1; rc=$?; tail -3 gpurun_out/r2_embed2_pytest.log | cut -c1-300
[ $rc -ne 0 ] && exit 1
for w in bge-base-q4_0-b512-s512 minilm-q4_0-b1738-var bge-small-f16-b256-s128; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/bench_r2_embed2_$w.json
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_embed2_$w.json')); print('$w', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['by_kernel_ms'], d['roofline'].get('hbm_bound_kernels',{}).get('embed_ln'))
PY
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"embed_ln_kernel" -s 6 -c 1 -o gpurun_out/r2_embed2_ncu -f python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/r2_embed2_ncu.log 2>&1; tail -1 gpurun_out/r2_embed2_ncu.log | cut -c1-200
timeout 600 ncu --set full --clock-control none -k regex:"gemm_tc_kernel" -s 196 -c 8 -o gpurun_out/r2_gemm262k_ncu -f python bench.py --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/r2_gemm262k_ncu.log 2>&1; tail -1 gpurun_out/r2_gemm262k_ncu.log | cut -c1-200
