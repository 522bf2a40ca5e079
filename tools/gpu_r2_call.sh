#!/bin/bash
# round 2 GPU call: fused-LayerNorm epilogue with the second residual box prefetched into the output staging buffer (BERT_LN_RES2),
# self-tests, kernel-level A/B against the round-1 order (variant build res2off), determinism at full size, default bench
mkdir -p gpurun_out
echo "== GEMM self-tests"; timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "gemm or full_size or golden" > gpurun_out/r2_res2_pytest.log 2>&1; rc=$?; tail -3 gpurun_out/r2_res2_pytest.log | cut -c1-300
[ $rc -ne 0 ] && exit 1
for rep in 1 2; do for v in "" res2off; do
  if [ -n "$v" ]; then export BERT_B200_LIB=$PWD/embeddings.cpp_b200/lib/variants/libbert_$v.so; else unset BERT_B200_LIB; fi
  timeout 120 python - <<PY
import __graft_entry__ as ge
pkg = ge.load_package()
out = []
for name, N, K, wt in (("o+LN", 768, 768, 2), ("ffo+LN", 768, 3072, 2), ("o+LN f16", 768, 768, 1), ("o+LN large", 1024, 1024, 2)):
    ms, tf = pkg.bench_gemm(262144, N, K, wt, 29, reps=20)
    out.append(f"{name} {ms:.4f} ms {tf:.0f} TF")
print("variant '$v':", " | ".join(out))
PY
done; done
unset BERT_B200_LIB
timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/bench_r2_res2.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_res2.json')); print('default', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), d['roofline']['by_kernel_ms'])
PY
