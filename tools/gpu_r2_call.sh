#!/bin/bash
# round 2 GPU call: packed spans in the head-size-64 kernel -- attention self-tests, parity tests, A/B on the variable-length zh workload
mkdir -p gpurun_out
echo "== attention self-tests"; timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention" > gpurun_out/r2_pack64_attn.log 2>&1; rc=$?; tail -3 gpurun_out/r2_pack64_attn.log | cut -c1-300
[ $rc -ne 0 ] && exit 1
echo "== rest of the GPU suite"; timeout 1500 python -m pytest tests -m gpu -x -q -k "not fused_attention" > gpurun_out/r2_pack64_pytest.log 2>&1; tail -4 gpurun_out/r2_pack64_pytest.log | cut -c1-300
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as ge; ge.smoke()" 2>&1 | grep -v "^bert_load" | tail -2
for pk in 0 1; do
for w in bge-small-zh-q4_1-b1024-var bge-base-q4_0-b512-s512; do
  BERT_ATTN_PACK=$pk timeout 300 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/bench_r2_pack${pk}_$w.json
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_pack${pk}_$w.json')); print('pack=$pk $w', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), d['roofline']['by_kernel_ms'])
PY
done; done
