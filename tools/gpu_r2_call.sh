#!/bin/bash
# round 2 GPU call: whole GPU suite + smoke after the packed-span attention, final bench lines of the small-model workloads
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r2_final_pytest.log 2>&1; tail -4 gpurun_out/r2_final_pytest.log | cut -c1-300
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as ge; ge.smoke()" 2>&1 | grep -v "^bert_load" | tail -2
for w in minilm-q4_0-b1738-var bge-small-f16-b256-s128 bge-small-zh-q4_1-b1024-var bge-base-q4_0-b512-s512; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/bench_r2_final_$w.json
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_final_$w.json')); print('$w', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['value']/d['value'],3), 'frac', round(d['roofline']['frac'],3), d['roofline']['by_kernel_ms'])
PY
done
