#!/bin/bash
# round 2, validation of the final state: the whole GPU suite, smoke(), bench lines with the CPU baseline, ncu launch list
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/r2_final_pytest.log 2>&1; tail -5 gpurun_out/r2_final_pytest.log | cut -c1-300
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as ge; ge.smoke()" 2>&1 | grep -v "^bert_load" | tail -2
echo "== bench default"; timeout 600 python bench.py 2>gpurun_out/r2_final_bench.err | grep '^{' > gpurun_out/bench_r2_final.json
echo "== bench bge-large"; timeout 600 python bench.py --workload bge-large-q4_0-b1024-s256 2>/dev/null | grep '^{' > gpurun_out/bench_r2_final_bge-large-q4_0-b1024-s256.json
python - <<'PY'
import json
for f in ('gpurun_out/bench_r2_final.json','gpurun_out/bench_r2_final_bge-large-q4_0-b1024-s256.json'):
    d=json.load(open(f)); print(d['config']['workload'], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), d['roofline']['by_kernel_ms'], d['clocks'], d.get('cpu_baseline'))
PY
echo "== ncu launch list"; timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches_r2_final.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2_final_ncu.log 2>&1; tail -2 gpurun_out/r2_final_ncu.log | cut -c1-200; wc -l gpurun_out/launches_r2_final.csv
