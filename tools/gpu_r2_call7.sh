#!/bin/bash
# round 2, GPU call 7: new host runtime (persistent workers / 3 streams) + widened parity + all bench workloads
mkdir -p gpurun_out; rm -f gpurun_out/parity_r2.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2c7_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2c7_pytest.log | cut -c1-200
cat gpurun_out/parity_r2.jsonl
timeout 600 python bench.py > gpurun_out/r2c7_bench_default.out 2> gpurun_out/r2c7_bench_default.err; echo "bench rc=$?"
for w in minilm-q4_0-b1738-var bge-small-f16-b256-s128 bge-small-zh-q4_1-b1024-var bge-large-q4_0-b1024-s256 bge-base-f16-b512-s512; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline > gpurun_out/r2c7_bench_$w.out 2> gpurun_out/r2c7_bench_$w.err; echo "bench $w rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2c7_bench_*.out')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f.split('bench_')[1][:-4], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ratio', round(d['e2e']['value']/d['value'],3), 'gemmTF', round(d['roofline']['achieved'],1), d['roofline']['by_kernel_ms'], d['host'], d['config'].get('padded_vs_valid_tokens'))
    except Exception as e:
        print(f, 'FAILED', e)
PY
timeout 300 python tools/latency.py > gpurun_out/r2c7_latency.log 2>&1; tail -8 gpurun_out/r2c7_latency.log
