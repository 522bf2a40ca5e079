#!/bin/bash
# round 2, GPU call 10: no-alloc quantiser entry point (alone in a fresh process, then at the end of the whole suite),
# small workloads after the chunk-rule change, latency path on one stream
mkdir -p gpurun_out
echo "== no-alloc alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4 | cut -c1-200
echo "== whole suite"; timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | cut -c1-200
cp gpurun_out/parity_r2.jsonl gpurun_out/r2c10_parity.jsonl 2>/dev/null
echo "== latency"; timeout 300 python tools/latency.py 2>&1 | tee gpurun_out/r2c10_latency.log | head -8
for w in bge-base-q4_0-b512-s512 bge-small-f16-b256-s128 minilm-q4_0-b1738-var bge-small-zh-q4_1-b1024-var; do timeout 600 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c10_bench_$w.json; done
timeout 300 python tools/e2e_trace.py minilm-q4_0-b1738-var > gpurun_out/r2c10_trace_minilm.log 2>&1; tail -30 gpurun_out/r2c10_trace_minilm.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2c10_bench_*.json')):
    try:
        d=json.load(open(f)); print(f.split('bench_')[1][:-5], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['value']/d['value'],3), d['roofline']['by_kernel_ms'])
    except Exception as ex: print(f, ex)
PY
