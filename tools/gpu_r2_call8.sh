#!/bin/bash
# round 2, GPU call 8: new attention kernel (column-split softmax, head size 32 on tcgen05): self-tests, A/B timing, e2e traces
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention" 2>&1 | tail -15 | cut -c1-200
echo "== tc2"; timeout 120 python tools/attn_bench.py 2>&1 | tail -6
echo "== tc1"; BERT_ATTN_IMPL=tc1 timeout 120 python tools/attn_bench.py 2>&1 | tail -6
echo "== tc2 d32"; timeout 120 python tools/attn_bench.py 512 256 12 32 2>&1 | tail -6
echo "== mma d32"; BERT_ATTN_IMPL=tc1 timeout 120 python tools/attn_bench.py 512 256 12 32 2>&1 | tail -6
timeout 900 python -m pytest tests/test_gpu.py -m gpu -x -q -k "golden or edge or named_model or full_size_properties or zero_layer or staged" 2>&1 | tail -8 | cut -c1-200
timeout 300 python tools/e2e_trace.py minilm-q4_0-b1738-var > gpurun_out/r2c8_trace_minilm.log 2>&1; tail -40 gpurun_out/r2c8_trace_minilm.log
timeout 300 python tools/e2e_trace.py bge-small-f16-b256-s128 > gpurun_out/r2c8_trace_small.log 2>&1; tail -30 gpurun_out/r2c8_trace_small.log
timeout 300 python tools/latency.py 2>&1 | head -4
for w in bge-base-q4_0-b512-s512 bge-small-f16-b256-s128 minilm-q4_0-b1738-var; do timeout 600 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c8_bench_$w.json; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2c8_bench_*.json')):
    d=json.load(open(f)); print(f.split('bench_')[1][:-5], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['value']/d['value'],3), d['roofline']['by_kernel_ms'])
PY
