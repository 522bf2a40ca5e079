#!/bin/bash
# round 2 GPU call: host-runtime changes (smaller tokenizer work items for long texts, text pointer table in numpy) and the
# removal of the mma.sync attention kernel: tests that go through bert_encode_batch, traces, bench lines
mkdir -p gpurun_out
echo "== tests"; timeout 900 python -m pytest tests/test_gpu.py tests/test_examples.py -m gpu -x -q -k "not perturbed and not gemm and not config" 2>&1 | tail -4 | cut -c1-300
for w in bge-base-q4_0-b512-s512 minilm-q4_0-b1738-var; do
  timeout 200 python tools/e2e_trace.py $w > gpurun_out/r2c_trace_$w.log 2>&1; echo "== $w"; grep -E "chunk_ready|encode_end|python side|device-resident" gpurun_out/r2c_trace_$w.log | head -8 | cut -c1-220
done
for w in bge-base-q4_0-b512-s512 minilm-q4_0-b1738-var bge-small-f16-b256-s128 bge-small-zh-q4_1-b1024-var; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c_bench_$w.json
  python - <<PY
import json
d=json.load(open('gpurun_out/r2c_bench_$w.json')); print('$w', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['value']/d['value'],3), d['roofline']['by_kernel_ms'])
PY
done
