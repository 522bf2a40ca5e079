#!/bin/bash
# round 2 GPU call: headline bench against the micro-batch size (BERT_MICROBATCH_TOKENS; default 131072), two alternations
mkdir -p gpurun_out
for round in 1 2; do
  for mb in 131072 262144 65536 32768 16384; do
    BERT_MICROBATCH_TOKENS=$mb timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c_bench_mb$mb.json; python - <<PY
import json
d=json.load(open('gpurun_out/r2c_bench_mb$mb.json')); print('mb $mb', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['by_kernel_ms'], d['clocks']['sm_mhz'])
PY
  done
done
