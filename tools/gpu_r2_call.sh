#!/bin/bash
# round 2 GPU call: the variable-length zh workload end to end after the tokenizer change
mkdir -p gpurun_out
timeout 50 python bench.py --workload bge-small-zh-q4_1-b1024-var --no-cpu-baseline --steps 10 --warmup 3 2>/dev/null | grep '^{' > gpurun_out/bench_r2_tok_zh.json
python - <<PY
import json
d=json.load(open('gpurun_out/bench_r2_tok_zh.json')); print('zh', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d.get('host'))
PY
