#!/bin/bash
# round 2 GPU call: fused LayerNorm for N = 384 (three 128-wide column blocks): self-tests, parity, A/B bench of the H = 384 models
mkdir -p gpurun_out
echo "== LN self-tests"; timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_layernorm" > gpurun_out/st.log 2>&1; rc=$?; tail -3 gpurun_out/st.log | cut -c1-300
if [ $rc -ne 0 ]; then echo "self-tests failed: stopping"; exit 1; fi
echo "== parity"; timeout 600 python -m pytest tests/test_gpu.py -m gpu -x -q -k "golden or named_model or baseline_configs or edge" 2>&1 | tail -3 | cut -c1-300
for round in 1 2; do
  for w in minilm-q4_0-b1738-var bge-small-f16-b256-s128; do
    for on in 0 1; do
      BERT_LN_BN128=$on timeout 200 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c_bench_ln$on_$w.json; python - <<PY
import json
d=json.load(open('gpurun_out/r2c_bench_ln$on_$w.json')); print('$w fused=$on', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['by_kernel_ms'])
PY
    done
  done
done
