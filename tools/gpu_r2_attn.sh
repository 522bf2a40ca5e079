#!/bin/bash
# round 2: attention_tc.cu with the per-item output tail moved to its own warpgroup, against the round-1 build of the same file
mkdir -p gpurun_out
V=embeddings.cpp_b200/lib/variants
echo "== attention self-tests (tc1)"; timeout 200 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention and round1" > gpurun_out/st.log 2>&1; rc=$?; tail -4 gpurun_out/st.log | cut -c1-300
if [ $rc -ne 0 ]; then echo "self-tests failed: stopping"; exit 1; fi
for round in 1 2; do
  echo "== tc1 round-1 build (round $round)"; BERT_B200_LIB=$PWD/$V/libbert_tc1old.so BERT_ATTN_IMPL=tc1 timeout 60 python tools/attn_bench.py 2>&1 | grep "^attention"
  echo "== tc1 with the output warpgroup (round $round)"; BERT_ATTN_IMPL=tc1 timeout 60 python tools/attn_bench.py 2>&1 | grep "^attention"
done
echo "== e2e parity subset (tc1)"; BERT_ATTN_IMPL=tc1 timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "golden or edge or full_size_properties" 2>&1 | tail -4 | cut -c1-300
echo "== bench (tc1)"; BERT_ATTN_IMPL=tc1 timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2attn_bench_default.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2attn_bench_default.json')); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['by_kernel_ms'], d['clocks'])
PY
