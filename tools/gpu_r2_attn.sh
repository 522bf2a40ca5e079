#!/bin/bash
# round 2: attention_tc.cu with the slimmed MMA-issuer loop (default) against the round-1 loop (variant mmaslim0)
mkdir -p gpurun_out
V=embeddings.cpp_b200/lib/variants
echo "== attention self-tests"; timeout 200 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention" > gpurun_out/st.log 2>&1; rc=$?; tail -3 gpurun_out/st.log | cut -c1-300
if [ $rc -ne 0 ]; then echo "self-tests failed: stopping"; exit 1; fi
for round in 1 2 3; do
  echo "== round-1 MMA loop (round $round)"; BERT_B200_LIB=$PWD/$V/libbert_mmaslim0.so timeout 60 python tools/attn_bench.py 2>&1 | grep "^attention"
  echo "== slim MMA loop (round $round)"; timeout 60 python tools/attn_bench.py 2>&1 | grep "^attention"
done
for round in 1 2; do
  echo "== head size 32, round-1 descriptors (round $round)"; BERT_B200_LIB=$PWD/$V/libbert_mma2slim0.so timeout 60 python tools/attn_bench.py 512 256 12 32 2>&1 | grep "^attention"
  echo "== head size 32, hoisted descriptors (round $round)"; timeout 60 python tools/attn_bench.py 512 256 12 32 2>&1 | grep "^attention"
done
echo "== e2e parity subset"; timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "golden or edge or full_size_properties" 2>&1 | tail -3 | cut -c1-300
for lib in mmaslim0 default mmaslim0 default; do
  if [ $lib = default ]; then unset BERT_B200_LIB; else export BERT_B200_LIB=$PWD/$V/libbert_$lib.so; fi
  timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2attn_bench_$lib.json; python - <<PY
import json
d=json.load(open('gpurun_out/r2attn_bench_$lib.json')); print('$lib', 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['by_kernel_ms'], d['clocks']['sm_mhz'])
PY
done
