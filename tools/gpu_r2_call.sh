#!/bin/bash
# round 2 GPU call: converter / loader tests, accuracy harness, attention self-tests of all kernels, default bench line
mkdir -p gpurun_out
echo "== converter"; timeout 300 python -m pytest tests/test_convert.py -m gpu -x -q 2>&1 | tail -4 | cut -c1-300
echo "== attention self-tests"; timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention" 2>&1 | tail -3 | cut -c1-300
echo "== pair similarity test"; timeout 600 python -m pytest tests/test_pair_similarity.py -m gpu -x -q 2>&1 | tail -4 | cut -c1-300
echo "== pair similarity harness"; rm -f gpurun_out/pair_similarity_r2.jsonl
timeout 900 python tests/pair_similarity.py --pairs 2000 --out gpurun_out/pair_similarity_r2.jsonl 2>/dev/null | grep '^{' | cut -c1-400
timeout 900 python tests/pair_similarity.py --shape bge-small-en-v1.5 --pairs 500 --out gpurun_out/pair_similarity_r2.jsonl 2>/dev/null | grep '^{' | cut -c1-400
echo "== bench"; timeout 300 python bench.py --no-cpu-baseline 2>/dev/null | grep '^{' > gpurun_out/r2c_bench_default.json; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c_bench_default.json')); print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['roofline']['by_kernel_ms'], d['clocks'])
PY
