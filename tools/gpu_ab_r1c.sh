# attention: row sums from the P x ones MMA (msum), with the polynomial exp2 every 4th pair / never (msum0) / every 8th (msum8)
L=embeddings.cpp_b200/lib
for v in libbert_msum libbert_msum0; do echo "== tests $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention or golden or named_model or full_size or edge_cases" 2>&1 | grep -E "^E  .*assert|passed|failed" | head -6; done
for r in 1 2; do for v in libbert libbert_msum libbert_msum0 libbert_msum8; do echo "== attn $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 120 python tools/attn_bench.py 2>&1 | head -2; done; done
for r in 1 2; do for v in libbert libbert_msum; do echo "== bench $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['by_kernel_ms'], d['clocks'])"; done; done
