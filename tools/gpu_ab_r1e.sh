L=embeddings.cpp_b200/lib
for v in libbert_q2k3; do echo "== tests $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention or golden or full_size or edge_cases" 2>&1 | grep -E "^E  .*assert|passed|failed" | head -4; done
for r in 1 2; do for v in libbert libbert_q2k3 libbert_q1k3 libbert_q2k2; do echo "== attn $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 120 python tools/attn_bench.py 2>&1 | head -4; done; done
