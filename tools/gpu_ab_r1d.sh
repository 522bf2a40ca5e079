L=embeddings.cpp_b200/lib
for v in libbert_spec; do echo "== tests $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "fused_attention or golden or named_model or full_size or edge_cases" 2>&1 | grep -E "^E  .*assert|passed|failed" | head -6; done
for r in 1 2 3; do for v in libbert libbert_spec; do echo "== attn $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 120 python tools/attn_bench.py 2>&1 | head -2; done; done
