L=embeddings.cpp_b200/lib
for r in 1 2 3; do
for v in libbert libbert_ld2s libbert_attw; do
echo "== $v round $r"
BERT_B200_LIB=$PWD/$L/$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "full_size_properties or tcgen05_gemm or fused_attention" 2>&1 | grep -E "^E  .*assert|passed|failed" | head -6
done; done
for r in 1 2; do for v in libbert libbert_ld2s; do echo "== gemm $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 120 python tools/prof_kernels.py 131072 qkv,ffn_in_h2 20; done; done
for r in 1 2; do for v in libbert libbert_attw; do echo "== attn $v"; BERT_B200_LIB=$PWD/$L/$v.so timeout 120 python tools/attn_bench.py 2>&1 | head -1; done; done
