L=embeddings.cpp_b200/lib
for r in 1 2 3; do
for v in libbert libbert_base; do
echo "== $v round $r"
BERT_B200_LIB=$PWD/$L/$v.so timeout 300 python -m pytest tests/test_gpu.py -m gpu -x -q -k "full_size_properties" 2>&1 | grep -E "^E |passed|failed|assert" | head -12
done; done
