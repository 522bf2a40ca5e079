L=embeddings.cpp_b200/lib
run() { echo "== $*"; env "$@" timeout 200 python -m pytest tests/test_gpu.py -m gpu -x -q -k "full_size" 2>&1 | grep -E "^E  .*assert|passed|failed" | head -3; }
for r in 1 2 3; do run BERT_B200_LIB=$PWD/$L/libbert_ldx1.so; run BERT_B200_LIB=$PWD/$L/libbert_ldx1s.so; done
