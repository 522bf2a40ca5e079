L=embeddings.cpp_b200/lib
run() { echo "== $*"; env "$@" BERT_B200_LIB=$PWD/$L/libbert_ldx.so timeout 200 python -m pytest tests/test_gpu.py -m gpu -x -q -k "full_size" 2>&1 | grep -E "^E  .*assert|passed|failed" | head -3; }
run A=1; run A=1
run BERT_PDL=0; run BERT_PDL=0
run BERT_ACT_QUANT=0; run BERT_ACT_QUANT=0
run BERT_GELU_H2=0; run BERT_FUSE_LN=0
