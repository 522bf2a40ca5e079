# round-1 late A/B: epilogue TMEM double load (ld2), attention polynomial share (p3 / p2 vs every 4th), new tests
L=embeddings.cpp_b200/lib
echo "== xu_mix"; timeout 60 tools/micro/xu_mix
timeout 300 python -m pytest tests -m gpu -x -q -k "cls_pooling or bge-large" 2>&1 | tail -3
for r in 1 2; do for v in base ld2; do echo "== gemm $v"; BERT_B200_LIB=$PWD/$L/libbert_$v.so timeout 120 python tools/prof_kernels.py 131072 qkv,ffn_in_h2,o_ln,ffn_out_ln 20; done; done
for r in 1 2; do for v in base p3 p2; do echo "== attn $v"; BERT_B200_LIB=$PWD/$L/libbert_$v.so timeout 120 python tools/attn_bench.py 2>&1 | head -2; done; done
for r in 1 2; do for v in base ld2 p3 p2; do echo "== bench $v"; BERT_B200_LIB=$PWD/$L/libbert_$v.so timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['by_kernel_ms'], d['clocks'])"; done; done
