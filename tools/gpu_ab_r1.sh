timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== EPB=2 selftest K=3072"; BERT_LN_EPB=2 python -c "
import __graft_entry__ as ge
pkg=ge.load_package()
for a in [(777,768,3072,3,29),(1000,768,3072,2,29),(600,1024,4096,2,29),(515,512,1536,2,29),(300,768,3072,1,21)]:
    print(a, pkg.selftest_gemm(a[0],a[1],a[2],a[3],a[4],seed=5))
"
for e in 0 1 2; do echo "== BERT_LN_EPB=$e"; BERT_LN_EPB=$e python tools/prof_kernels.py 131072 o_ln,ffn_out_ln 20; done
echo "== gelu"; python tools/prof_kernels.py 131072 ffn_in,ffn_in_h2,ffn_in,ffn_in_h2 20
for cfg in "0 0" "1 1" "0 0" "1 1"; do set -- $cfg; echo "== EPB=$1 GELU_H2=$2"; BERT_LN_EPB=$1 BERT_GELU_H2=$2 timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline']['by_kernel_ms'], d['clocks'])"; done
