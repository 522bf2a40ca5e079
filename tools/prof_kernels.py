"""Small driver for `ncu --set full`: one launch of each hot kernel at the bge-base shapes
(q4_0 GEMMs at M tokens, tcgen05 attention over 64 x 512-token sentences).
    python tools/prof_kernels.py [M]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
M = int(sys.argv[1]) if len(sys.argv) > 1 else 37888           # 148 pair tiles of 256 rows
H, I = 768, 3072
for name, N, K, epi in (("qkv", 3 * H, H, 1), ("o", H, H, 5), ("ffn_in", I, H, 11), ("ffn_out", H, I, 5)):
    ms, tf = pkg.bench_gemm(M, N, K, 2, epi, reps=1)
    print(name, round(ms, 4), "ms", round(tf, 1), "TF")
print("attn", pkg.selftest_attention(64, 512, 12, 64, 1))
