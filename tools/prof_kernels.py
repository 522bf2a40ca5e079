"""Small driver for `ncu --set full`: one launch of each hot kernel at the bge-base shapes
(q4_0 GEMMs at M tokens, tcgen05 attention over 64 x 512-token sentences).
    python tools/prof_kernels.py [M [names]]     names: comma list of qkv,o,ffn_in,ffn_in_h2,ffn_out,o_ln,ffn_out_ln,attn"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
M = int(sys.argv[1]) if len(sys.argv) > 1 else 37888           # 148 pair tiles of 256 rows
H, I = 768, 3072
names = sys.argv[2].split(",") if len(sys.argv) > 2 else ["qkv", "o", "ffn_in", "ffn_out", "attn"]
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
shapes = {"qkv": (3 * H, H, 1), "o": (H, H, 5), "ffn_in": (I, H, 11), "ffn_out": (H, I, 5),
          "o_ln": (H, H, 29), "ffn_out_ln": (H, I, 29), "ffn_in_h2": (I, H, 43)}
for name in names:
    if name == "attn":
        print("attn", pkg.selftest_attention(64, 512, 12, 64, 1))
        continue
    N, K, epi = shapes[name]
    ms, tf = pkg.bench_gemm(M, N, K, 2, epi, reps=reps)
    print(name, round(ms, 4), "ms", round(tf, 1), "TF")
