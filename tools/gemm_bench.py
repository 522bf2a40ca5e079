"""GEMM micro-benchmark on the GPU box: the four GEMM shapes of a model at a given
micro-batch, fp16-weight (TMA B) vs fused q4 dequant.  python tools/gemm_bench.py [H I M]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
H = int(sys.argv[1]) if len(sys.argv) > 1 else 768
I = int(sys.argv[2]) if len(sys.argv) > 2 else 3072
M = int(sys.argv[3]) if len(sys.argv) > 3 else 32768
shapes = [("qkv", 3 * H, H, 1), ("o", H, H, 5), ("ffn_in", I, H, 3), ("ffn_out", H, I, 5)]
for wt, wn in ((1, "f16"), (2, "q4_0"), (3, "q4_1")):
    tot_ms, tot_fl = 0.0, 0.0
    for name, N, K, epi in shapes:
        e = epi | (8 if (wt > 1 and name == "ffn_in") else 0)
        ms, tf = pkg.bench_gemm(M, N, K, wt, e, reps=10)
        tot_ms += ms
        tot_fl += 2.0 * M * N * K
        print(json.dumps({"w": wn, "gemm": name, "M": M, "N": N, "K": K, "epi": e, "ms": round(ms, 4), "tflops": round(tf, 1)}), flush=True)
    print(json.dumps({"w": wn, "layer_gemm_ms": round(tot_ms, 4), "tflops": round(tot_fl / tot_ms / 1e9, 1)}), flush=True)
