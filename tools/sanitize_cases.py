"""Small kernel self-tests for compute-sanitizer runs (racecheck / synccheck / memcheck), many tiles per persistent
cluster at small M via BERT_GEMM_MAX_CLUSTERS:
    BERT_GEMM_MAX_CLUSTERS=2 compute-sanitizer --tool racecheck python tools/sanitize_cases.py [gemm|attn|e2e ...]
Every instantiation the forward uses is covered: q4_0 / q4_1 / f16 weights x {bias, bias+GELU+q8, bias+residual,
fused LayerNorm (cluster of 6 / 8 CTAs, with and without the short-K variant)}, attention head sizes 64 and 32."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
what = sys.argv[1:] or ["gemm", "attn"]
GEMM = [  # wtype, M, N, K, epilogue
    (2, 2048, 2304, 768, 1), (2, 2048, 3072, 768, 43), (3, 1536, 1536, 384, 43), (1, 2048, 2304, 768, 1),
    (2, 1024, 768, 768, 5), (2, 2048, 768, 768, 29), (2, 1536, 768, 3072, 29), (3, 1024, 1024, 1024, 29),
    (1, 1536, 768, 768, 21), (2, 700, 1152, 384, 1), (2, 600, 192, 384, 11), (2, 256, 2304, 768, 1),
]
ATTN = [(6, 512, 12, 64), (9, 130, 12, 32), (33, 17, 12, 64), (3, 300, 4, 32)]
bad = 0
if "gemm" in what:
    for wt, M, N, K, epi in GEMM:
        rc, err, ref = pkg.selftest_gemm(M, N, K, wt, epi, seed=M + N)
        ok = rc == 0 and err <= 4e-3 * max(ref, 1.0) + (1.1 * ref / 127 if epi & 8 else 0.0)
        bad += not ok
        print(f"gemm wtype {wt} M {M} N {N} K {K} epilogue {epi}: rc {rc} max err {err:.4g} (ref max {ref:.3g}) {'ok' if ok else 'FAIL'}", flush=True)
if "attn" in what:
    for n, L, nh, d in ATTN:
        rc, err, ref = pkg.selftest_attention(n, L, nh, d, seed=L)
        ok = rc == 0 and err <= 6e-3 * max(ref, 1.0)
        bad += not ok
        print(f"attention sentences {n} max_len {L} heads {nh} x {d}: rc {rc} max err {err:.4g} {'ok' if ok else 'FAIL'}", flush=True)
if "e2e" in what:
    import numpy as np
    from tools import synth
    from oracle import fileformat
    for shape, ft in (("tiny-d64", "q4_0"), ("tiny-d32", "q4_1")):
        path = synth.model_path(shape, ft)
        mf = fileformat.read_model(path)
        m = pkg.BertModel(path)
        e = m.encode(synth.texts_var_len(mf.vocab, 40, 2, 300, seed=3, cjk=True))
        print(f"e2e {shape} {ft}: finite {bool(np.isfinite(e).all())} norm err {float(np.abs(np.linalg.norm(e, axis=1) - 1).max()):.2e}", flush=True)
        m.close()
print("cases failed:", bad)
sys.exit(1 if bad else 0)
