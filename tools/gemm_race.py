"""Element-wise bisect of the persistent tcgen05 GEMM at full size (profiles/r2_gemm_race.md).

    BERT_B200_LIB=<variant .so> python tools/gemm_race.py [case ...]     case = M,N,K,wtype,epilogue[,runs]

Runs bertx_debug_gemm (SIMT validation kernel once, tcgen05 kernel `runs` times on the same device-generated
operands) and reports, per case: run-to-run bitwise equality, elements off against the validation kernel, and
where they sit in the kernel's own coordinates -- tile sequence number inside the persistent cluster, CTA rank of
the pair, TMEM lane quadrant, 64-column box, 32-column half, lane -- plus whether a wrong value equals the value the
same cluster computes at the same in-tile position 1 or 2 tiles earlier / later (stale or early-overwritten
accumulator).
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

DEFAULT_CASES = ["131072,2304,768,2,1,3", "131072,3072,768,2,43,3"]


def pick_bn(N):
    for bn in (256, 192, 128, 64):
        if N % bn == 0:
            return bn
    return 0


MATRIX = [  # every GEMM instantiation the five named models launch, at many tiles per persistent cluster
    "32768,2304,768,2,1", "32768,2304,768,2,0", "32768,3072,768,2,43", "32768,768,768,2,29", "32768,768,3072,2,29", "32768,768,768,2,5",
    "32768,1152,384,2,1", "32768,1536,384,2,43", "32768,384,384,2,5", "32768,384,1536,2,5", "32768,960,768,2,1", "32768,320,768,2,1",
    "32768,1536,512,3,1", "32768,2048,512,3,43", "32768,512,512,3,29", "32768,512,2048,3,29",
    "32768,3072,1024,2,1", "32768,1024,1024,2,29", "32768,4096,1024,2,43", "32768,1024,4096,2,29",
    "32768,2304,768,1,1", "32768,768,768,1,21", "32768,3072,768,1,3", "32768,768,3072,1,21",
]


def analyse(M, N, K, wt, epi, ref, tc, ncl, brief=False):
    bn = pick_bn(N)
    num_n, num_m = N // bn, (M + 255) // 256
    r32 = ref.astype(np.float32)
    tol = 6e-3 * np.maximum(1.0, np.abs(r32))
    if epi & 8:                                     # q8 fake-quant: one step of the 32-wide block may flip
        amax = np.abs(r32).reshape(M, N // 32, 32).max(2, keepdims=True)
        tol = (tol.reshape(M, N // 32, 32) + 1.1 * amax / 127).reshape(M, N)
    out = []
    for r in range(tc.shape[0]):
        t32 = tc[r].astype(np.float32)
        bad = ~(np.abs(t32 - r32) <= tol)           # NaN counts as bad
        nbad = int(bad.sum())
        same = None if r == 0 else int((tc[r].view(np.uint16) != tc[0].view(np.uint16)).sum())
        line = f"  run {r}: off vs validation kernel {nbad} of {M * N}"
        if same is not None:
            line += f"; elements differing from run 0: {same}"
        out.append(line)
        if nbad == 0 or brief or (epi & 16):
            if nbad and (epi & 16):
                rr, cc = np.nonzero(bad)
                out.append(f"    fused LayerNorm: 256-row stripes hit {len(np.unique(rr // 256))}, column blocks hit {sorted(set((cc // bn).tolist()))}, "
                           f"max |err| {float(np.nanmax(np.abs(t32 - r32)[bad])):.4g}")
            continue
        rr, cc = np.nonzero(bad)
        mb, nb = rr // 256, cc // bn
        tile = mb * num_n + nb
        total = num_m * num_n
        n_cl = min(ncl, total)
        cid, seq = tile % n_cl, tile // n_cl
        my_tiles = (total - cid + n_cl - 1) // n_cl
        rank, quad, lane = (rr // 128) & 1, (rr % 128) // 32, rr % 32
        box, half = (cc % bn) // 64, (cc % 64) // 32

        def hist(name, v, n=None):
            h = np.bincount(v, minlength=n or 0)
            nz = {int(i): int(c) for i, c in enumerate(h) if c}
            out.append(f"    by {name}: {nz if len(nz) <= 24 else str(dict(list(nz.items())[:24])) + ' ...'}")
        out.append(f"    tiles hit {len(np.unique(tile))} of {total}; clusters hit {len(np.unique(cid))} of {n_cl}; "
                   f"max |err| {float(np.nanmax(np.abs(t32 - r32)[bad])):.4g}; NaN/inf {int((~np.isfinite(t32[bad])).sum())}")
        hist("rank", rank, 2); hist("quadrant", quad, 4); hist("box", box); hist("half", half, 2)
        hist("accumulator (seq&1)", seq & 1, 2)
        hist("seq (tile index in its cluster)", seq)
        hist("tiles left after this one", my_tiles - 1 - seq)
        hist("lane", lane, 32)
        hist("column in half", cc % 32, 32)
        # per (tile, box-warp) count: whole 32x32 sub-boxes or scattered elements?
        key = ((tile * 8 + rank * 4 + quad) * 8 + box * 2 + half)
        _, cnt = np.unique(key, return_counts=True)
        out.append(f"    per 32x32 sub-box: {len(cnt)} sub-boxes, elements off min/median/max {cnt.min()}/{int(np.median(cnt))}/{cnt.max()} of 1024")
        for d in (-2, -1, 1, 2):                    # same in-tile position, d tiles away in the same cluster
            t2 = tile + d * n_cl
            ok = (t2 >= 0) & (t2 < total)
            if not ok.any():
                continue
            r2 = (t2[ok] // num_n) * 256 + rr[ok] % 256
            c2 = (t2[ok] % num_n) * bn + cc[ok] % bn
            inb = r2 < M
            hit = np.abs(t32[rr[ok][inb], cc[ok][inb]] - r32[r2[inb], c2[inb]]) <= tol[r2[inb], c2[inb]]
            out.append(f"    equals the validation value of tile seq{d:+d}: {int(hit.sum())} of {int(inb.sum())}")
        # which columns of each hit tile are off (B-operand rows: a dequant warp owns 16 rows + the 16 rows 64 further)
        for t in np.unique(tile)[:12]:
            sel = tile == t
            cols = np.unique(cc[sel] % bn)
            runs_, start = [], None
            for c in range(bn + 1):
                inside = c < bn and c in set(cols.tolist())
                if inside and start is None:
                    start = c
                if not inside and start is not None:
                    runs_.append(f"{start}-{c - 1}"); start = None
            rows_hit = len(np.unique(rr[sel] % 256))
            e = (t32 - r32)[rr[sel], cc[sel]]
            out.append(f"    tile {int(t)} (seq {int(t) // n_cl}): columns off {runs_}, rows hit {rows_hit} of 256, rms err {float(np.sqrt(np.mean(e * e))):.3f}")
        ex = np.random.default_rng(0).choice(len(rr), size=min(6, len(rr)), replace=False)
        for i in ex:
            out.append(f"    e.g. row {rr[i]} col {cc[i]} tile(m {mb[i]}, n {nb[i]}) cid {cid[i]} seq {seq[i]} rank {rank[i]} quad {quad[i]} box {box[i]} "
                       f"half {half[i]}: got {t32[rr[i], cc[i]]:.5f} want {r32[rr[i], cc[i]]:.5f}")
    return out


def solve(M, N, K, epi, ref, tc, A, W, ncl):
    """Which k-stage of which B rows went wrong?  For every corrupted (tile, 16-column group): least-squares fit of the
    error on the per-stage partial products (a coefficient of -1 = that stage's true contribution is missing), then the
    remainder is correlated with the partial products of OTHER weight blocks (same rows / other k-stage, other n tile)."""
    bn = pick_bn(N)
    num_n = N // bn
    nk = K // 64
    out = []
    r32 = ref.astype(np.float32)
    A32 = None
    done = 0
    for r in range(tc.shape[0]):
        t32 = tc[r].astype(np.float32)
        bad = ~(np.abs(t32 - r32) <= 6e-3 * np.maximum(1.0, np.abs(r32)))
        if not bad.any():
            continue
        rr, cc = np.nonzero(bad)
        tile = (rr // 256) * num_n + cc // bn
        for t in np.unique(tile):
            sel = tile == t
            mb, nb = int(t) // num_n, int(t) % num_n
            groups = np.unique((cc[sel] % bn) // 8)
            for g in groups:
                if done >= 24:
                    return out
                done += 1
                c0 = nb * bn + int(g) * 8
                rows = slice(mb * 256, mb * 256 + 256)
                E = (t32 - r32)[rows, c0:c0 + 8].astype(np.float64)                      # [256, 8]
                At = A[rows].astype(np.float64).reshape(256, nk, 64)                    # [256, nk, 64]
                Wn = W[c0:c0 + 8].astype(np.float64).reshape(8, nk, 64)                 # [8, nk, 64]
                P = np.einsum("rks,nks->krn", At, Wn)                                   # [nk, 256, 8]
                X = P.reshape(nk, -1).T
                coef, *_ = np.linalg.lstsq(X, E.reshape(-1), rcond=None)
                kb = int(np.argmin(coef))
                line = (f"  run {r} tile {int(t)} (m {mb}, n {nb}, seq {int(t) // ncl}) cols {c0 % bn}-{c0 % bn + 7}: rms err {np.sqrt((E * E).mean()):.3f}; "
                        f"stage coefficients {np.round(coef, 2).tolist()}")
                out.append(line)
                # remainder after putting the missing stages back: what was used instead?
                miss = [k for k in range(nk) if coef[k] < -0.5]
                for k in miss:
                    R = E + P[k]                                                        # = A_k . W_used^T
                    # candidates: every aligned 8-row group of W x every k-stage, multiplied by the A of stage k
                    Wall = W.astype(np.float64).reshape(N, nk, 64)
                    Z = np.einsum("rs,nks->nkr", At[:, k, :], Wall)                      # [N, nk, 256]
                    Z = Z.reshape(N // 8, 8, nk, 256).transpose(0, 2, 3, 1)             # [N/8, nk, 256, 8]
                    num = (Z * R[None, None]).sum((2, 3))
                    den = np.sqrt((Z * Z).sum((2, 3)) * (R * R).sum()) + 1e-30
                    corr = num / den                                                    # [N/8, nk]
                    idx = np.argsort(corr.reshape(-1))[::-1][:3]
                    best = [(float(corr.reshape(-1)[i]), f"rows {int(i // nk) * 8}-{int(i // nk) * 8 + 7} (n tile {int(i // nk) * 8 // bn}, in-tile {int(i // nk) * 8 % bn})", int(i % nk)) for i in idx]
                    best.sort(reverse=True)
                    out.append(f"    stage {k} missing; remainder rms {np.sqrt((R * R).mean()):.3f}; best matches (corr, n tile, k-stage of the weights used instead): "
                               f"{[(round(c, 3), a, b) for c, a, b in best[:3]]}")
    return out


def main():
    pkg = ge.load_package()
    L = pkg.lib()
    L.bertx_debug_gemm.restype = ctypes.c_int32
    args = sys.argv[1:]
    brief = "--brief" in args
    do_solve = "--solve" in args
    args = [a for a in args if a not in ("--brief", "--solve")]
    cases = MATRIX if args == ["--matrix"] else (args or DEFAULT_CASES)
    ncl = int(os.environ.get("BERT_GEMM_MAX_CLUSTERS", "0")) or 74
    print(f"lib {os.path.basename(pkg.LIB_PATH)}  clusters {ncl}", flush=True)
    for c in cases:
        v = [int(x) for x in c.split(",")]
        M, N, K, wt, epi = v[:5]
        runs = v[5] if len(v) > 5 else 3
        twin = (epi & 16) and (epi & 8)
        ref = np.zeros((M, N), dtype=np.float16)
        tc = np.zeros((runs, M, N), dtype=np.float16)
        ref2 = np.zeros((M, N), dtype=np.float16) if twin else None
        tc2 = np.zeros((runs, M, N), dtype=np.float16) if twin else None
        ms = np.zeros(runs, dtype=np.float32)
        vp = lambda x: ctypes.c_void_p(x.ctypes.data) if x is not None else None  # noqa: E731
        A = np.zeros((M, K), dtype=np.float16) if do_solve else None
        W = np.zeros((N, K), dtype=np.float32) if do_solve else None
        rc = L.bertx_debug_gemm(M, N, K, wt, epi, 11, runs, vp(ref), vp(tc), vp(ms), vp(ref2), vp(tc2), vp(A), vp(W))
        print(f"case M={M} N={N} K={K} wtype={wt} epilogue={epi}: rc {rc}, kernel ms {[round(float(x), 4) for x in ms]}", flush=True)
        if rc == 0:
            print("\n".join(analyse(M, N, K, wt, epi, ref, tc, ncl, brief)), flush=True)
            if do_solve:
                print("\n".join(solve(M, N, K, epi, ref, tc, A, W, ncl)), flush=True)
            if twin:
                print("  q8 twin:\n" + "\n".join(analyse(M, N, K, wt, 8, ref2, tc2, ncl, True)), flush=True)


if __name__ == "__main__":
    main()
