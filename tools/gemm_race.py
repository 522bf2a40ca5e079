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


def analyse(M, N, K, wt, epi, ref, tc, ncl):
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
        if nbad == 0:
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
        ex = np.random.default_rng(0).choice(len(rr), size=min(6, len(rr)), replace=False)
        for i in ex:
            out.append(f"    e.g. row {rr[i]} col {cc[i]} tile(m {mb[i]}, n {nb[i]}) cid {cid[i]} seq {seq[i]} rank {rank[i]} quad {quad[i]} box {box[i]} "
                       f"half {half[i]}: got {t32[rr[i], cc[i]]:.5f} want {r32[rr[i], cc[i]]:.5f}")
    return out


def main():
    pkg = ge.load_package()
    L = pkg.lib()
    L.bertx_debug_gemm.restype = ctypes.c_int32
    cases = sys.argv[1:] or DEFAULT_CASES
    ncl = int(os.environ.get("BERT_GEMM_MAX_CLUSTERS", "0")) or 74
    print(f"lib {os.path.basename(pkg.LIB_PATH)}  clusters {ncl}", flush=True)
    for c in cases:
        v = [int(x) for x in c.split(",")]
        M, N, K, wt, epi = v[:5]
        runs = v[5] if len(v) > 5 else 3
        ref = np.zeros((M, N), dtype=np.float16)
        tc = np.zeros((runs, M, N), dtype=np.float16)
        ms = np.zeros(runs, dtype=np.float32)
        rc = L.bertx_debug_gemm(M, N, K, wt, epi, 11, runs, ctypes.c_void_p(ref.ctypes.data), ctypes.c_void_p(tc.ctypes.data),
                                ctypes.c_void_p(ms.ctypes.data))
        print(f"case M={M} N={N} K={K} wtype={wt} epilogue={epi}: rc {rc}, kernel ms {[round(float(x), 4) for x in ms]}", flush=True)
        if rc == 0:
            print("\n".join(analyse(M, N, K, wt, epi, ref, tc, ncl)), flush=True)


if __name__ == "__main__":
    main()
