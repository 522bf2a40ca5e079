"""Stress driver: repeats a GEMM micro-benchmark in fresh processes and counts failures.
python tools/stress_gemm.py trials M N K wtype epilogue reps"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
trials, M, N, K, wt, epi, reps = [int(x) for x in sys.argv[1:8]]
code = f"""
import sys
sys.path.insert(0, {ROOT!r})
import __graft_entry__ as ge
pkg = ge.load_package()
for _ in range(4):
    print(pkg.bench_gemm({M}, {N}, {K}, {wt}, {epi}, reps={reps}))
"""
bad = 0
last = ""
for t in range(trials):
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    if r.returncode != 0:
        bad += 1
        last = (r.stdout + r.stderr)[-400:]
    else:
        last_ok = r.stdout.strip().splitlines()[-1]
print(f"M={M} N={N} K={K} wt={wt} epi={epi}: {bad}/{trials} failed", locals().get("last_ok", ""))
if bad:
    print(last)
