"""Attention micro-benchmark on the GPU box: python tools/attn_bench.py [n_sent len n_head d_head]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
a = [int(x) for x in sys.argv[1:]] or [512, 512, 12, 64]
for length in (a[1], a[1] // 2, a[1] // 4, 100):
    ns = a[0] * a[1] // length
    ms, tf = pkg.bench_attention(ns, length, a[2], a[3], reps=10)
    print(f"attention n_sent={ns} len={length} heads={a[2]} d={a[3]}: {ms:.4f} ms  {tf:.1f} TFLOP/s", flush=True)
