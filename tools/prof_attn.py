"""Driver for ncu: one warm-up + one timed launch of the attention kernel at the headline shape."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
print(pkg.bench_attention(512, 512, 12, 64, reps=1))
