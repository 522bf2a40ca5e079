"""Where the end-to-end time of one bert_encode_batch call goes (BERT_TRACE=1 host-side events + the Python-side pieces).
    python tools/e2e_trace.py [workload]"""
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["BERT_TRACE"] = "1"
os.environ.setdefault("BERT_CUDA_DEVICES", "0")
import numpy as np  # noqa: E402
import bench  # noqa: E402
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "minilm-q4_0-b1738-var"
shape, ftype, n_sent, n_tok, kind, scaling = bench.WORKLOADS[name]
texts = bench.make_texts(shape, kind, n_sent, n_tok)
m = bench.quiet_load(pkg, bench.model_file(shape, ftype))
sys.stderr.flush()
devnull = os.open(os.devnull, os.O_WRONLY)
saved = os.dup(2)
os.dup2(devnull, 2)                      # warm-up calls: traces discarded
for _ in range(3):
    m.encode(texts)
os.dup2(saved, 2)
n = len(texts)
t0 = time.perf_counter()
out = np.zeros((n, m.n_embd), dtype=np.float32)
t1 = time.perf_counter()
arr = (ctypes.c_char_p * n)(*texts)
t2 = time.perf_counter()
rows = pkg._row_ptrs(out)
t3 = time.perf_counter()
m.lib.bert_encode_batch(m.ctx, m.n_threads, 16, n, arr, rows)
t4 = time.perf_counter()
print(f"{name}: python side: zeros {1e6 * (t1 - t0):.0f} us, char_p array {1e6 * (t2 - t1):.0f} us, row pointers {1e6 * (t3 - t2):.0f} us, "
      f"bert_encode_batch {1e6 * (t4 - t3):.0f} us")
ids, lens = m.tokenize_batch(texts)
m.stage(ids, lens)
m.run_staged(reps=3)
ms, _ = m.run_staged(reps=10)
print(f"device-resident forward: {100 * ms:.0f} us per pass")
