"""Single-request latency of bert_encode (the TCP server's unit of work):
    python tools/latency.py [shape ftype]      median / p99 over 300 calls for 8-, 32- and 128-token sentences"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import __graft_entry__ as ge  # noqa: E402
from tools import synth  # noqa: E402

pkg = ge.load_package()
shape = sys.argv[1] if len(sys.argv) > 1 else "bge-base-en-v1.5"
ftype = sys.argv[2] if len(sys.argv) > 2 else "q4_0"
path = synth.model_path(shape, ftype)
vocab = synth.make_vocab(synth.SHAPES[shape][4], 0)
m = pkg.BertModel(path)
for n_tok in (8, 32, 128):
    text = synth.texts_exact_len(vocab, 1, n_tok, seed=n_tok)
    for _ in range(20):
        m.encode(text)
    ts = []
    for _ in range(300):
        t0 = time.perf_counter()
        m.encode(text)
        ts.append((time.perf_counter() - t0) * 1e3)
    ts = np.sort(ts)
    print(f"{shape} {ftype} {n_tok:4d} tokens: median {ts[150]:.3f} ms  p99 {ts[297]:.3f} ms  min {ts[0]:.3f} ms", flush=True)

# where the time goes: device time of the staged forward (CUDA events, no host work between launches) and the
# per-kernel-class event times of one profiled forward
for n_sent, n_tok in ((1, 32), (16, 32), (64, 32), (128, 32), (64, 128)):
    ids, lens = m.tokenize_batch(synth.texts_exact_len(vocab, n_sent, n_tok, seed=32))
    m.stage(ids, lens)
    m.run_staged(reps=20)
    ms, _ = m.run_staged(reps=200)
    prof = m.profile_staged()
    print(f"staged forward, {n_sent} x {n_tok} tokens: {ms / 200:.3f} ms per forward (stream time);  per class (events around each launch): "
          + ", ".join(f"{k} {v[0]:.3f} ms / {v[1]} launches" for k, v in prof.items() if v[1]), flush=True)
