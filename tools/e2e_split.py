"""Where the end-to-end time of bert_encode_batch goes at the headline shape: tokenizer vs forward vs
Python marshalling.   python tools/e2e_split.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
from tools import synth  # noqa: E402
from oracle import fileformat  # noqa: E402

path = synth.model_path("bge-base-en-v1.5", "q4_0")
mf = fileformat.read_model(path)
texts = synth.texts_exact_len(mf.vocab, 512, 512, seed=0)
print("bytes per text", sum(len(t) for t in texts) / len(texts))
m = pkg.BertModel(path)


def t(fn, n=5):
    fn()
    t0 = time.perf_counter()
    for _ in range(n):
        r = fn()
    return (time.perf_counter() - t0) / n * 1e3, r


ms_tok, (ids, lens) = t(lambda: m.tokenize_batch(texts))
toks = np.split(np.asarray(ids), np.cumsum(lens)[:-1])
ms_fwd, _ = t(lambda: m.forward_tokens(toks))
ms_e2e, _ = t(lambda: m.encode(texts))
bt = [s.decode() for s in texts]
ms_e2e_b, _ = t(lambda: m.encode(bt))
print(f"tokenize_batch {ms_tok:.2f} ms   forward_tokens {ms_fwd:.2f} ms   encode(bytes) {ms_e2e:.2f} ms   encode(str) {ms_e2e_b:.2f} ms")
