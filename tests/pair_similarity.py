"""Offline stand-in for the reference's accuracy harness (benchmarks/run_mteb.py:34-106, STSBenchmark over the modes
f32 / f16 / q4_0 / q4_1).  MTEB and its data are unavailable on this box, so there are no human similarity labels; what
the harness guards instead is that the GPU library ranks sentence pairs the way the reference does:

  * N synthetic sentence pairs (B = A with a random share of its words replaced, so cosine similarities spread over
    roughly [0.2, 1]) are encoded by libbert.so (GPU) and by the compiled reference (oracle/_ref/libbert_ref.so, CPU);
  * per weight type: Pearson and Spearman correlation between the two engines' pair similarities, and the largest
    |difference|;
  * the STS-style score of every mode against the f32 reference's similarities as the "gold" labels -- the number
    run_mteb.py would print per mode -- for both engines side by side.

TEST INFRASTRUCTURE (it runs the oracle): used by tests/test_pair_similarity.py and from the command line,
    python tests/pair_similarity.py [--shape all-MiniLM-L6-v2] [--pairs 2000] [--out profiles/pair_similarity_r2.jsonl]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from oracle import fileformat, reflib  # noqa: E402
from tools import synth  # noqa: E402

MODES = ("f32", "f16", "q4_0", "q4_1")


def make_pairs(vocab, n_pairs, seed=0, lo=4, hi=40):
    rng = np.random.default_rng(4242 + seed)
    words = synth.whole_words(vocab)
    a_texts, b_texts = [], []
    for _ in range(n_pairs):
        n = int(rng.integers(lo, hi + 1))
        a = [words[int(i)] for i in rng.integers(0, len(words), n)]
        share = rng.random()
        b = [words[int(rng.integers(0, len(words)))] if rng.random() < share else w for w in a]
        if rng.random() < 0.25:                                    # some pairs also differ in length
            b = b[: max(2, int(len(b) * rng.uniform(0.5, 1.0)))]
        a_texts.append(b" ".join(a))
        b_texts.append(b" ".join(b))
    return a_texts, b_texts


def pair_cos(emb, n):
    a, b = emb[:n], emb[n:]
    return (a * b).sum(1) / np.linalg.norm(a, axis=1) / np.linalg.norm(b, axis=1)


def pearson(x, y):
    x, y = x - x.mean(), y - y.mean()
    return float((x * y).sum() / np.sqrt((x * x).sum() * (y * y).sum()))


def spearman(x, y):
    rank = lambda v: np.argsort(np.argsort(v, kind="stable"), kind="stable").astype(np.float64)  # noqa: E731
    return pearson(rank(x), rank(y))


def run(shape="all-MiniLM-L6-v2", n_pairs=2000, modes=MODES, seed=0, log=None):
    pkg = ge.load_package()
    vocab = fileformat.read_model(synth.model_path(shape, "f32")).vocab
    a, b = make_pairs(vocab, n_pairs, seed)
    texts = a + b
    sims = {}
    for mode in modes:
        path = synth.model_path(shape, mode)
        m = pkg.BertModel(path)
        gpu = pair_cos(m.encode(texts), n_pairs)
        m.close()
        ref_m = reflib.RefBert(path)
        ref = pair_cos(ref_m.encode(texts, batch_size=32, max_len_hint=64), n_pairs)
        ref_m.close()
        sims[mode] = (gpu, ref)
    gold = sims["f32"][1] if "f32" in sims else None
    rows = []
    for mode in modes:
        gpu, ref = sims[mode]
        row = {"shape": shape, "weights": mode, "pairs": n_pairs,
               "similarity_range": [round(float(ref.min()), 3), round(float(ref.max()), 3)],
               "pearson_gpu_vs_reference": pearson(gpu, ref), "spearman_gpu_vs_reference": spearman(gpu, ref),
               "max_abs_similarity_diff": float(np.abs(gpu - ref).max())}
        if gold is not None:
            row["sts_style_spearman_vs_f32_reference"] = {"gpu": spearman(gpu, gold), "reference": spearman(ref, gold)}
        rows.append(row)
        if log:
            log(json.dumps(row))
    return rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="all-MiniLM-L6-v2")
    ap.add_argument("--pairs", type=int, default=2000)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    out = open(args.out, "a") if args.out else None

    def log(line):
        print(line, flush=True)
        if out:
            out.write(line + "\n")
            out.flush()
    run(args.shape, args.pairs, log=log)
