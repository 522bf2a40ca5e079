"""Synthetic model files and texts for the five BASELINE.json configs (test/bench input
generator -- no real checkpoints or vocab files exist on the boxes).

Model files are written in the reference's own format through oracle.fileformat
(a restatement of models/convert-to-ggml.py:68-110 and models/quantize.cpp:154-217),
so both the compiled reference and libbert.so load the very same bytes.
Everything is seeded and deterministic; files are cached under $BERT_SYNTH_CACHE
(default /tmp/bert_synth).
"""
import hashlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import fileformat  # noqa: E402  (generator is test infrastructure)

# name: (H, I, heads, layers, vocab)   -- public HF config.json values
SHAPES = {
    "all-MiniLM-L6-v2": (384, 1536, 12, 6, 30522),
    "bge-small-en-v1.5": (384, 1536, 12, 12, 30522),
    "bge-base-en-v1.5": (768, 3072, 12, 12, 30522),
    "bge-small-zh-v1.5": (512, 2048, 8, 4, 21128),
    "bge-large-en-v1.5": (1024, 4096, 16, 24, 30522),
    # tiny shapes for fast CPU tests / committed golden fixtures
    "tiny-d32": (64, 256, 2, 2, 600),
    "tiny-d64": (128, 512, 2, 2, 600),
}

CACHE = os.environ.get("BERT_SYNTH_CACHE", "/tmp/bert_synth")

_LETTERS = "abcdefghijklmnopqrstuvwxyz"


def make_vocab(n_vocab, seed=0):
    """A WordPiece-shaped vocabulary: specials at the BERT ids (100 [UNK], 101 [CLS],
    102 [SEP]), single ASCII chars, '##' single chars, CJK / kana single chars, then
    seeded random whole words and '##' suffixes.  A few deliberate duplicates exercise
    the first-wins / last-wins map rules (bert.cpp:483-493)."""
    rng = np.random.default_rng(1000 + seed)
    v = [b"[PAD]"] + [f"[unused{i}]".encode() for i in range(1, 100)]
    v += [b"[UNK]", b"[CLS]", b"[SEP]", b"[MASK]"]
    v += [bytes([c]) for c in range(33, 127)]                       # printable ASCII singles
    v += [b"##" + bytes([c]) for c in range(ord("a"), ord("z") + 1)]
    v += [b"##" + bytes([c]) for c in range(ord("0"), ord("9") + 1)]
    n_cjk = min(3000, max(60, n_vocab // 8))
    cjk = [chr(0x4E00 + i).encode() for i in range(n_cjk)]
    v += cjk
    v += [b"##" + c for c in cjk[: n_cjk // 4]]
    v += ["，".encode(), "。".encode(), "！".encode(), "（".encode(), "）".encode(), "、".encode()]
    kana = [chr(0x3041 + i).encode() for i in range(80)]
    v += kana + [b"##" + c for c in kana]
    v += [b"hello", b"world", b"##ing", b"##ed", b"##s", b"the", b"a", b"dup", b"dup", b"##dup", b"##dup"]
    seen = set(v)
    while len(v) < n_vocab:
        ln = int(rng.integers(2, 9))
        w = "".join(_LETTERS[int(i)] for i in rng.integers(0, 26, ln)).encode()
        if rng.random() < 0.3:
            w = b"##" + w
        if w in seen:
            continue
        seen.add(w)
        v.append(w)
    return v[:n_vocab]


def make_weights(H, I, n_layer, n_vocab, n_max_tokens=512, seed=0, outliers=False):
    """Random weights with BERT-like statistics: unit-variance attention logits (Q/K sigma
    1/sqrt(H)) and sub-layer outputs at about half the residual stream's scale (O and FFN-out
    sigma 0.5/sqrt(fan_in)), LayerNorm gamma near 1.  Sharper/larger choices make a 12-layer
    random-weight stack chaotic: the reference's own q8 activation rounding then flips enough
    decisions that the reference disagrees WITH ITSELF across batch compositions
    (bert_forward_batch of 4 vs of 1: cosine 0.99988 on the bge-base shape, measured), which
    would make a 0.9999 parity bar meaningless.  With these statistics the reference is
    self-consistent to 0.999999 and parity is a real test."""
    rng = np.random.default_rng(seed)

    def mat(o, i, s):
        return (rng.standard_normal((o, i), dtype=np.float32) * np.float32(s))

    def vec(n, s, mean=0.0):
        return (rng.standard_normal(n, dtype=np.float32) * np.float32(s) + np.float32(mean))

    t = {
        "embeddings.word_embeddings.weight": mat(n_vocab, H, 0.05),
        "embeddings.position_embeddings.weight": mat(n_max_tokens, H, 0.02),
        "embeddings.token_type_embeddings.weight": mat(2, H, 0.02),
        "embeddings.LayerNorm.weight": vec(H, 0.1, 1.0),
        "embeddings.LayerNorm.bias": vec(H, 0.05),
    }
    sa = 1.0 / np.sqrt(H)
    for i in range(n_layer):
        p = f"encoder.layer.{i}."
        t[p + "attention.self.query.weight"] = mat(H, H, 1.0 * sa)
        t[p + "attention.self.query.bias"] = vec(H, 0.1)
        t[p + "attention.self.key.weight"] = mat(H, H, 1.0 * sa)
        t[p + "attention.self.key.bias"] = vec(H, 0.1)
        t[p + "attention.self.value.weight"] = mat(H, H, 1.0 * sa)
        t[p + "attention.self.value.bias"] = vec(H, 0.05)
        t[p + "attention.output.dense.weight"] = mat(H, H, 0.5 * sa)
        t[p + "attention.output.dense.bias"] = vec(H, 0.05)
        t[p + "attention.output.LayerNorm.weight"] = vec(H, 0.1, 1.0)
        t[p + "attention.output.LayerNorm.bias"] = vec(H, 0.05)
        t[p + "intermediate.dense.weight"] = mat(I, H, 1.0 * sa)
        t[p + "intermediate.dense.bias"] = vec(I, 0.1)
        t[p + "output.dense.weight"] = mat(H, I, 0.5 / np.sqrt(I))
        t[p + "output.dense.bias"] = vec(H, 0.05)
        t[p + "output.LayerNorm.weight"] = vec(H, 0.1, 1.0)
        t[p + "output.LayerNorm.bias"] = vec(H, 0.05)
    if outliers:
        # Outlier-stress variant: real BERT / BGE checkpoints carry a handful of hidden channels whose LayerNorm gain and
        # dense biases are 20-50x the typical magnitude, so the residual stream has a few huge coordinates (the
        # real-weight caveat of 8c).  Four fixed channels get that treatment in every LayerNorm and every output bias.
        orng = np.random.default_rng(9000 + seed)
        ch = orng.choice(H, 4, replace=False)
        for name, w in t.items():
            if name.endswith("LayerNorm.weight"):
                w[ch] = orng.uniform(20.0, 50.0, 4).astype(np.float32) * orng.choice([-1.0, 1.0], 4).astype(np.float32)
            elif name.endswith("LayerNorm.bias"):
                w[ch] = orng.uniform(1.0, 2.5, 4).astype(np.float32) * orng.choice([-1.0, 1.0], 4).astype(np.float32)
            elif name.endswith("output.dense.bias"):
                w[ch] = orng.uniform(1.0, 2.5, 4).astype(np.float32) * orng.choice([-1.0, 1.0], 4).astype(np.float32)
    return t


def model_path(shape, ftype, seed=0, n_layer=None, n_max_tokens=512, outliers=False):
    """Return the path of the (cached) synthetic model file, generating it if needed."""
    H, I, nh, nl, nv = SHAPES[shape]
    if n_layer is not None:
        nl = n_layer
    ft = fileformat.FTYPE_IDS[ftype] if isinstance(ftype, str) else ftype
    os.makedirs(CACHE, exist_ok=True)
    tag = f"{shape}-L{nl}-P{n_max_tokens}-s{seed}-v2{'-outl' if outliers else ''}-{fileformat.FTYPE_NAMES[ft]}"
    path = os.path.join(CACHE, f"ggml-model-{tag}.bin")
    if os.path.exists(path):
        return path
    hp = fileformat.Hparams(nv, n_max_tokens, H, I, nh, nl, ft)
    vocab = make_vocab(nv, seed)
    w = make_weights(H, I, nl, nv, n_max_tokens, seed, outliers)
    tmp = path + f".tmp{os.getpid()}"
    fileformat.write_model(tmp, hp, vocab, w, ft)
    os.replace(tmp, path)
    return path


def whole_words(vocab):
    """Vocabulary entries that tokenize to exactly themselves as one id: alphabetic whole
    words (no '##', no punctuation), first occurrence only."""
    out, seen = [], set()
    for i, w in enumerate(vocab):
        if i < 104 or w.startswith(b"##") or not w.isalpha() or not w.islower() or w in seen:
            continue
        seen.add(w)
        out.append(w)
    return out


def texts_exact_len(vocab, n_texts, n_tokens, seed=0):
    """Texts that tokenize to exactly n_tokens ids ([CLS] + n_tokens-2 whole words + [SEP])."""
    rng = np.random.default_rng(77 + seed)
    words = whole_words(vocab)
    out = []
    for _ in range(n_texts):
        idx = rng.integers(0, len(words), n_tokens - 2)
        out.append(b" ".join(words[int(i)] for i in idx))
    return out


def texts_var_len(vocab, n_texts, lo, hi, seed=0, cjk=False):
    """Variable-length texts; with cjk=True mixes CJK chars, kana, full-width punctuation."""
    rng = np.random.default_rng(99 + seed)
    words = whole_words(vocab)
    cj = [w for w in vocab if len(w) == 3 and 0xE4 <= w[0] <= 0xE9 and not w.startswith(b"##")]
    punct = [b",", b".", b"!", b"?", "，".encode(), "。".encode(), "！".encode()]
    out = []
    for _ in range(n_texts):
        n = int(rng.integers(lo, hi + 1))
        parts = []
        for _ in range(max(n - 2, 0)):
            r = rng.random()
            if cjk and r < 0.5 and cj:
                parts.append(cj[int(rng.integers(0, len(cj)))])          # no space needed
            elif r < 0.6:
                parts.append(punct[int(rng.integers(0, len(punct) if cjk else 4))])
            else:
                parts.append(b" " + words[int(rng.integers(0, len(words)))] + b" ")
        out.append(b"".join(parts))
    return out


def sample_client_texts(vocab, n_lines=1738, seed=1738):
    """Stand-in for the missing examples/sample_client_texts.txt (1738 lines; clipped
    log-normal word count, median 18, max 60 -- SURVEY.md section 8d config 1)."""
    rng = np.random.default_rng(seed)
    words = whole_words(vocab)
    out = []
    for _ in range(n_lines):
        n = int(np.clip(np.round(rng.lognormal(np.log(18), 0.5)), 3, 60))
        idx = rng.integers(0, len(words), n)
        s = b" ".join(words[int(i)] for i in idx)
        if rng.random() < 0.5:
            s += b"?" if rng.random() < 0.3 else b"."
        out.append(s[:1].upper() + s[1:])
    return out


def sha16(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()[:16]


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("shape", choices=list(SHAPES))
    ap.add_argument("ftype", choices=["f32", "f16", "q4_0", "q4_1"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--layers", type=int, default=None)
    a = ap.parse_args()
    print(model_path(a.shape, a.ftype, a.seed, a.layers))
