"""Generates the committed golden fixtures from the COMPILED, UNMODIFIED reference
(oracle/_ref/libbert_ref.so, built by oracle/Makefile from /root/reference).

    python tests/golden/make_golden.py

For every (tiny shape, weight type): token ids from the reference's bert_tokenize and
embeddings from the reference's bert_encode_batch / bert_forward_batch on a fixed text set
(the reference's own examples/test_prompts.txt, the three sentences of
examples/test_batch_encode.cpp:39-43, plus seeded synthetic multilingual texts).  The model
files are regenerated deterministically by tools/synth.py; their sha256 prefix is stored so
that drift is detected.  Output: tests/golden/<shape>-<ftype>.npz (a few KB each).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import fileformat, reflib  # noqa: E402
from tools import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
PROMPTS = os.path.join(HERE, "test_prompts.txt")
BATCH_ENCODE_SENTENCES = ["你好世界".encode(), "こんにちは、世界！".encode(), b"hello world"]


def text_set(vocab):
    texts = [l.rstrip(b"\n") for l in open(PROMPTS, "rb")]
    texts += BATCH_ENCODE_SENTENCES
    texts += synth.texts_var_len(vocab, 10, 2, 48, seed=21, cjk=True)
    texts += synth.texts_var_len(vocab, 6, 2, 90, seed=22, cjk=False)
    texts += [b"H\xc3\x89llo W\xc3\x96RLD \xc3\xa7a va?", b"dup dupdup ##dup", b"a", b"\xf0\x9f\x98\x80 emoji hello"]
    return texts


def main():
    assert reflib.available(), "build oracle/_ref first (make -C oracle)"
    for shape in ("tiny-d32", "tiny-d64"):
        for ft in ("f32", "f16", "q4_0", "q4_1"):
            path = synth.model_path(shape, ft)
            mf = fileformat.read_model(path)
            texts = text_set(mf.vocab)
            ref = reflib.RefBert(path)
            toks = [ref.tokenize(t) for t in texts]
            emb_fwd = ref.forward_tokens(toks, chunk=4)
            emb_enc = ref.encode(texts, batch_size=4)
            flat = np.concatenate([np.asarray(t, dtype=np.int32) for t in toks])
            lens = np.asarray([len(t) for t in toks], dtype=np.int32)
            out = os.path.join(HERE, f"{shape}-{ft}.npz")
            np.savez_compressed(out, texts=np.array(texts, dtype=object), tokens=flat, lens=lens,
                                emb_forward=emb_fwd, emb_encode=emb_enc, model_sha=synth.sha16(path))
            print(out, len(texts), "texts", os.path.getsize(out), "bytes")
            ref.close()


if __name__ == "__main__":
    main()
