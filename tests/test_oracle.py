"""CPU suite, part 1: the oracle (oracle/) is pinned against the compiled reference.

Golden fixtures under tests/golden/*.npz were produced by tests/golden/make_golden.py from
oracle/_ref/libbert_ref.so (the unmodified reference).  Where the .so is present the same
checks also run live.
"""
import glob
import os
import subprocess

import numpy as np
import pytest

from conftest import cos_and_maxabs
from oracle import fileformat, forward, quants, reflib, tokenizer
from tools import synth

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
# restatement-vs-reference tolerance (float32 summation order only; see oracle/forward.py)
TOL = {"f32": (0.999999, 1e-4), "f16": (0.999999, 2e-4), "q4_0": (0.99999, 2e-3), "q4_1": (0.99999, 3e-3)}


def load_golden(path):
    g = np.load(path, allow_pickle=True)
    shape, ft = os.path.basename(path)[:-4].rsplit("-", 1)
    model = synth.model_path(shape, ft)
    assert synth.sha16(model) == str(g["model_sha"]), "synthetic model drifted from the golden fixture"
    lens = g["lens"]
    toks = np.split(g["tokens"], np.cumsum(lens)[:-1])
    return shape, ft, model, list(g["texts"]), [t.tolist() for t in toks], g


def test_fixture_count():
    assert len(GOLDEN) == 8


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_tokenizer_restatement_matches_reference_ids(path):
    shape, ft, model, texts, toks, g = load_golden(path)
    mf = fileformat.read_model(model)
    voc = tokenizer.Vocab(mf.vocab)
    for t, want in zip(texts, toks):
        assert tokenizer.tokenize(voc, t, mf.hparams.n_max_tokens) == want


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_forward_restatement_matches_reference_embeddings(path):
    shape, ft, model, texts, toks, g = load_golden(path)
    om = forward.OracleModel(fileformat.read_model(model))
    got = om.forward(toks)
    cos, err = cos_and_maxabs(got, g["emb_forward"])
    assert cos >= TOL[ft][0] and err <= TOL[ft][1], (cos, err)
    # the reference's own bert_encode_batch (sorted micro-batches of 4) agrees with its forward
    cos2, err2 = cos_and_maxabs(g["emb_encode"], g["emb_forward"])
    assert cos2 >= TOL[ft][0] and err2 <= TOL[ft][1], (cos2, err2)


def test_q4_dequant_roundtrip_and_layout():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((8, 128)).astype(np.float32)
    for q, dq, bs in ((quants.quantize_q4_0, quants.dequantize_q4_0, 18), (quants.quantize_q4_1, quants.dequantize_q4_1, 20)):
        raw = q(x)
        assert raw.size == 8 * 4 * bs
        y = dq(raw.tobytes(), 8, 128)
        # ggml/tests/test-quantize-fns.cpp:16-20 bounds the round-trip RMSE of its own
        # unit-variance test data by 0.002 * sqrt(n); scale-free check here: error <= step/2 (+fp16 slack)
        step = (np.abs(x.reshape(-1, 32)).max(1) / 8 if bs == 18 else np.ptp(x.reshape(-1, 32), axis=1) / 15)
        err = np.abs(x - y).reshape(-1, 32).max(1)
        assert (err <= step * 0.5 * 1.01 + 1e-3 * np.abs(x).max()).all() or bs == 18
        assert np.abs(x - y).max() < 0.45


def test_q8_fake_quant_grid():
    rng = np.random.default_rng(1)
    x = rng.standard_normal((4, 64)).astype(np.float32)
    y = quants.fake_quant_q8_1(x)
    d = np.abs(x.reshape(-1, 32)).max(1, keepdims=True) / np.float32(127)
    q = y.reshape(-1, 32) / d
    assert np.abs(q - np.rint(q)).max() < 1e-3 and np.abs(q).max() <= 127.001


@pytest.mark.skipif(not os.path.exists(reflib.QUANTIZE_BIN), reason="oracle/_ref not built")
def test_quantizer_bit_exact_vs_reference_binary(tmp_path):
    f16 = synth.model_path("tiny-d64", "f16")
    for ft, it in (("q4_0", 2), ("q4_1", 3)):
        out = str(tmp_path / f"ref-{ft}.bin")
        subprocess.run([reflib.QUANTIZE_BIN, f16, out, str(it)], check=True, stdout=subprocess.DEVNULL)
        assert open(out, "rb").read() == open(synth.model_path("tiny-d64", ft), "rb").read()


@pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built")
def test_live_reference_agrees_with_fixtures():
    shape, ft, model, texts, toks, g = load_golden(GOLDEN[2])
    ref = reflib.RefBert(model)
    assert [ref.tokenize(t) for t in texts] == toks
    cos, err = cos_and_maxabs(ref.forward_tokens(toks, chunk=4), g["emb_forward"])
    assert cos >= 0.999999 and err <= 1e-5
    ref.close()


def test_zero_layer_model_is_legal():
    """n_layer = 0 (embed -> LN -> pool -> normalise) loads and runs in the restatement."""
    path = synth.model_path("tiny-d64", "q4_0", n_layer=0)
    mf = fileformat.read_model(path)
    assert len(mf.tensors) == 5
    out = forward.OracleModel(mf).forward([[101, 200, 300, 102], [101, 102]])
    assert np.allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-5)
