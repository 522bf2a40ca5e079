"""GPU suite (-m gpu): parity of the CUDA path against the oracle, through the C ABI.

Acceptance bar (BASELINE.json north_star): per-embedding cosine >= 0.9999 against the
reference's ggml CPU path on the same weights and inputs, with a stated max-abs error per
weight type.  The golden fixtures were produced by the compiled reference; where
oracle/_ref/libbert_ref.so travelled to this box the reference itself is also run live.
"""
import os

import numpy as np
import pytest

from conftest import cos_and_maxabs, has_gpu
from oracle import fileformat, forward, reflib, tokenizer
from test_oracle import GOLDEN, load_golden
from tools import synth

pytestmark = pytest.mark.gpu

# stated max-abs error per weight type (unit-norm embeddings), cosine floor 0.9999 for all
MAX_ABS = {"f32": 1.5e-3, "f16": 1.5e-3, "q4_0": 8e-3, "q4_1": 8e-3}
COS_MIN = 0.9999


@pytest.fixture(scope="module", autouse=True)
def _need_gpu():
    assert has_gpu(), "the -m gpu suite needs a B200; the product has no CPU path"


@pytest.fixture(scope="module")
def models(pkg):
    cache = {}

    def get(path):
        if path not in cache:
            cache[path] = pkg.BertModel(path)
        return cache[path]
    yield get
    for m in cache.values():
        m.close()


# ------------------------------------------------------------------ kernel-level self-tests
@pytest.mark.parametrize("wt,M,N,K,epi", [
    (1, 128, 64, 64, 0), (1, 300, 192, 256, 5), (1, 1000, 768, 768, 3), (1, 4096, 3072, 768, 3),
    (2, 128, 64, 64, 0), (2, 300, 256, 256, 1), (2, 1000, 768, 768, 7), (2, 4096, 768, 3072, 5),
    (3, 128, 128, 128, 1), (3, 2048, 1024, 512, 7), (2, 131, 2304, 768, 1), (3, 77, 1152, 384, 1),
    (1, 1000, 768, 768, 35), (2, 900, 3072, 768, 35),          # 32 = EPI_GELU_H2: GELU in packed half arithmetic
])
def test_tcgen05_gemm_matches_simt_reference(pkg, wt, M, N, K, epi):
    rc, err, ref = pkg.selftest_gemm(M, N, K, wt, epi, seed=M + N)
    assert rc == 0
    assert np.isfinite(err) and err <= 4e-3 * max(ref, 1.0), (err, ref)       # fp16 output rounding


@pytest.mark.parametrize("wt,M,N,K,epi", [(2, 515, 192, 384, 11), (3, 777, 1024, 512, 15), (2, 640, 3072, 768, 11),
                                           (2, 640, 3072, 768, 43), (3, 515, 1536, 384, 43)])
def test_tcgen05_gemm_with_q8_fakequant_epilogue(pkg, wt, M, N, K, epi):
    rc, err, ref = pkg.selftest_gemm(M, N, K, wt, epi, seed=3)
    assert rc == 0
    # a value on a rounding boundary may land one q8 step (amax/127) apart in the two kernels
    assert np.isfinite(err) and err <= 1.1 * ref / 127 + 4e-3 * ref, (err, ref)


# EPI_LN = 16: LayerNorm of (acc + bias + residual) inside the GEMM epilogue, statistics exchanged between the
# CTAs of a cluster through DSMEM; checked against validation GEMM + the stand-alone LayerNorm row kernel
@pytest.mark.parametrize("wt,M,N,K,epi", [
    (1, 600, 768, 768, 21), (2, 1000, 768, 768, 29), (3, 777, 768, 3072, 29), (2, 515, 512, 1536, 29),
    (2, 300, 256, 256, 21), (1, 131, 128, 512, 29), (2, 70000, 768, 768, 29), (2, 600, 1024, 1024, 29), (1, 300, 1024, 512, 21),
    (2, 900, 384, 384, 29), (2, 1300, 384, 1536, 29), (1, 700, 384, 384, 21), (1, 515, 384, 1536, 29), (3, 40000, 384, 1536, 29),   # three 128-wide blocks
])
def test_tcgen05_gemm_with_fused_layernorm_epilogue(pkg, wt, M, N, K, epi):
    rc, err, ref = pkg.selftest_gemm(M, N, K, wt, epi, seed=5)
    assert rc == 0
    tol = 4e-3 * max(ref, 1.0) + (1.1 * ref / 127 if epi & 8 else 0.0)
    assert np.isfinite(err) and err <= tol, (err, ref)


@pytest.mark.parametrize("wt,M,N,K", [(1, 600, 768, 768), (2, 515, 2304, 768), (3, 777, 512, 2048), (2, 1000, 384, 1536)])
def test_tcgen05_gemm_against_numpy(pkg, wt, M, N, K):
    """Per-kernel check against host arithmetic, not against another CUDA kernel: bertx_debug_gemm hands back the fp16 A
    operand and the weights as decoded from the 4-bit planes (f32; the mainloop rounds d * q + m to fp16 once); numpy forms A W^T in f32 (what ggml's mul_mat computes for
    these operands, ggml.c:11030-11258) and the tcgen05 kernel's fp16 output must match it to fp16 rounding."""
    import ctypes
    L = pkg.lib()
    L.bertx_debug_gemm.restype = ctypes.c_int32
    tc = np.zeros((1, M, N), dtype=np.float16)
    ms = np.zeros(1, dtype=np.float32)
    A = np.zeros((M, K), dtype=np.float16)
    W = np.zeros((N, K), dtype=np.float32)
    vp = lambda x: ctypes.c_void_p(x.ctypes.data)  # noqa: E731
    rc = L.bertx_debug_gemm(M, N, K, wt, 0, 5, 1, None, vp(tc), vp(ms), None, None, vp(A), vp(W))
    assert rc == 0
    want = A.astype(np.float32) @ W.astype(np.float16).astype(np.float32).T     # the B operand is the fp16-rounded weight
    got = tc[0].astype(np.float32)
    assert np.isfinite(got).all() and np.abs(want).max() > 0.5
    assert np.all(np.abs(got - want) <= 2e-3 * np.maximum(1.0, np.abs(want))), float(np.abs(got - want).max())


ATTN_SHAPES = [(5, 64, 2, 64), (7, 200, 4, 64), (3, 512, 12, 64), (9, 130, 12, 32), (4, 512, 2, 32), (33, 17, 12, 64),
               (1, 1, 2, 32), (2, 2, 2, 64), (6, 97, 3, 64), (40, 384, 12, 64),
               (200, 30, 12, 32), (64, 70, 12, 32), (300, 9, 6, 32), (150, 128, 4, 32),     # many short sentences: packed spans
               (200, 30, 12, 64), (64, 70, 4, 64), (300, 9, 6, 64), (150, 128, 4, 64)]


@pytest.mark.parametrize("n_sent,L,nh,d", ATTN_SHAPES)
@pytest.mark.parametrize("flags", [0, 1024, 2048, 2048 + 1024, 4096, 4096 + 1024, 8192, 8192 + 1024, 4096 + 16384, 2048 + 16384],
                         ids=["default", "default-ramp", "tc1", "tc1-ramp", "8warp", "8warp-ramp", "3buf", "3buf-ramp", "8warp-unpacked", "tc1-unpacked"])
def test_fused_attention_matches_reference_kernel(pkg, n_sent, L, nh, d, flags):
    """Seed bit 10: key rows grow along the sentence, so the running max keeps leaving the 2^8 window (the rescale / redo path);
    bit 11: attention_tc.cu (head size 64; head size 32 always runs attention_tc2.cu); bit 12: attention_tc2.cu; bit 13: attention_tc3.cu; none: the
    engine's dispatch.  attention_tc.cu and attention_tc2.cu pack short sentences into spans of <= 128 tokens as the engine does
    (bit 14: one work item per sentence)."""
    rc, err, ref = pkg.selftest_attention(n_sent, L, nh, d, seed=(L % 1024) | flags)
    assert rc == 0
    assert np.isfinite(err) and err <= 6e-3 * max(ref, 1.0), (err, ref)


ATTN_NP_CASES = [([5, 64, 1, 130, 17], 4, 64), ([33] * 20 + [7] * 30, 12, 32), ([512, 3], 2, 64), ([200, 100, 60, 60], 6, 32)]


@pytest.mark.parametrize("lens,nh,d", ATTN_NP_CASES, ids=["ragged-d64", "short-d32", "long-d64", "mixed-d32"])
@pytest.mark.parametrize("flags", [0, 16384], ids=["packed", "unpacked"])
def test_attention_against_numpy(pkg, lens, nh, d, flags):
    """The production attention kernels against an independent numpy soft-max attention on the same fp16 inputs (per sentence
    and head: scores / sqrt(d), soft-max over the sentence's own keys, weighted sum of V; reference bert.cpp:1018-1036) -- not
    against another CUDA kernel of this repo.  Tolerance: the fp16 rounding of P and of the output, as in the self-test."""
    qkv, out = pkg.debug_attention(lens, nh, d, seed=11, flags=flags)
    q, k, v = (qkv[:, i, :].astype(np.float32) for i in range(3))
    want = np.zeros(out.shape, np.float32)
    t0 = 0
    for L in lens:
        for h in range(nh):
            sl = slice(h * d, (h + 1) * d)
            s = q[t0:t0 + L, sl] @ k[t0:t0 + L, sl].T / np.sqrt(np.float32(d))
            p = np.exp(s - s.max(axis=1, keepdims=True))
            want[t0:t0 + L, sl] = (p / p.sum(axis=1, keepdims=True)) @ v[t0:t0 + L, sl]
        t0 += L
    got = out.astype(np.float32)
    assert np.isfinite(got).all() and np.abs(want).max() > 0.5
    assert np.abs(got - want).max() <= 6e-3 * max(float(np.abs(want).max()), 1.0), float(np.abs(got - want).max())


# ------------------------------------------------------------------ end-to-end parity
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_encode_matches_reference_golden_embeddings(models, path):
    shape, ft, model, texts, toks, g = load_golden(path)
    m = models(model)
    assert [m.tokenize(t) for t in texts] == toks
    got = m.encode(texts, batch_size=4)
    for want in (g["emb_encode"], g["emb_forward"]):
        cos, err = cos_and_maxabs(got, want)
        assert cos >= COS_MIN and err <= MAX_ABS[ft], (ft, cos, err)
    got2 = m.forward_tokens(toks)
    cos, err = cos_and_maxabs(got2, g["emb_forward"])
    assert cos >= COS_MIN and err <= MAX_ABS[ft], (ft, cos, err)
    # batch-composition invariance: packing replaces padding, so only rounding may differ.  (Short sentences
    # share attention work items, so where a sentence's keys fall in the 64-key blocks -- and with it the reference max its
    # fp16 P values are rounded against -- depends on its neighbours: differences at the level of that rounding, 1e-4.)
    got3 = np.concatenate([m.forward_tokens([t]) for t in toks[:6]])
    assert np.abs(got3 - got2[:6]).max() <= (3e-4 if ft in ("f32", "f16") else 4e-3)
    assert np.array_equal(m.forward_tokens(toks[:5], fake_batch=True), m.forward_tokens(toks[:5]))


@pytest.mark.parametrize("ft", ["f16", "q4_0", "q4_1"])
def test_edge_cases_and_error_conventions(models, ft, capfd):
    model = synth.model_path("tiny-d64", ft)
    mf = fileformat.read_model(model)
    m = models(model)
    om = forward.OracleModel(mf)
    vocab = mf.vocab
    # empty text -> [CLS][SEP]; single word; exactly n_max_tokens ids; ragged mix in one call
    texts = [b"", b"hello", synth.texts_exact_len(vocab, 1, 512, seed=5)[0], b"hello world !",
             synth.texts_exact_len(vocab, 1, 511, seed=6)[0], synth.texts_exact_len(vocab, 1, 129, seed=7)[0]]
    voc = tokenizer.Vocab(vocab)
    toks = [tokenizer.tokenize(voc, t, 512) for t in texts]
    assert [len(t) for t in toks] == [2, 3, 512, 5, 511, 129]
    got = m.encode(texts)
    want = om.forward(toks)
    cos, err = cos_and_maxabs(got, want)
    assert cos >= COS_MIN and err <= MAX_ABS[ft], (cos, err)
    assert np.allclose(np.linalg.norm(got, axis=1), 1.0, atol=1e-4)
    # a single string instead of a list: one vector, as the reference's Python binding returns (sample_dylib.py:42-62)
    one = m.encode("hello world !")
    assert one.shape == (got.shape[1],) and np.array_equal(one, m.encode([b"hello world !"])[0])
    # too many tokens: message on stderr, outputs untouched (reference bert.cpp:867-871)
    before = m.kernel_launches()
    out = m.forward_tokens([[101] + [200] * 600 + [102], [101, 102]])
    assert (out == 0).all() and m.kernel_launches() == before
    assert "Too many tokens" in capfd.readouterr().err
    # over-long TEXT is truncated by the tokenizer (documented deviation) and still encodes
    long_text = b" ".join([b"hello"] * 700)
    e = m.encode([long_text])
    assert np.isfinite(e).all() and abs(np.linalg.norm(e[0]) - 1.0) < 1e-4


def test_zero_layer_model(pkg):
    """embed -> LN -> pool -> normalise only: isolates K0 + K4 against the oracle."""
    for ft in ("f32", "f16", "q4_0", "q4_1"):
        model = synth.model_path("tiny-d64", ft, n_layer=0)
        mf = fileformat.read_model(model)
        m = pkg.BertModel(model)
        toks = [[101, 300, 301, 302, 102], [101, 102], [101] + list(range(200, 460)) + [102]]
        got = m.forward_tokens(toks)
        want = forward.OracleModel(mf).forward(toks)
        cos, err = cos_and_maxabs(got, want)
        assert cos >= 0.999999 and err <= 2e-4, (ft, cos, err)
        m.close()


@pytest.mark.parametrize("ft", ["f16", "q4_0"])
def test_cls_pooling_extension(pkg, ft):
    """Opt-in [CLS] pooling (bertx_set_pooling / BERT_POOLING=cls; north_star (e)).  The reference has
    no such mode, so the check is the numpy restatement with row 0 in place of the masked mean; switching back
    restores the reference's mean bit for bit."""
    model = synth.model_path("tiny-d64", ft)
    mf = fileformat.read_model(model)
    m = pkg.BertModel(model)
    om = forward.OracleModel(mf)
    toks = [[101, 300, 301, 302, 102], [101, 102], [101] + list(range(200, 330)) + [102]]
    mean0 = m.forward_tokens(toks)
    assert m.set_pooling("cls") == "mean"
    got = m.forward_tokens(toks)
    want = om.forward(toks, pooling="cls")
    cos, err = cos_and_maxabs(got, want)
    assert cos >= COS_MIN and err <= MAX_ABS[ft], (cos, err)
    assert np.allclose(np.linalg.norm(got, axis=1), 1.0, atol=1e-4)
    assert np.abs(got - mean0).max() > 1e-3                      # it really is a different pooling
    assert m.set_pooling("mean") == "cls"
    assert np.array_equal(m.forward_tokens(toks), mean0)
    m.close()


def test_staged_job_equals_encode_and_multi_microbatch(pkg):
    """Device-resident path (bench `value`) gives the same embeddings as bert_encode_batch,
    also when the job is cut into several micro-batches."""
    os.environ["BERT_MICROBATCH_TOKENS"] = "1024"
    try:
        model = synth.model_path("tiny-d32", "q4_0")
        mf = fileformat.read_model(model)
        m = pkg.BertModel(model)
        texts = synth.texts_var_len(mf.vocab, 150, 2, 120, seed=4, cjk=True)
        ids, lens = m.tokenize_batch(texts)
        assert int(lens.sum()) > 4 * 1024
        m.stage(ids, lens)
        ms, staged = m.run_staged(reps=2, fetch=True)
        enc = m.encode(texts)
        # the two paths cut the job into different micro-batches; with short sentences packed into shared attention work
        # items (head size 32) a sentence's neighbours decide where its keys fall in the 64-key blocks, so the results agree
        # to the rounding of the fp16 P values / one q8 step, not bit for bit ...
        assert ms > 0 and np.abs(staged - enc).max() <= 4e-3
        assert np.array_equal(m.encode(texts), enc)                  # ... while the same call is bit-identical from run to run
        prof = m.profile_staged()
        assert prof["gemm"][1] > 0 and prof["attention"][1] > 0 and m.kernel_launches() > 0
        m.close()
        # ... and without the packing every sentence is computed on its own: any cut gives the same bits
        os.environ["BERT_ATTN_PACK"] = "0"
        try:
            m = pkg.BertModel(model)
            m.stage(ids, lens)
            _, staged = m.run_staged(reps=1, fetch=True)
            assert np.array_equal(staged, m.encode(texts))
            m.close()
        finally:
            del os.environ["BERT_ATTN_PACK"]
    finally:
        del os.environ["BERT_MICROBATCH_TOKENS"]
    m2 = pkg.BertModel(model)
    one = m2.encode(texts)
    cos, err = cos_and_maxabs(one, enc)
    assert cos >= 0.99999 and err <= 4e-3
    m2.close()


@pytest.mark.skipif(not reflib.available(), reason="oracle/_ref did not travel to this box")
@pytest.mark.parametrize("shape,ft,n,lo,hi", [
    ("all-MiniLM-L6-v2", "q4_0", 24, 3, 70),
    ("bge-small-en-v1.5", "f16", 16, 100, 128),
    ("bge-small-zh-v1.5", "q4_1", 24, 2, 256),
    ("bge-base-en-v1.5", "q4_0", 6, 200, 512),
    ("bge-large-en-v1.5", "q4_0", 4, 120, 256),
])
def test_named_model_shapes_against_live_reference(pkg, shape, ft, n, lo, hi):
    """BASELINE.json configs at the named shapes (synthetic weights), live compiled reference."""
    model = synth.model_path(shape, ft)
    vocab = synth.make_vocab(synth.SHAPES[shape][4], 0)
    texts = synth.texts_var_len(vocab, n, lo, hi, seed=9, cjk="zh" in shape)
    if shape.startswith("bge-base"):
        texts[0] = synth.texts_exact_len(vocab, 1, 512, seed=1)[0]
    m = pkg.BertModel(model)
    got = m.encode(texts)
    ref = reflib.RefBert(model)
    toks = [ref.tokenize(t) for t in texts]
    assert [m.tokenize(t) for t in texts] == toks
    want = ref.forward_tokens(toks, chunk=2 if hi > 256 else 8)
    cos, err = cos_and_maxabs(got, want)
    print(f"\n{shape} {ft}: min cosine {cos:.7f} max abs {err:.2e}")
    assert cos >= COS_MIN and err <= MAX_ABS[ft], (cos, err)
    m.close()
    ref.close()


PARITY_LOG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "parity_r2.jsonl")


def _record(row):
    """One line per parity case for the table in DESIGN.md section 4 (written on the GPU box, merged back by gpurun)."""
    import json
    try:
        os.makedirs(os.path.dirname(PARITY_LOG), exist_ok=True)
        with open(PARITY_LOG, "a") as f:
            f.write(json.dumps(row) + "\n")
    except OSError:
        pass


# BASELINE.json's five configs at the sample sizes SURVEY 8(d) / BASELINE.md section 4 name, against the LIVE compiled
# reference (oracle/_ref/libbert_ref.so, the unmodified ggml CPU path), plus bge-base in f32 and f16 at 512 tokens.
CONFIG_PARITY = [
    # id, shape, ftype, text kind, sentences, tokens (0 = variable), reference chunk
    ("config1-minilm-q4_0-1738-lines", "all-MiniLM-L6-v2", "q4_0", "client", 1738, 0, 16),
    ("config2-bge-small-f16-256x128", "bge-small-en-v1.5", "f16", "exact", 256, 128, 8),
    ("config3-bge-base-q4_0-64x512", "bge-base-en-v1.5", "q4_0", "exact", 64, 512, 2),
    ("config4-bge-small-zh-q4_1-256-var", "bge-small-zh-v1.5", "q4_1", "cjkvar", 256, 0, 8),
    ("config5-bge-large-q4_0-32x256", "bge-large-en-v1.5", "q4_0", "exact", 32, 256, 4),
    ("bge-base-f32-8x512", "bge-base-en-v1.5", "f32", "exact", 8, 512, 2),
    ("bge-base-f16-8x512", "bge-base-en-v1.5", "f16", "exact", 8, 512, 2),
]


@pytest.mark.skipif(not reflib.available(), reason="oracle/_ref did not travel to this box")
@pytest.mark.parametrize("cid,shape,ft,kind,n,n_tok,chunk", CONFIG_PARITY, ids=[c[0] for c in CONFIG_PARITY])
def test_baseline_configs_against_live_reference(pkg, cid, shape, ft, kind, n, n_tok, chunk):
    import bench
    model = synth.model_path(shape, ft)
    texts = bench.make_texts(shape, kind, n, n_tok)
    m = pkg.BertModel(model)
    got = m.encode(texts)
    ref = reflib.RefBert(model)
    toks = [ref.tokenize(t) for t in texts]
    ids, lens = m.tokenize_batch(texts)
    assert ids.tolist() == [i for t in toks for i in t]                      # token ids bit-exact on the whole set
    order = np.argsort([len(t) for t in toks], kind="stable")                # the reference pads to the chunk's longest
    want = np.zeros_like(got)
    want[order] = ref.forward_tokens([toks[i] for i in order], chunk=chunk)
    cos, err = cos_and_maxabs(got, want)
    _record({"case": cid, "weights": ft, "sentences": n, "tokens": int(lens.sum()), "min_cosine": cos, "max_abs": err})
    print(f"\n{cid}: min cosine {cos:.7f} max abs {err:.2e}")
    assert cos >= COS_MIN and err <= MAX_ABS[ft], (cos, err)
    m.close()
    ref.close()


@pytest.mark.skipif(not reflib.available(), reason="oracle/_ref did not travel to this box")
@pytest.mark.parametrize("ft", ["f16", "q4_0", "q4_1"])
def test_outlier_channels_against_live_reference(pkg, ft):
    """Weights with a few hidden channels at 20-50x the typical LayerNorm gain / bias (what real BERT checkpoints have and
    the default synthetic weights do not): the fp16 residual stream and the q8 activation twin must still
    follow the reference's f32 activations."""
    shape = "bge-base-en-v1.5"
    model = synth.model_path(shape, ft, outliers=True)
    vocab = synth.make_vocab(synth.SHAPES[shape][4], 0)
    texts = synth.texts_var_len(vocab, 10, 8, 300, seed=31)
    m = pkg.BertModel(model)
    got = m.encode(texts)
    ref = reflib.RefBert(model)
    toks = [ref.tokenize(t) for t in texts]
    want = ref.forward_tokens(toks, chunk=2)
    self_check = ref.forward_tokens(toks[:4], chunk=1)                       # the reference against itself, other batching
    cos, err = cos_and_maxabs(got, want)
    scos, serr = cos_and_maxabs(self_check, want[:4])
    _record({"case": "outlier-channels-bge-base", "weights": ft, "sentences": len(texts), "min_cosine": cos, "max_abs": err,
             "reference_self_min_cosine": scos, "reference_self_max_abs": serr})
    print(f"\noutlier channels {ft}: min cosine {cos:.7f} max abs {err:.2e} (reference vs itself: {scos:.7f} / {serr:.2e})")
    assert np.isfinite(got).all()
    assert cos >= COS_MIN and err <= MAX_ABS[ft], (cos, err)
    m.close()
    ref.close()


def test_full_size_properties_bge_base(pkg):
    """Size-independent properties at BASELINE's full size (512 x 512 tokens): unit norm,
    permutation equivariance of the batch, duplicate sentences give identical rows."""
    model = synth.model_path("bge-base-en-v1.5", "q4_0")
    vocab = synth.make_vocab(30522, 0)
    texts = synth.texts_exact_len(vocab, 512, 512, seed=0)
    texts[7] = texts[3]
    m = pkg.BertModel(model)
    e = m.encode(texts)
    assert np.isfinite(e).all() and np.allclose(np.linalg.norm(e, axis=1), 1.0, atol=2e-4)
    assert np.array_equal(e[7], e[3]) or np.abs(e[7] - e[3]).max() < 1e-6
    perm = np.random.default_rng(0).permutation(64)
    e2 = m.encode([texts[i] for i in perm])
    assert np.abs(e2 - e[perm]).max() <= 4e-3
    # run-to-run determinism at full size (many tiles per persistent CTA: a barrier / TMEM ordering bug shows up here
    # long before it shows in the small kernel self-tests)
    for _ in range(2):
        assert np.array_equal(m.encode(texts), e)
    m.close()


def test_full_size_determinism_under_perturbed_builds(pkg, tmp_path):
    """The round-2 mainloop race (profiles/r2_gemm_race.md) only showed when the timing of the GEMM CTA shifted, so
    the full-size check also runs on three timing-perturbed builds of the kernel (build.py PERTURBED: one epilogue TMEM load
    at a time / every dequant thread arrives itself / rings one stage shorter) and with PDL off.  None of them changes
    the arithmetic: every build must give the default build's embeddings bit for bit, twice in a row."""
    import subprocess
    import sys
    libs = {"default": pkg.LIB_PATH}
    for path in pkg.build_perturbed():
        libs[os.path.basename(path)[len("libbert_"):-3]] = path
    script = (
        "import sys, numpy as np\n"
        f"sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})\n"
        "import __graft_entry__ as ge\n"
        "from tools import synth\n"
        "pkg = ge.load_package()\n"
        "m = pkg.BertModel(synth.model_path('bge-base-en-v1.5', 'q4_0'))\n"
        "texts = synth.texts_exact_len(synth.make_vocab(30522, 0), 512, 512, seed=0)\n"
        "e = m.encode(texts)\n"
        "assert np.array_equal(m.encode(texts), e), 'run-to-run difference'\n"
        "np.save(sys.argv[1], e)\n")
    runs = [(name, {}) for name in libs] + [("default", {"BERT_PDL": "0"})]
    want = None
    for i, (name, extra) in enumerate(runs):
        out = str(tmp_path / f"e{i}.npy")
        env = dict(os.environ, BERT_B200_LIB=libs[name], **extra)
        r = subprocess.run([sys.executable, "-c", script, out], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (name, extra, r.stdout[-400:], r.stderr[-800:])
        e = np.load(out)
        assert np.isfinite(e).all()
        if want is None:
            want = e
        else:
            assert np.array_equal(e, want), (name, extra, float(np.abs(e - want).max()))


def test_in_process_sharding_over_all_visible_gpus(pkg):
    """One process, every visible GPU (what a drop-in caller of bert_encode_batch gets on a multi-GPU box):
    micro-batches are dealt to the replicas with no collective; results equal the single-GPU ones bit for bit."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run under gpurun --gpus 2)")
    model = synth.model_path("tiny-d64", "q4_0")
    mf = fileformat.read_model(model)
    texts = synth.texts_var_len(mf.vocab, 700, 2, 200, seed=21, cjk=True)
    old = {k: os.environ.get(k) for k in ("BERT_CUDA_DEVICES", "BERT_MICROBATCH_TOKENS")}
    os.environ["BERT_MICROBATCH_TOKENS"] = "2048"
    try:
        os.environ["BERT_CUDA_DEVICES"] = "0"
        m1 = pkg.BertModel(model)
        assert m1.n_devices() == 1
        one = m1.encode(texts)
        m1.close()
        del os.environ["BERT_CUDA_DEVICES"]
        mN = pkg.BertModel(model)
        assert mN.n_devices() == torch.cuda.device_count()
        many = mN.encode(texts)
        again = mN.encode(texts[:5])
        mN.close()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    assert np.isfinite(many).all() and np.allclose(np.linalg.norm(many, axis=1), 1.0, atol=2e-4)
    # chunking differs (the chunk target scales with the GPU count), so micro-batch composition differs: rounding only
    assert np.abs(many - one).max() <= 4e-3
    assert np.abs(again - one[:5]).max() <= 4e-3


@pytest.mark.parametrize("shape", ["tiny-d64", "all-MiniLM-L6-v2"])
def test_gpu_quantizer_writes_the_reference_quantizers_bytes(pkg, shape, tmp_path):
    """bertx_quantize_file (CUDA kernel) == models/quantize.cpp output, byte for byte: against the files the
    numpy restatement writes (itself pinned bit-exact to the reference binary, tests/test_oracle.py) and, where
    oracle/_ref/quantize travelled, against the reference binary run live (f16 and f32 sources)."""
    import subprocess
    f16 = synth.model_path(shape, "f16")
    for itype, ft in ((2, "q4_0"), (3, "q4_1")):
        out = str(tmp_path / f"{ft}.bin")
        pkg.quantize_file(f16, out, itype)
        assert open(out, "rb").read() == open(synth.model_path(shape, ft), "rb").read(), (shape, ft)
    if os.path.exists(reflib.QUANTIZE_BIN) and shape == "tiny-d64":
        f32 = synth.model_path(shape, "f32")
        for src in (f16, f32):
            for itype in (2, 3):
                a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
                subprocess.run([reflib.QUANTIZE_BIN, src, a, str(itype)], check=True, capture_output=True)
                pkg.quantize_file(src, b, itype)
                assert open(a, "rb").read() == open(b, "rb").read(), (src, itype)
    # the quantised file loads and encodes
    m = pkg.BertModel(str(tmp_path / "q4_1.bin"))
    e = m.encode([b"hello world"])
    assert np.isfinite(e).all() and abs(np.linalg.norm(e[0]) - 1.0) < 1e-4
    m.close()
