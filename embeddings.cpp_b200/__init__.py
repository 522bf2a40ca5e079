"""embeddings.cpp_b200 -- B200-native BERT embedding engine behind the reference's bert.h ABI.

The product is ``lib/libbert.so`` (hand-written CUDA for sm_100a + C++ host runtime, built
by ``build.py``).  This module is the Python mirror of the reference's ctypes consumer
(/root/reference/examples/sample_dylib.py:15-62, benchmarks/run_mteb.py:34-83): the same
``BertModel(fname).encode(sentences, batch_size)`` surface bound to the same C symbols.
There is no CPU fallback: loading fails loudly when the library or a B200 is missing.

The directory name contains a dot, so import it through the ``load_package()`` helper in
``__graft_entry__.py`` (importlib by path) rather than ``import embeddings.cpp_b200``.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BERT_B200_LIB") or os.path.join(HERE, "lib", "libbert.so")   # env: A/B runs of two builds (tools/ab.py)

_lib = None


def build(force=False):
    import importlib.util
    spec = importlib.util.spec_from_file_location("_b200_build", os.path.join(HERE, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build(force=force)


def build_perturbed(force=False):
    """The timing-perturbed builds of the GEMM kernel (build.py PERTURBED) -> list of library paths."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_b200_build", os.path.join(HERE, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.build_perturbed(force=force)


def lib():
    """The loaded C-ABI library with argtypes set (raises if it has not been built)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python embeddings.cpp_b200/build.py` "
                           "(there is no fallback implementation)")
    L = ctypes.cdll.LoadLibrary(LIB_PATH)
    c_i32, c_i64, vp, cp = ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p, ctypes.c_char_p
    pf, pi = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32)
    L.bert_load_from_file.restype = vp
    L.bert_load_from_file.argtypes = [cp]
    L.bertx_load_host_only.restype = vp
    L.bertx_load_host_only.argtypes = [cp]
    L.bert_free.argtypes = [vp]
    L.bert_n_embd.restype = c_i32
    L.bert_n_embd.argtypes = [vp]
    L.bert_n_max_tokens.restype = c_i32
    L.bert_n_max_tokens.argtypes = [vp]
    L.bert_vocab_id_to_token.restype = cp
    L.bert_vocab_id_to_token.argtypes = [vp, c_i32]
    L.bert_encode.argtypes = [vp, c_i32, cp, pf]
    L.bert_encode_batch.argtypes = [vp, c_i32, c_i32, c_i32, ctypes.POINTER(cp), ctypes.POINTER(pf)]
    L.bert_tokenize.argtypes = [vp, cp, pi, pi, c_i32]
    L.bert_forward.argtypes = [vp, c_i32, pi, c_i32, pf]
    L.bert_forward_batch.argtypes = [vp, c_i32, c_i32, ctypes.POINTER(pi), pi, ctypes.POINTER(pf)]
    L.bert_forward_fake_batch.argtypes = L.bert_forward_batch.argtypes
    L.bertx_n_devices.restype = c_i32
    L.bertx_n_devices.argtypes = [vp]
    L.bertx_hparam.restype = c_i32
    L.bertx_hparam.argtypes = [vp, c_i32]
    L.bertx_set_pooling.restype = c_i32
    L.bertx_set_pooling.argtypes = [vp, c_i32]
    L.bertx_kernel_launches.restype = c_i64
    L.bertx_kernel_launches.argtypes = [vp]
    L.bertx_plan.restype = c_i32
    L.bertx_plan.argtypes = [vp, c_i32, c_i32, pi, pi, pi, c_i32]
    L.bertx_pack_spans.restype = c_i32
    L.bertx_pack_spans.argtypes = [c_i32, pi, pi]
    L.bertx_tokenize_batch.restype = c_i64
    L.bertx_tokenize_batch.argtypes = [vp, c_i32, c_i32, ctypes.POINTER(cp), pi, c_i64, pi]
    L.bertx_stage.restype = c_i32
    L.bertx_stage.argtypes = [vp, c_i32, c_i32, pi, pi]
    L.bertx_run_staged.restype = c_i32
    L.bertx_run_staged.argtypes = [vp, c_i32, c_i32, pf, pf]
    L.bertx_profile_staged.restype = c_i32
    L.bertx_profile_staged.argtypes = [vp, c_i32, pf, ctypes.POINTER(c_i64)]
    L.bertx_selftest_gemm.restype = c_i32
    L.bertx_selftest_gemm.argtypes = [c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, pf, pf]
    L.bertx_bench_gemm.restype = c_i32
    L.bertx_bench_gemm.argtypes = [c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, pf]
    L.bertx_bench_attention.restype = c_i32
    L.bertx_bench_attention.argtypes = [c_i32, c_i32, c_i32, c_i32, c_i32, pf]
    L.bertx_selftest_attention.restype = c_i32
    L.bertx_selftest_attention.argtypes = [c_i32, c_i32, c_i32, c_i32, c_i32, pf, pf]
    L.bertx_debug_attention.restype = c_i32
    L.bertx_debug_attention.argtypes = [c_i32, pi, c_i32, c_i32, c_i32, c_i32, vp, vp]
    L.bertx_quantize_file.restype = c_i32
    L.bertx_quantize_file.argtypes = [cp, cp, c_i32]
    L.bertx_quantize_rows_noalloc.restype = c_i32
    L.bertx_quantize_rows_noalloc.argtypes = [vp, c_i32, c_i64, c_i32, vp]
    L.bertx_convert_safetensors.restype = c_i32
    L.bertx_convert_safetensors.argtypes = [cp, cp, cp, c_i32, cp]
    L.bertx_load_from_safetensors.restype = vp
    L.bertx_load_from_safetensors.argtypes = [cp, cp, cp, c_i32]
    _lib = L
    return L


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _row_ptrs(out):
    """float** over the rows of a 2-D float32 array: the addresses are computed in numpy (one vector op) instead
    of one ctypes cast per row (examples/sample_dylib.py:50-51 does the latter; 1738 rows cost milliseconds)."""
    addr = out.ctypes.data + np.arange(out.shape[0], dtype=np.uint64) * np.uint64(out.strides[0])
    out_ptrs = np.ascontiguousarray(addr, dtype=np.uint64)
    p = ctypes.cast(out_ptrs.ctypes.data, ctypes.POINTER(ctypes.POINTER(ctypes.c_float)))
    p._keep = out_ptrs                      # the pointer table must outlive the call
    return p


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))


def _text_ptrs(sentences):
    """const char** over the texts.  Many short texts: they are joined into ONE NUL-separated buffer and the pointer table is
    base + offsets, computed in numpy -- `(c_char_p * n)(*texts)` converts item by item (348 us for 1738 short lines, a tenth
    of that call).  Long texts (or few): the item-by-item table, which does not copy the bytes."""
    n = len(sentences)
    probe = sentences[:16]
    if n < 256 or sum(len(s) for s in probe) > 256 * len(probe):
        return (ctypes.c_char_p * n)(*[s if isinstance(s, bytes) else s.encode("utf-8") for s in sentences])
    try:
        buf = b"\0".join(sentences) + b"\0"
        bs = sentences
    except TypeError:                                              # some (or all) are str
        bs = [s if isinstance(s, bytes) else s.encode("utf-8") for s in sentences]
        buf = b"\0".join(bs) + b"\0"
    offs = np.zeros(n, dtype=np.uint64)
    lens = np.fromiter(map(len, bs), dtype=np.uint64, count=n)
    np.cumsum(lens[:-1] + np.uint64(1), out=offs[1:])
    addr = offs + np.uint64(ctypes.cast(ctypes.c_char_p(buf), ctypes.c_void_p).value)
    p = ctypes.cast(addr.ctypes.data, ctypes.POINTER(ctypes.c_char_p))
    p._keep = (addr, buf)                                          # table and text bytes must outlive the call
    return p


class BertModel:
    """Same surface as the reference's examples/sample_dylib.py BertModel."""

    def __init__(self, fname, host_only=False, n_threads=None, _ctx=None):
        self.lib = lib()
        if _ctx is not None:
            self.ctx = _ctx
        else:
            fn = fname.encode() if isinstance(fname, str) else fname
            self.ctx = (self.lib.bertx_load_host_only if host_only else self.lib.bert_load_from_file)(fn)
        if not self.ctx:
            raise RuntimeError(f"bert_load_from_file('{fname}') failed (see stderr)")
        self.n_embd = self.lib.bert_n_embd(self.ctx)
        self.n_max_tokens = self.lib.bert_n_max_tokens(self.ctx)
        self.n_threads = n_threads or min(os.cpu_count() or 1, 32)

    @classmethod
    def from_safetensors(cls, model_dir, ftype=2, n_threads=None):
        """Load a Hugging Face checkpoint directory (model.safetensors, config.json, vocab.txt) without an
        intermediate model file; ftype 0 f32, 1 f16, 2 q4_0, 3 q4_1 (bertx_load_from_safetensors)."""
        L = lib()
        ctx = L.bertx_load_from_safetensors(*[os.fsencode(os.path.join(model_dir, f))
                                              for f in ("model.safetensors", "config.json", "vocab.txt")], ftype)
        if not ctx:
            raise RuntimeError(f"bertx_load_from_safetensors('{model_dir}') failed (see stderr)")
        return cls(model_dir, n_threads=n_threads, _ctx=ctx)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.bert_free(self.ctx)
            self.ctx = None

    # -- reference surface -----------------------------------------------------------
    def encode(self, sentences, batch_size=16):
        """bert_encode_batch; returns [n, n_embd] float32 -- or, like the reference's binding, one [n_embd] vector when
        `sentences` is a single string (sample_dylib.py:42-62)."""
        single = isinstance(sentences, (str, bytes))
        if single:
            sentences = [sentences]
        n = len(sentences)
        out = np.zeros((n, self.n_embd), dtype=np.float32)
        if n == 0:
            return out
        self.lib.bert_encode_batch(self.ctx, self.n_threads, batch_size, n, _text_ptrs(sentences), _row_ptrs(out))
        return out[0] if single else out

    def tokenize(self, text, n_max=None):
        n_max = self.n_max_tokens if n_max is None else n_max
        buf = np.zeros(max(n_max, 1) + 8, dtype=np.int32)
        n = ctypes.c_int32(0)
        self.lib.bert_tokenize(self.ctx, text if isinstance(text, bytes) else text.encode("utf-8"),
                               _ip(buf), ctypes.byref(n), n_max)
        return buf[:n.value].tolist()

    def id_to_token(self, i):
        return self.lib.bert_vocab_id_to_token(self.ctx, i)

    def forward_tokens(self, batch_tokens, fake_batch=False):
        """bert_forward_batch on lists of ids; returns [n, n_embd] float32."""
        n = len(batch_tokens)
        out = np.zeros((n, self.n_embd), dtype=np.float32)
        arrs = [np.ascontiguousarray(t, dtype=np.int32) for t in batch_tokens]
        ptrs = (ctypes.POINTER(ctypes.c_int32) * n)(*[_ip(a) for a in arrs])
        lens = np.array([len(a) for a in arrs], dtype=np.int32)
        fn = self.lib.bert_forward_fake_batch if fake_batch else self.lib.bert_forward_batch
        fn(self.ctx, self.n_threads, n, ptrs, _ip(lens), _row_ptrs(out))
        return out

    # -- measurement extensions (bert_b200.h) -----------------------------------------
    def n_devices(self):
        return self.lib.bertx_n_devices(self.ctx)

    def hparams(self):
        k = ["n_vocab", "n_max_tokens", "n_embd", "n_intermediate", "n_head", "n_layer", "ftype"]
        return {name: self.lib.bertx_hparam(self.ctx, i) for i, name in enumerate(k)}

    def set_pooling(self, mode):
        """'mean' (reference behaviour) or 'cls' (opt-in extension); returns the previous mode."""
        prev = self.lib.bertx_set_pooling(self.ctx, {"mean": 0, "cls": 1}[mode])
        if prev < 0:
            raise ValueError(mode)
        return ["mean", "cls"][prev]

    def kernel_launches(self):
        return int(self.lib.bertx_kernel_launches(self.ctx))

    def plan(self, lens, n_devices=1):
        """The micro-batches one encode() call would cut for sentences of these token counts on n_devices GPUs
        (bertx_plan; works on a host-only context).  Returns (micro-batch index per sentence, tokens per micro-batch)."""
        lens = np.ascontiguousarray(lens, dtype=np.int32)
        mb_of = np.zeros(len(lens), dtype=np.int32)
        mb_tok = np.zeros(len(lens) + 1, dtype=np.int32)
        n = self.lib.bertx_plan(self.ctx, n_devices, len(lens), _ip(lens), _ip(mb_of), _ip(mb_tok), len(mb_tok))
        if n < 0:
            raise RuntimeError(f"bertx_plan failed ({n})")
        return mb_of, mb_tok[:n].copy()

    def tokenize_batch(self, sentences):
        n = len(sentences)
        cap = n * self.n_max_tokens
        ids = np.empty(cap, dtype=np.int32)                    # (filled by the call; untouched pages are never mapped)
        lens = np.zeros(n, dtype=np.int32)
        tot = self.lib.bertx_tokenize_batch(self.ctx, self.n_threads, n, _text_ptrs(sentences), _ip(ids), cap, _ip(lens))
        if tot < 0:
            raise RuntimeError("bertx_tokenize_batch failed")
        return ids[:tot].copy(), lens

    def stage(self, ids, lens, dev=0):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        lens = np.ascontiguousarray(lens, dtype=np.int32)
        rc = self.lib.bertx_stage(self.ctx, dev, len(lens), _ip(ids), _ip(lens))
        if rc != 0:
            raise RuntimeError(f"bertx_stage failed ({rc})")
        self._staged = len(lens)

    def run_staged(self, reps=1, dev=0, fetch=False):
        """Returns (milliseconds for all reps, embeddings or None)."""
        ms = ctypes.c_float(0)
        out = np.zeros((self._staged, self.n_embd), dtype=np.float32) if fetch else None
        rc = self.lib.bertx_run_staged(self.ctx, dev, reps, ctypes.byref(ms), _fp(out) if fetch else None)
        if rc != 0:
            raise RuntimeError(f"bertx_run_staged failed ({rc})")
        return ms.value, out

    def profile_staged(self, dev=0):
        ms = np.zeros(8, dtype=np.float32)
        cnt = np.zeros(8, dtype=np.int64)
        rc = self.lib.bertx_profile_staged(self.ctx, dev, _fp(ms), cnt.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)))
        if rc != 0:
            raise RuntimeError(f"bertx_profile_staged failed ({rc})")
        names = ["embed_ln", "gemm", "attention", "add_ln", "pool_norm", "other"]
        return {n: (float(ms[i]), int(cnt[i])) for i, n in enumerate(names)}


def pack_spans(lens):
    """Attention work units of one micro-batch (bertx_pack_spans, host only): token offsets of the spans that consecutive
    sentences of these token counts are grouped into (at most 128 tokens per span; a longer sentence is its own span)."""
    lens = np.ascontiguousarray(lens, dtype=np.int32)
    span = np.zeros(len(lens) + 1, dtype=np.int32)
    n = lib().bertx_pack_spans(len(lens), _ip(lens), _ip(span))
    if n < 0:
        raise RuntimeError(f"bertx_pack_spans failed ({n})")
    return span[:n + 1].copy()


def convert_safetensors(model_dir, fname_out, ftype):
    """HF checkpoint directory -> the reference's model file (what convert-to-ggml.py, then quantize, would write)."""
    rc = lib().bertx_convert_safetensors(*[os.fsencode(os.path.join(model_dir, f))
                                           for f in ("model.safetensors", "config.json", "vocab.txt")], ftype, os.fsencode(fname_out))
    if rc != 0:
        raise RuntimeError(f"bertx_convert_safetensors failed ({rc}, see stderr)")


def quantize_file(fname_inp, fname_out, itype):
    """models/quantize.cpp's `quantize in out type` (2 = q4_0, 3 = q4_1) on the GPU; raises on failure."""
    rc = lib().bertx_quantize_file(os.fsencode(fname_inp), os.fsencode(fname_out), itype)
    if rc != 0:
        raise RuntimeError(f"bertx_quantize_file failed ({rc}, see stderr)")


def quantize_rows_noalloc(values, itype):
    """q4_0 (itype 2) / q4_1 (itype 3) blocks of a float32 or float16 array through bertx_quantize_rows_noalloc (fixed
    device scratch, no allocation; at most 1 Mi values per call).  Returns the block bytes as a uint8 array."""
    a = np.ascontiguousarray(values)
    if a.dtype not in (np.float32, np.float16):
        raise TypeError("float32 or float16 values")
    out = np.empty(max(a.size // 32, 1) * (18 if itype == 2 else 20), dtype=np.uint8)
    rc = lib().bertx_quantize_rows_noalloc(a.ctypes.data, 0 if a.dtype == np.float32 else 1, a.size, itype, out.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"bertx_quantize_rows_noalloc failed ({rc}, see stderr)")
    return out[:a.size // 32 * (18 if itype == 2 else 20)]


def selftest_gemm(M, N, K, wtype, epilogue, seed=0):
    err, ref = ctypes.c_float(0), ctypes.c_float(0)
    rc = lib().bertx_selftest_gemm(M, N, K, wtype, epilogue, seed, ctypes.byref(err), ctypes.byref(ref))
    return rc, err.value, ref.value


def selftest_attention(n_sent, max_len, n_head, d_head, seed=0):
    err, ref = ctypes.c_float(0), ctypes.c_float(0)
    rc = lib().bertx_selftest_attention(n_sent, max_len, n_head, d_head, seed, ctypes.byref(err), ctypes.byref(ref))
    return rc, err.value, ref.value


def debug_attention(lens, n_head, d_head, seed=0, flags=0):
    """bertx_debug_attention: (qkv [T, 3, H] float16, out [T, H] float16) of the production attention kernel on random inputs."""
    lens = np.ascontiguousarray(lens, dtype=np.int32)
    T, H = int(lens.sum()), n_head * d_head
    qkv = np.zeros((T, 3, H), dtype=np.float16)
    out = np.zeros((T, H), dtype=np.float16)
    rc = lib().bertx_debug_attention(len(lens), _ip(lens), n_head, d_head, seed, flags, qkv.ctypes.data, out.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"bertx_debug_attention failed ({rc})")
    return qkv, out


def bench_attention(n_sent, length, n_head, d_head, reps=10):
    """(ms per launch, TFLOP/s) of the attention kernel on n_sent sentences of `length` tokens."""
    ms = ctypes.c_float(0)
    rc = lib().bertx_bench_attention(n_sent, length, n_head, d_head, reps, ctypes.byref(ms))
    if rc != 0:
        raise RuntimeError(f"bertx_bench_attention failed ({rc})")
    per = ms.value / reps
    return per, 4.0 * n_sent * length * length * n_head * d_head / (per * 1e-3) / 1e12


def bench_gemm(M, N, K, wtype, epilogue, reps=20):
    """(ms per launch, TFLOP/s) of the tcgen05 GEMM kernel on random data."""
    ms = ctypes.c_float(0)
    rc = lib().bertx_bench_gemm(M, N, K, wtype, epilogue, reps, ctypes.byref(ms))
    if rc != 0:
        raise RuntimeError(f"bertx_bench_gemm failed ({rc})")
    per = ms.value / reps
    return per, 2.0 * M * N * K / (per * 1e-3) / 1e12
