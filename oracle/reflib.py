"""TEST INFRASTRUCTURE ONLY -- ctypes driver for the compiled reference (oracle/_ref).

Binds libbert_ref.so exactly like /root/reference/examples/sample_dylib.py:15-62 does
for the reference's libbert.so, plus the one helper from oracle/ref_wrapper.cpp that
pre-sizes the ggml arena so long inputs do not crash (see that file's header).
Nothing here reads /root/reference at run time; the .so is prebuilt.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libbert_ref.so")
QUANTIZE_BIN = os.path.join(REF_DIR, "quantize")


def available():
    return os.path.exists(REF_SO)


class RefBert:
    """The unmodified reference engine (ggml CPU path)."""

    def __init__(self, model_path, so_path=REF_SO):
        self.lib = ctypes.cdll.LoadLibrary(so_path)
        L = self.lib
        L.bert_load_from_file.restype = ctypes.c_void_p
        L.bert_load_from_file.argtypes = [ctypes.c_char_p]
        L.bert_n_embd.restype = ctypes.c_int32
        L.bert_n_embd.argtypes = [ctypes.c_void_p]
        L.bert_n_max_tokens.restype = ctypes.c_int32
        L.bert_n_max_tokens.argtypes = [ctypes.c_void_p]
        L.bert_free.argtypes = [ctypes.c_void_p]
        L.bert_encode_batch.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                        ctypes.POINTER(ctypes.c_char_p),
                                        ctypes.POINTER(ctypes.POINTER(ctypes.c_float))]
        L.bert_forward_batch.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                         ctypes.POINTER(ctypes.POINTER(ctypes.c_int32)),
                                         ctypes.POINTER(ctypes.c_int32),
                                         ctypes.POINTER(ctypes.POINTER(ctypes.c_float))]
        L.bert_tokenize.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int32),
                                    ctypes.POINTER(ctypes.c_int32), ctypes.c_int32]
        L.bert_vocab_id_to_token.restype = ctypes.c_char_p
        L.bert_vocab_id_to_token.argtypes = [ctypes.c_void_p, ctypes.c_int32]
        L.oracle_ref_reserve.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32]
        L.oracle_ref_mem_per_token.restype = ctypes.c_int64
        L.oracle_ref_mem_per_token.argtypes = [ctypes.c_void_p]
        L.oracle_ref_n_head.restype = ctypes.c_int32
        L.oracle_ref_n_head.argtypes = [ctypes.c_void_p]
        L.oracle_ref_n_layer.restype = ctypes.c_int32
        L.oracle_ref_n_layer.argtypes = [ctypes.c_void_p]
        self.ctx = L.bert_load_from_file(model_path.encode())
        if not self.ctx:
            raise RuntimeError(f"reference failed to load {model_path}")
        self.n_embd = L.bert_n_embd(self.ctx)
        self.n_max_tokens = L.bert_n_max_tokens(self.ctx)
        self.n_head = L.oracle_ref_n_head(self.ctx)
        self.n_layer = L.oracle_ref_n_layer(self.ctx)
        self.n_threads = os.cpu_count() or 1

    def close(self):
        if self.ctx:
            self.lib.bert_free(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- arena sizing -----------------------------------------------------------------
    def _reserve(self, bs, max_len):
        """Arena big enough for a [bs, max_len] forward: the reference's own per-token
        probe (bert.cpp:771-781) covers the O(L) tensors; add the O(L^2*n_head) ones
        (mask + repeat + KQ + scale + add + softmax, ~7 f32 copies per layer)."""
        mpt = self.lib.oracle_ref_mem_per_token(self.ctx)
        lin = int(1.3 * mpt * max_len * bs)
        quad = int(5 * 4 * max_len * max_len * self.n_head * bs * (self.n_layer + 1))
        self.lib.oracle_ref_reserve(self.ctx, lin + quad + (64 << 20), max(bs, 1))

    # -- API --------------------------------------------------------------------------
    def tokenize(self, text: bytes, n_max=None):
        n_max = self.n_max_tokens if n_max is None else n_max
        buf = (ctypes.c_int32 * (n_max + 4096))()       # slack: the reference may overrun
        n = ctypes.c_int32(0)
        self.lib.bert_tokenize(self.ctx, text, buf, ctypes.byref(n), n_max)
        return list(buf[:n.value])

    def forward_tokens(self, batch_tokens, chunk=8):
        """bert_forward_batch on token id lists, `chunk` sentences per call."""
        n = len(batch_tokens)
        out = np.zeros((n, self.n_embd), dtype=np.float32)
        for s in range(0, n, chunk):
            part = batch_tokens[s:s + chunk]
            bs = len(part)
            L = max(len(t) for t in part)
            self._reserve(bs, L)
            arrs = [(ctypes.c_int32 * len(t))(*t) for t in part]
            ptrs = (ctypes.POINTER(ctypes.c_int32) * bs)(*[ctypes.cast(a, ctypes.POINTER(ctypes.c_int32)) for a in arrs])
            lens = (ctypes.c_int32 * bs)(*[len(t) for t in part])
            outs = (ctypes.POINTER(ctypes.c_float) * bs)(
                *[out[s + i].ctypes.data_as(ctypes.POINTER(ctypes.c_float)) for i in range(bs)])
            self.lib.bert_forward_batch(self.ctx, self.n_threads, bs, ptrs, lens, outs)
        return out

    def encode(self, texts, batch_size=8, max_len_hint=None):
        """bert_encode_batch (tokenise + length-sort + forward), sample_dylib.py:42-62."""
        n = len(texts)
        out = np.zeros((n, self.n_embd), dtype=np.float32)
        self._reserve(min(batch_size, n), max_len_hint or self.n_max_tokens)
        tarr = (ctypes.c_char_p * n)(*[t if isinstance(t, bytes) else t.encode() for t in texts])
        outs = (ctypes.POINTER(ctypes.c_float) * n)(
            *[out[i].ctypes.data_as(ctypes.POINTER(ctypes.c_float)) for i in range(n)])
        self.lib.bert_encode_batch(self.ctx, self.n_threads, batch_size, n, tarr, outs)
        return out
