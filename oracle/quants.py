"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the ggml block formats on the path.

Follows /root/reference/ggml/src/ggml.c:
  block_q4_0 / block_q4_1 layouts ............ ggml.c:893-907
  quantize_row_q4_0_reference ................ ggml.c:942-977
  quantize_row_q4_1_reference ................ ggml.c:983-1018
  dequantize_row_q4_0 / q4_1 ................. ggml.c:1524-1563
  quantize_row_q8_0 (AVX2 branch) ............ ggml.c:1146-1300 (d = max/127 -> fp16,
                                               id = 127/max, round-half-even)
  quantize_row_q8_1 (AVX2 branch) ............ ggml.c:1425-1480 (d f32, s = d*sum(q))
  ggml_quantize_q4_0 / q4_1 (row loop) ....... ggml.c:19184-19228
Arithmetic is float32 as in the C code; float64 appears only to emulate a single-rounding FMA.
"""
import numpy as np

QK = 32
F32 = np.float32


def _as_blocks(x):
    x = np.ascontiguousarray(x, dtype=F32)
    assert x.shape[-1] % QK == 0, "row length must be a multiple of 32"
    return x.reshape(-1, QK)


def quantize_q4_0(x):
    """rows of f32 -> raw block_q4_0 bytes (uint8 array, 18 B per 32 weights)."""
    b = _as_blocks(x)
    idx = np.argmax(np.abs(b), axis=1)               # first max wins (strict '<' in the C loop)
    mx = b[np.arange(b.shape[0]), idx]
    d = (mx / F32(-8.0)).astype(F32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, F32(1.0) / d, F32(0.0)).astype(F32)
    # gcc -O3 -mfma contracts `x*id + 8.5f` into one FMA (single rounding); emulate via f64
    q = (b.astype(np.float64) * idv[:, None].astype(np.float64) + 8.5).astype(F32)
    q = np.minimum(15, q.astype(np.int8).astype(np.int32)).astype(np.uint8)   # (int8_t) cast truncates
    out = np.empty((b.shape[0], 18), dtype=np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(-1)


def quantize_q4_1(x):
    """rows of f32 -> raw block_q4_1 bytes (20 B per 32 weights)."""
    b = _as_blocks(x)
    mn = b.min(axis=1)
    mx = b.max(axis=1)
    d = ((mx - mn) / F32(15.0)).astype(F32)
    with np.errstate(divide="ignore"):
        idv = np.where(d != 0, F32(1.0) / d, F32(0.0)).astype(F32)
    q = ((b - mn[:, None]).astype(F32).astype(np.float64) * idv[:, None].astype(np.float64) + 0.5).astype(F32)
    q = np.minimum(15, q.astype(np.int8).astype(np.int32)).astype(np.uint8)
    out = np.empty((b.shape[0], 20), dtype=np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = mn.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 4:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(-1)


def dequantize_q4_0(raw, n_rows, n_cols):
    """raw block bytes -> f32 [n_rows, n_cols]; w[j] = (lo-8)*d, w[j+16] = (hi-8)*d."""
    blk = np.frombuffer(raw, dtype=np.uint8).reshape(-1, 18)
    d = blk[:, 0:2].copy().view(np.float16).astype(F32)           # [nb,1]
    qs = blk[:, 2:]
    lo = (qs & 0x0F).astype(np.int32) - 8
    hi = (qs >> 4).astype(np.int32) - 8
    w = np.concatenate([lo, hi], axis=1).astype(F32) * d
    return w.reshape(n_rows, n_cols)


def dequantize_q4_1(raw, n_rows, n_cols):
    """w[j] = lo*d + m, w[j+16] = hi*d + m."""
    blk = np.frombuffer(raw, dtype=np.uint8).reshape(-1, 20)
    d = blk[:, 0:2].copy().view(np.float16).astype(F32)
    m = blk[:, 2:4].copy().view(np.float16).astype(F32)
    qs = blk[:, 4:]
    lo = (qs & 0x0F).astype(F32)
    hi = (qs >> 4).astype(F32)
    w = np.concatenate([lo, hi], axis=1) * d + m
    return w.astype(F32).reshape(n_rows, n_cols)


def fake_quant_q8_0(x):
    """Activation rows -> q8_0 and straight back to f32 (what vec_dot_q4_0_q8_0 sees)."""
    b = _as_blocks(x)
    amax = np.abs(b).max(axis=1)
    d = (amax / F32(127.0)).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        idv = np.where(amax != 0, F32(127.0) / amax, F32(0.0)).astype(F32)
    q = np.rint((b * idv[:, None]).astype(F32))                   # _MM_ROUND_NEAREST = half-even
    d16 = d.astype(np.float16).astype(F32)                        # block_q8_0.d is fp16
    return (q * d16[:, None]).astype(F32).reshape(np.shape(x))


def fake_quant_q8_1(x):
    """Activation rows -> q8_1 and back (d stays f32: block_q8_1, ggml.c:934-939)."""
    b = _as_blocks(x)
    amax = np.abs(b).max(axis=1)
    d = (amax / F32(127.0)).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        idv = np.where(amax != 0, F32(127.0) / amax, F32(0.0)).astype(F32)
    q = np.rint((b * idv[:, None]).astype(F32))
    return (q * d[:, None]).astype(F32).reshape(np.shape(x))


def raw_bytes_per_row(ftype, n_cols):
    return {0: 4 * n_cols, 1: 2 * n_cols, 2: n_cols // 32 * 18, 3: n_cols // 32 * 20}[ftype]


def decode_tensor(raw, ftype, n_rows, n_cols):
    """Raw file bytes of a 2-D tensor -> f32 [n_rows, n_cols] exactly as ggml dequantises it."""
    if ftype == 0:
        return np.frombuffer(raw, dtype=F32).reshape(n_rows, n_cols).copy()
    if ftype == 1:
        return np.frombuffer(raw, dtype=np.float16).reshape(n_rows, n_cols).astype(F32)
    if ftype == 2:
        return dequantize_q4_0(raw, n_rows, n_cols)
    if ftype == 3:
        return dequantize_q4_1(raw, n_rows, n_cols)
    raise ValueError(f"bad ftype {ftype}")
