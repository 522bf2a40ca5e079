"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference forward pass.

Restates /root/reference/bert.cpp:827-1147 (bert_forward_batch) with the ggml CPU
numerics the graph actually reaches (/root/reference/ggml/src/ggml.c):

  get_rows (dequantise gathered rows) ........ ggml.c:11614-11696   -> bert.cpp:963-974
  norm (eps 1e-5, f64 accumulators) .......... ggml.c:10605-10653   -> bert.cpp:977-984
  mul_mat: activations converted to the weight type's vec_dot_type
     f32 -> f32, f16 -> f16, q4_0 -> q8_0, q4_1 -> q8_1 ... ggml.c:1644-1700, 11135-11152
  gelu via f16 lookup table of tanh-GELU ..... ggml.c:3561-3580, 4332
  soft_max: max-subtract, exp via f16 table, f64 sum ..... ggml.c:12008-12077, 4335
  attention mask (pad x pad - 1) * 1e5 ....... bert.cpp:957-961
  masked-mean pooling + L2 normalise ......... bert.cpp:894-922, 1087-1095

Summation ORDER inside dot products is not restated (numpy/BLAS float32), so the
restatement agrees with the compiled reference to float32 rounding, not bit-exactly;
tests pin it at cosine >= 0.999999 / max-abs <= 2e-5 against oracle/_ref outputs.

`exact_activations=True` drops the reference's activation rounding (q8 blocks, f16
rows, f16 LUTs) and computes everything in float32 from the dequantised weights: the
"ideal" result that both the reference and the CUDA path approximate.
"""
import numpy as np

from . import quants
from .fileformat import ModelFile

F32 = np.float32
SQRT_2_OVER_PI = F32(0.79788456080286535587989211986876)
GELU_COEF_A = F32(0.044715)


def _f16_round(x):
    return x.astype(np.float16).astype(F32)


def gelu_lut(x):
    """ggml_vec_gelu_f32 with GGML_GELU_FP16: y = f32(table[f16(x)]), table = f16(gelu_f32(f))."""
    with np.errstate(over="ignore"):
        f = _f16_round(x)
        g = (F32(0.5) * f * (F32(1.0) + np.tanh(SQRT_2_OVER_PI * f * (F32(1.0) + GELU_COEF_A * f * f)))).astype(F32)
        return _f16_round(g)


def gelu_exact(x):
    x = x.astype(F32)
    return (F32(0.5) * x * (F32(1.0) + np.tanh(SQRT_2_OVER_PI * x * (F32(1.0) + GELU_COEF_A * x * x)))).astype(F32)


def layer_norm(x, g, b):
    x64 = x.astype(np.float64)
    mean = (x64.sum(axis=-1, keepdims=True) / x.shape[-1]).astype(F32)
    v = (x - mean).astype(F32)
    var = ((v.astype(np.float64) ** 2).sum(axis=-1, keepdims=True) / x.shape[-1]).astype(F32)
    scale = (F32(1.0) / np.sqrt(var + F32(1e-5))).astype(F32)
    return ((v * scale) * g + b).astype(F32)


class OracleModel:
    def __init__(self, mf: ModelFile):
        self.hp = mf.hparams
        self.ftype = mf.hparams.ftype
        t = mf.tensors
        self.w = {name: rt.f32() for name, rt in t.items()}
        self.wt = {name: rt.ftype for name, rt in t.items()}

    def _linear(self, x, wname, bname, exact):
        w = self.w[wname]
        ft = self.wt[wname]
        if not exact:
            if ft == 1:
                x = _f16_round(x)
            elif ft == 2:
                x = quants.fake_quant_q8_0(x)
            elif ft == 3:
                x = quants.fake_quant_q8_1(x)
        y = x @ w.T
        return (y + self.w[bname]).astype(F32)

    def forward(self, batch_tokens, exact_activations=False, return_hidden=False, pooling="mean"):
        """batch_tokens: list of lists of ids.  Returns [B, n_embd] float32.
        pooling="cls" is NOT in the reference (it only has the masked mean, bert.cpp:1087-1089): it checks the
        opt-in [CLS] extension of the CUDA path -- same hidden states, row 0 instead of the mean."""
        hp = self.hp
        ex = exact_activations
        H, nh = hp.n_embd, hp.n_head
        d = H // nh
        B = len(batch_tokens)
        lens = [len(t) for t in batch_tokens]
        L = max(lens)
        if L > hp.n_max_tokens:
            raise ValueError("Too many tokens")
        tok = np.full((B, L), 101, dtype=np.int64)                 # pad id 101 (bert.cpp:916)
        pad = np.zeros((B, L), dtype=F32)
        wsum = np.zeros((B, L), dtype=F32)
        for b, t in enumerate(batch_tokens):
            tok[b, :len(t)] = t
            pad[b, :len(t)] = 1.0
            wsum[b, :len(t)] = F32(1.0) / F32(len(t))
        mask = ((pad[:, :, None] * pad[:, None, :] - F32(1.0)) * F32(100000.0)).astype(F32)   # [B,L,L]

        W = self.w
        x = (W["embeddings.word_embeddings.weight"][tok]
             + W["embeddings.token_type_embeddings.weight"][0]
             + W["embeddings.position_embeddings.weight"][np.arange(L)][None]).astype(F32)
        x = layer_norm(x, W["embeddings.LayerNorm.weight"], W["embeddings.LayerNorm.bias"])
        scale = F32(1.0 / np.sqrt(float(d)))
        for il in range(hp.n_layer):
            p = f"encoder.layer.{il}."
            x2 = x.reshape(B * L, H)
            q = self._linear(x2, p + "attention.self.query.weight", p + "attention.self.query.bias", ex)
            k = self._linear(x2, p + "attention.self.key.weight", p + "attention.self.key.bias", ex)
            v = self._linear(x2, p + "attention.self.value.weight", p + "attention.self.value.bias", ex)
            q = q.reshape(B, L, nh, d).transpose(0, 2, 1, 3)
            k = k.reshape(B, L, nh, d).transpose(0, 2, 1, 3)
            v = v.reshape(B, L, nh, d).transpose(0, 2, 1, 3)
            s = (np.matmul(q, k.transpose(0, 1, 3, 2)) * scale).astype(F32) + mask[:, None]
            m = s.max(axis=-1, keepdims=True)
            if ex:
                e = np.exp((s - m).astype(F32)).astype(F32)
            else:
                with np.errstate(over="ignore"):
                    e = _f16_round(np.exp(_f16_round((s - m).astype(F32))).astype(F32))
            inv = (1.0 / e.astype(np.float64).sum(axis=-1, keepdims=True)).astype(F32)
            pr = (e * inv).astype(F32)
            ctx = np.matmul(pr, v).transpose(0, 2, 1, 3).reshape(B * L, H).astype(F32)
            a = self._linear(ctx, p + "attention.output.dense.weight", p + "attention.output.dense.bias", ex)
            a = layer_norm((a + x2).astype(F32), W[p + "attention.output.LayerNorm.weight"],
                           W[p + "attention.output.LayerNorm.bias"])
            h = self._linear(a, p + "intermediate.dense.weight", p + "intermediate.dense.bias", ex)
            h = gelu_exact(h) if ex else gelu_lut(h)
            o = self._linear(h, p + "output.dense.weight", p + "output.dense.bias", ex)
            x = layer_norm((a + o).astype(F32), W[p + "output.LayerNorm.weight"],
                           W[p + "output.LayerNorm.bias"]).reshape(B, L, H)
        pooled = x[:, 0, :].astype(F32) if pooling == "cls" else np.einsum("blh,bl->bh", x, wsum).astype(F32)
        norm = np.sqrt((pooled * pooled).sum(axis=-1, keepdims=True)).astype(F32)
        out = (pooled / norm).astype(F32)
        if return_hidden:
            return out, x
        return out
