"""TEST INFRASTRUCTURE ONLY -- reader/writer for the reference's legacy-ggml model file.

Format (all little-endian), as written by /root/reference/models/convert-to-ggml.py:68-110
and parsed by /root/reference/bert.cpp:434-468 (header), 471-495 (vocab), 656-752 (tensors):

  u32 magic 0x67676d6c
  i32 n_vocab, n_max_tokens, n_embd, n_intermediate, n_head, n_layer, f16(ftype 0..3)
  n_vocab x { u32 len, bytes }
  until EOF: { i32 n_dims, i32 name_len, i32 ftype, i32 ne[n_dims] (fastest dim first),
               name bytes, raw data }
Tensor order is the HF state_dict order; 1-D tensors are always f32; every 2-D
'*.weight' (including the three embedding tables) is in the file's weight type
(models/quantize.cpp:154-167).
"""
import struct
from dataclasses import dataclass, field

import numpy as np

from . import quants

MAGIC = 0x67676D6C
FTYPE_NAMES = {0: "f32", 1: "f16", 2: "q4_0", 3: "q4_1"}
FTYPE_IDS = {v: k for k, v in FTYPE_NAMES.items()}


@dataclass
class Hparams:
    n_vocab: int
    n_max_tokens: int
    n_embd: int
    n_intermediate: int
    n_head: int
    n_layer: int
    ftype: int


@dataclass
class RawTensor:
    name: str
    ftype: int
    ne: tuple          # (ne0[, ne1]) fastest dim first, as stored
    raw: bytes

    def f32(self):
        """Decoded to float32, shaped [ne1, ne0] (2-D) or [ne0] (1-D)."""
        if len(self.ne) == 1:
            assert self.ftype == 0
            return np.frombuffer(self.raw, dtype=np.float32).copy()
        return quants.decode_tensor(self.raw, self.ftype, self.ne[1], self.ne[0])


@dataclass
class ModelFile:
    hparams: Hparams
    vocab: list                      # list[bytes], index = id
    tensors: dict = field(default_factory=dict)   # name -> RawTensor (insertion = file order)


def tensor_names(n_layer):
    """HF state_dict order minus position_ids / pooler (convert-to-ggml.py:84-87)."""
    names = [
        ("embeddings.word_embeddings.weight", 2),
        ("embeddings.position_embeddings.weight", 2),
        ("embeddings.token_type_embeddings.weight", 2),
        ("embeddings.LayerNorm.weight", 1),
        ("embeddings.LayerNorm.bias", 1),
    ]
    for i in range(n_layer):
        p = f"encoder.layer.{i}."
        names += [
            (p + "attention.self.query.weight", 2), (p + "attention.self.query.bias", 1),
            (p + "attention.self.key.weight", 2), (p + "attention.self.key.bias", 1),
            (p + "attention.self.value.weight", 2), (p + "attention.self.value.bias", 1),
            (p + "attention.output.dense.weight", 2), (p + "attention.output.dense.bias", 1),
            (p + "attention.output.LayerNorm.weight", 1), (p + "attention.output.LayerNorm.bias", 1),
            (p + "intermediate.dense.weight", 2), (p + "intermediate.dense.bias", 1),
            (p + "output.dense.weight", 2), (p + "output.dense.bias", 1),
            (p + "output.LayerNorm.weight", 1), (p + "output.LayerNorm.bias", 1),
        ]
    return names


def read_model(path):
    with open(path, "rb") as f:
        buf = f.read()
    off = 0
    (magic,) = struct.unpack_from("<I", buf, off); off += 4
    if magic != MAGIC:
        raise ValueError("bad magic")
    hp = Hparams(*struct.unpack_from("<7i", buf, off)); off += 28
    vocab = []
    for _ in range(hp.n_vocab):
        (n,) = struct.unpack_from("<I", buf, off); off += 4
        vocab.append(bytes(buf[off:off + n])); off += n
    mf = ModelFile(hp, vocab)
    while off < len(buf):
        n_dims, name_len, ftype = struct.unpack_from("<3i", buf, off); off += 12
        ne = struct.unpack_from(f"<{n_dims}i", buf, off); off += 4 * n_dims
        name = buf[off:off + name_len].decode(); off += name_len
        if n_dims == 1:
            nbytes = 4 * ne[0] if ftype == 0 else 2 * ne[0]
        else:
            nbytes = quants.raw_bytes_per_row(ftype, ne[0]) * ne[1]
        mf.tensors[name] = RawTensor(name, ftype, tuple(ne), bytes(buf[off:off + nbytes])); off += nbytes
    return mf


def write_model(path, hp, vocab, tensors_f32, ftype=None):
    """Write a model file.  tensors_f32: name -> float32 ndarray ([out,in] for 2-D).

    ftype 0/1 follow convert-to-ggml.py:93-98; ftype 2/3 additionally apply the
    reference quantiser's rule (models/quantize.cpp:154-217: every 2-D '*weight'
    tensor, rows of ne[0]) using oracle.quants (bit-exact restatement, see tests).
    q4 files made from an f16 file differ from ones made from f32 only by the f16
    rounding of the source; here 2/3 quantise the f16-ROUNDED weights, which is what
    `quantize ggml-model-f16.bin` (models/run_conversions.sh) would produce.
    """
    ftype = hp.ftype if ftype is None else ftype
    with open(path, "wb") as f:
        f.write(struct.pack("<I", MAGIC))
        f.write(struct.pack("<7i", hp.n_vocab, hp.n_max_tokens, hp.n_embd, hp.n_intermediate,
                            hp.n_head, hp.n_layer, ftype))
        assert len(vocab) == hp.n_vocab
        for tok in vocab:
            f.write(struct.pack("<I", len(tok)))
            f.write(tok)
        for name, nd in tensor_names(hp.n_layer):
            a = np.ascontiguousarray(tensors_f32[name], dtype=np.float32)
            assert a.ndim == nd, name
            is_w = nd == 2 and name.endswith(".weight")
            t = ftype if is_w else 0
            nb = name.encode()
            f.write(struct.pack("<3i", nd, len(nb), t))
            for i in range(nd):
                f.write(struct.pack("<i", a.shape[nd - 1 - i]))
            f.write(nb)
            if t == 0:
                a.tofile(f)
            elif t == 1:
                a.astype(np.float16).tofile(f)
            else:
                src = a.astype(np.float16).astype(np.float32)
                q = quants.quantize_q4_0(src) if t == 2 else quants.quantize_q4_1(src)
                q.tofile(f)
