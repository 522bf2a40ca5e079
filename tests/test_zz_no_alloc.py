"""bertx_quantize_rows_noalloc: the block quantiser with its device staging in a fixed __device__ array (no
cudaMalloc / cudaFree) against the existing entry point, bertx_quantize_file, on the same values.

The file is named so that pytest collects it last (`pytest --collect-only`).
The integer / byte bar of this repo applies: the blocks must be IDENTICAL bytes (tolerance 0), on two calls in a row
(the second call finds the first call's bytes in the shared array).
"""
import os

import numpy as np
import pytest

from conftest import has_gpu
from oracle import fileformat, quants
from tools import synth

gpu = pytest.mark.gpu


def _tensors(path):
    return fileformat.read_model(path).tensors


@gpu
def test_no_alloc_quantiser_matches_the_file_quantiser_twice_in_a_row(pkg, tmp_path):
    shape = "tiny-d64"
    results = []
    # two calls in a row per source type, of different sizes and block formats, biggest first so that the second call
    # runs over the leftovers of the first
    for src_ft, np_dt in (("f32", np.float32), ("f16", np.float16)):
        src_path = synth.model_path(shape, src_ft)
        src = _tensors(src_path)
        want = {}
        for itype, ft in ((2, "q4_0"), (3, "q4_1")):
            out = str(tmp_path / f"{src_ft}-{ft}.bin")
            pkg.quantize_file(src_path, out, itype)                      # existing entry point (cudaMalloc'ed staging)
            want[itype] = _tensors(out)
        calls = (("embeddings.word_embeddings.weight", 3), ("encoder.layer.1.intermediate.dense.weight", 2),
                 ("encoder.layer.0.attention.self.query.weight", 3), ("embeddings.word_embeddings.weight", 2))
        for name, itype in calls:
            t = src[name]
            vals = np.frombuffer(t.raw, dtype=np_dt)
            assert vals.size == t.ne[0] * t.ne[1] and vals.size <= (1 << 20)
            got = pkg.quantize_rows_noalloc(vals, itype)
            ref = np.frombuffer(want[itype][name].raw, dtype=np.uint8)
            assert want[itype][name].ftype == itype
            assert got.shape == ref.shape, (name, itype)
            assert np.array_equal(got, ref), (src_ft, name, itype, int((got != ref).sum()))
            results.append(got.size)
    assert len(results) == 8 and min(results) > 0


@gpu
def test_no_alloc_quantiser_limit_and_bad_arguments(pkg, capfd):
    big = np.zeros((1 << 20) + 32, dtype=np.float32)
    with pytest.raises(RuntimeError, match=r"\(9,"):
        pkg.quantize_rows_noalloc(big, 2)
    assert "exceed the fixed scratch" in capfd.readouterr().err
    with pytest.raises(RuntimeError, match=r"\(6,"):
        pkg.quantize_rows_noalloc(np.zeros(33, dtype=np.float32), 2)     # not whole blocks
    with pytest.raises(RuntimeError, match=r"\(1,"):
        pkg.quantize_rows_noalloc(np.zeros(32, dtype=np.float32), 5)     # no such block format
    # exactly the maximum works: all-zero blocks come out as d = 0 / -8 = -0.0 and q = 8 (ggml.c:942-977), the bytes
    # the numpy restatement of the reference quantiser gives; a second full-size call of random values right after it
    # (every byte of the array reused) matches the restatement as well
    full = pkg.quantize_rows_noalloc(np.zeros(1 << 20, dtype=np.float32), 2)
    assert full.size == (1 << 20) // 32 * 18
    blk = full.reshape(-1, 18)
    assert (blk[:, :2].copy().view(np.float16) == 0).all() and (blk[:, 2:] == 0x88).all()
    assert np.array_equal(full, quants.quantize_q4_0(np.zeros(1 << 20, dtype=np.float32)))
    x = np.random.default_rng(7).standard_normal(1 << 20).astype(np.float32)
    assert np.array_equal(pkg.quantize_rows_noalloc(x, 3), quants.quantize_q4_1(x))
    assert np.array_equal(pkg.quantize_rows_noalloc(x[:4096], 2), quants.quantize_q4_0(x[:4096]))


def test_no_alloc_entry_point_rejects_oversize_without_touching_a_gpu(pkg, capfd):
    """Argument checks come before any CUDA call, so the limit is enforced (and reported) on a host without a GPU too."""
    big = np.zeros((1 << 20) + 32, dtype=np.float32)
    with pytest.raises(RuntimeError, match=r"\(9,"):
        pkg.quantize_rows_noalloc(big, 3)
    assert "split the rows over several calls" in capfd.readouterr().err
    if not has_gpu():
        with pytest.raises(RuntimeError, match=r"\(2,"):
            pkg.quantize_rows_noalloc(np.zeros(64, dtype=np.float32), 2)
        assert "no CPU path" in capfd.readouterr().err
