"""CPU suite, part 2: the C-ABI library loads and exports every declared symbol, and the host
logic (tokenizer, model-file parser, error conventions, sharding) behaves like the reference.
No compute call is made without a GPU."""
import ctypes
import glob
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from oracle import fileformat, tokenizer
from test_oracle import GOLDEN, load_golden
from tools import synth


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        names |= set(re.findall(r"BERT_API[^;(]*?\b(bertx?_[a-z_0-9]+)\s*\(", src, flags=re.S))
    return names


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()
    names = declared_symbols()
    assert {"bert_load_from_file", "bert_encode_batch", "bert_n_embd", "bert_tokenize", "bert_forward_batch",
            "bert_forward_fake_batch", "bert_forward", "bert_encode", "bert_free", "bert_n_max_tokens",
            "bert_vocab_id_to_token", "bert_params_parse"} <= names
    assert len(names) >= 21
    for n in names:
        assert hasattr(L, n), f"libbert.so does not export {n}"


def test_library_is_self_contained(pkg):
    out = subprocess.run(["ldd", pkg.LIB_PATH], capture_output=True, text=True).stdout
    assert "libcudart" not in out and "libggml" not in out and "torch" not in out


@pytest.mark.parametrize("path", GOLDEN[:2] + GOLDEN[4:6], ids=lambda p: os.path.basename(p)[:-4])
def test_host_tokenizer_matches_reference_golden_ids(pkg, path):
    shape, ft, model, texts, toks, g = load_golden(path)
    m = pkg.BertModel(model, host_only=True)
    for t, want in zip(texts, toks):
        assert m.tokenize(t) == want
    ids, lens = m.tokenize_batch(texts)
    assert lens.tolist() == [len(t) for t in toks]
    assert ids.tolist() == [i for t in toks for i in t]
    m.close()


def test_host_tokenizer_fuzz_against_oracle(pkg):
    import random
    model = synth.model_path("tiny-d32", "f16")
    mf = fileformat.read_model(model)
    voc = tokenizer.Vocab(mf.vocab)
    m = pkg.BertModel(model, host_only=True)
    rnd = random.Random(7)
    alphabet = [b"a", b"b", b"Z", b" ", b"!", b"#", "你".encode(), "好".encode(), b"\xc3\x89", b"\xc3", "こ".encode(),
                b"\x80", b"\xff", b"\t", b"hello", b"dup", b"1", "！".encode(), b"\xf0\x9f\x98\x80", b"\n", b"ing"]
    for _ in range(2000):
        t = b"".join(rnd.choice(alphabet) for _ in range(rnd.randint(0, 48)))
        assert m.tokenize(t) == tokenizer.tokenize(voc, t, 512), t
    # documented deviation: inputs needing more than n_max ids are truncated, [SEP] kept
    long = b" ".join([b"hello"] * 700)
    got = m.tokenize(long)
    assert len(got) == 512 and got[0] == 101 and got[-1] == 102
    assert m.tokenize(long) == tokenizer.tokenize(voc, long, 512, safe=True)
    assert m.tokenize(b"hello world", 4) == tokenizer.tokenize(voc, b"hello world", 4, safe=True)
    # id -> token, including the reference's literal for unknown ids (bert.cpp:121-134)
    assert m.id_to_token(101) == b"[CLS]" and m.id_to_token(-1) == b"[UNK TOKEN from bert_vocab]"
    assert m.id_to_token(mf.hparams.n_vocab) == b"[UNK TOKEN from bert_vocab]"
    m.close()


def test_loader_error_conventions(pkg, tmp_path, capfd):
    L = pkg.lib()
    assert not L.bertx_load_host_only(str(tmp_path / "missing.bin").encode())
    bad = tmp_path / "bad.bin"
    bad.write_bytes(b"\0" * 64)
    assert not L.bertx_load_host_only(str(bad).encode())
    good = open(synth.model_path("tiny-d32", "q4_0"), "rb").read()
    trunc = tmp_path / "trunc.bin"
    trunc.write_bytes(good[: len(good) // 2])
    assert not L.bertx_load_host_only(str(trunc).encode())
    err = capfd.readouterr().err
    assert "failed to open" in err and "bad magic" in err


@pytest.mark.skipif(has_gpu(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback(pkg, capfd):
    """Without a GPU the product must fail loudly, never compute on the host."""
    model = synth.model_path("tiny-d32", "f16")
    assert not pkg.lib().bert_load_from_file(model.encode())
    assert "no CUDA device" in capfd.readouterr().err
    m = pkg.BertModel(model, host_only=True)
    out = m.encode([b"hello world"])
    assert (out == 0).all()                       # outputs untouched
    assert "no GPU replica" in capfd.readouterr().err
    rc, _, _ = pkg.selftest_gemm(128, 64, 64, 1, 0)
    assert rc < 0
    m.close()


def test_product_never_touches_the_oracle():
    pkgdir = os.path.join(ROOT, "embeddings.cpp_b200")
    for dirpath, _, files in os.walk(pkgdir):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the oracle", ""), f
                assert "/root/reference" not in src or f == "__init__.py", f


WORKER = r'''
import os, sys, json
sys.path.insert(0, %(root)r)
import torch, torch.distributed as dist
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 37
lo, hi = bench.shard_range(n, rank, world)
# every rank "encodes" its shard (a deterministic stand-in) and the shards tile the job exactly
mine = torch.zeros(n, dtype=torch.float64)
mine[lo:hi] = torch.arange(lo, hi, dtype=torch.float64) * 3 + 1
dist.all_reduce(mine)
assert torch.equal(mine, torch.arange(n, dtype=torch.float64) * 3 + 1)
t = torch.tensor([float(rank + 1)])
dist.all_reduce(t, op=dist.ReduceOp.MAX)          # the bench's max-over-ranks timing rule
assert t.item() == world
if rank == 0: print("OK", lo, hi)
dist.destroy_process_group()
'''


def test_sentence_sharding_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % {"root": ROOT})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29571", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK 0 19" in r.stdout, r.stdout + r.stderr


def test_shard_range_properties():
    import bench
    for n in (0, 1, 7, 512, 1000003):
        for w in (1, 2, 4, 8):
            parts = [bench.shard_range(n, r, w) for r in range(w)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.skipif(__import__("conftest").has_gpu(), reason="checks the no-GPU failure path")
def test_quantizer_fails_loudly_without_a_gpu(pkg, tmp_path, capfd):
    from tools import synth
    with pytest.raises(RuntimeError):
        pkg.quantize_file(synth.model_path("tiny-d64", "f16"), str(tmp_path / "o.bin"), 2)
    assert "no CUDA device" in capfd.readouterr().err
    with pytest.raises(RuntimeError):
        pkg.quantize_file("/nonexistent", str(tmp_path / "o.bin"), 7)      # bad type is rejected first
