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


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libbert_ref.so")),
                    reason="oracle/_ref (the compiled reference) is not on this box")
def test_host_tokenizer_fuzz_100k_against_compiled_reference(pkg):
    """10^5 fuzzed UTF-8 strings through bert_tokenize of the COMPILED reference (bert.cpp:297-417 as
    built by oracle/Makefile) and through the product's host tokenizer; ids must be identical.  English and Chinese
    vocabularies; strings stay short enough that the reference does not overrun its buffer (bert.cpp:408-415)."""
    import random
    from oracle import reflib
    total = 0
    for shape, seed in (("all-MiniLM-L6-v2", 11), ("bge-small-zh-v1.5", 12)):
        model = synth.model_path(shape, "q4_0" if "Mini" in shape else "q4_1")
        vocab = synth.make_vocab(synth.SHAPES[shape][4], 0)
        ref = reflib.RefBert(model)
        m = pkg.BertModel(model, host_only=True)
        rnd = random.Random(seed)
        words = [w for w in vocab[104:] if 0 < len(w) < 12]
        pieces = [b" ", b" ", b" ", b"!", b"#", b"##", b".", b",", b"?", b"\t", b"\n", b"A", b"Zq", b"\xc3\x89", b"\xc3\xa9", b"\xc3",
                  b"\x80", b"\xff", b"\xe4\xbd", b"\xf0\x9f\x98\x80", "你".encode(), "好，".encode(), "こ".encode(), "！".encode(),
                  "（".encode(), b"123", b"dup", b"##dup", b"[CLS]", b"[unused5]"]
        texts = []
        for _ in range(50000):
            k = rnd.randint(0, 14)
            texts.append(b"".join(rnd.choice(words) if rnd.random() < 0.5 else rnd.choice(pieces) for _ in range(k)).replace(b"\0", b" "))
        ids, lens = m.tokenize_batch(texts)
        off = 0
        for t, n in zip(texts, lens.tolist()):
            want = ref.tokenize(t)
            assert ids[off:off + n].tolist() == want, (shape, t)
            off += n
        total += len(texts)
        m.close()
        ref.close()
    assert total == 100000


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
import numpy as np
import torch, torch.distributed as dist
import bench
import __graft_entry__ as ge
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# the N > 1 path of bench.py on CPU: the global batch is generated once from one seed, every rank takes its
# shard_range() slice, tokenises it with the product's host tokenizer and cuts it into micro-batches (bertx_plan on a
# host-only context); the shards must tile the batch exactly and every sentence must land in exactly one micro-batch
pkg = ge.load_package()
shape, kind, total = "tiny-d64", "client", 1201
texts = bench.make_texts(shape, kind, total, 0)
lo, hi = bench.shard_range(total, rank, world)
m = pkg.BertModel(bench.model_file(shape, "q4_0"), host_only=True)
ids, lens = m.tokenize_batch(texts[lo:hi])
mb_of, mb_tok = m.plan(lens, n_devices=1)
assert len(mb_of) == hi - lo and mb_of.min() == 0 and mb_of.max() == len(mb_tok) - 1
assert np.array_equal(np.bincount(mb_of, weights=lens, minlength=len(mb_tok)).astype(np.int64), mb_tok.astype(np.int64))
cover = torch.zeros(total, dtype=torch.float64)
cover[lo:hi] = 1.0
tok = torch.tensor([float(lens.sum()), float(hi - lo), float(len(mb_tok))], dtype=torch.float64)
dist.all_reduce(cover)
dist.all_reduce(tok)
assert torch.equal(cover, torch.ones(total, dtype=torch.float64))          # the shards tile the job exactly
_, all_lens = m.tokenize_batch(texts)
assert tok[0].item() == float(all_lens.sum()) and tok[1].item() == total
t = torch.tensor([float(rank + 1)])
dist.all_reduce(t, op=dist.ReduceOp.MAX)          # the bench's max-over-ranks timing rule
assert t.item() == world
if rank == 0: print("OK", lo, hi, int(tok[2].item()))
m.close()
dist.destroy_process_group()
'''


def test_sentence_sharding_world_size_2_gloo(tmp_path):
    synth.model_path("tiny-d64", "q4_0")                  # generate once, before two ranks race on the cache
    script = tmp_path / "w.py"
    script.write_text(WORKER % {"root": ROOT})
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29571", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK 0 601" in r.stdout, r.stdout + r.stderr


def test_micro_batch_planner_properties(pkg, monkeypatch):
    """The scheduler behind bert_encode_batch (csrc/pipeline.cu; reference bert.cpp:1394-1443 re-designed) on token counts
    alone, through a host-only context: deterministic chunks on 16-sentence boundaries with a ramp (small first chunks
    start the GPUs early), sentences sorted by length inside a chunk, micro-batches packed to the token budget and
    split over the GPUs."""
    monkeypatch.setenv("BERT_MICROBATCH_TOKENS", "8192")
    m = pkg.BertModel(synth.model_path("tiny-d64", "q4_0"), host_only=True)
    rng = np.random.default_rng(3)
    for n, nd in ((5, 1), (16, 2), (17, 1), (1000, 1), (5000, 1), (5000, 2), (5000, 8), (40000, 8)):
        lens = rng.integers(2, 513, n).astype(np.int32)
        mb_of, mb_tok = m.plan(lens, n_devices=nd)
        mb_of2, mb_tok2 = m.plan(lens, n_devices=nd)
        assert np.array_equal(mb_of, mb_of2) and np.array_equal(mb_tok, mb_tok2)        # a function of the counts only
        assert np.array_equal(np.bincount(mb_of, weights=lens, minlength=len(mb_tok)).astype(np.int64), mb_tok)
        assert mb_tok.max() <= 8192 and mb_tok.min() >= 2                               # token budget respected
        first = np.array([np.nonzero(mb_of == k)[0].min() for k in range(len(mb_tok))])
        last = np.array([np.nonzero(mb_of == k)[0].max() for k in range(len(mb_tok))])
        # micro-batches of a chunk interleave inside it, chunks do not: chunk starts are multiples of 16 sentences
        pmax = np.maximum.accumulate(last)
        smin = np.minimum.accumulate(first[::-1])[::-1]
        chunk_start = [0] + [int(smin[k]) for k in range(1, len(mb_tok)) if pmax[k - 1] < smin[k]]
        assert all(c % 16 == 0 for c in chunk_start)
        if n >= 5000:
            # inside a chunk a micro-batch never holds a shorter sentence than a later micro-batch's longest
            k0 = [k for k in range(len(mb_tok)) if first[k] >= chunk_start[1] and (len(chunk_start) < 3 or last[k] < chunk_start[2])]
            for a, b in zip(k0[:-1], k0[1:]):
                assert lens[mb_of == a].min() >= lens[mb_of == b].max()
        if nd > 1 and n >= 5000:
            # a chunk is split over the GPUs: at least nd micro-batches share every full chunk
            sizes = np.diff(chunk_start + [n])
            assert len(mb_tok) >= nd * (len(sizes) - 1)
    m.close()


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


def test_text_pointer_table_round_trips(pkg):
    """const char** built for bert_encode_batch: the numpy path (many short texts: one joined buffer + offsets) and the
    item-by-item path (long or few texts) hand the same bytes to C."""
    cases = [
        [b"hello world", "café 一".encode(), b"", b"x" * 50] * 100,              # many short byte strings -> joined buffer
        ["str one", b"bytes two", "你好"] * 120,                                  # str and bytes mixed
        [b"y" * 1000] * 300,                                                              # long texts -> item by item
        [b"few", b"texts"],                                                               # few texts -> item by item
        [b"embedded\0nul", b"after"] * 200,                                               # C sees the text up to the first NUL
    ]
    for texts in cases:
        p = pkg._text_ptrs(texts)
        for i, t in enumerate(texts):
            want = t if isinstance(t, bytes) else t.encode("utf-8")
            assert p[i] == want.split(b"\0")[0]


def test_span_packing_matches_its_definition(pkg):
    """bertx_pack_spans (the host half of the packed attention spans): greedy, in order, spans of several sentences hold at most
    128 tokens, a longer sentence stays alone, every sentence is in exactly one span -- against a direct Python restatement."""
    def restate(lens):
        cu = np.concatenate([[0], np.cumsum(lens)])
        out, s = [0], 0
        while s < len(lens):
            e = s + 1
            while e < len(lens) and cu[e + 1] - cu[s] <= 128:
                e += 1
            out.append(int(cu[e])); s = e
        return np.array(out, dtype=np.int32)
    rng = np.random.default_rng(5)
    cases = [[], [1], [128], [129], [64, 64], [64, 65], [127, 1, 1], [512, 3, 3, 500, 2], [1] * 300,
             list(rng.integers(1, 40, 500)), list(rng.integers(1, 513, 200)), list(rng.integers(100, 140, 50))]
    for lens in cases:
        got = pkg.pack_spans(lens)
        assert np.array_equal(got, restate(lens)), lens[:10]
        cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        assert got[0] == 0 and got[-1] == cu[-1] and np.all(np.diff(got) > 0) or len(lens) == 0
        assert np.isin(got, cu).all()                                   # spans end on sentence boundaries
        for a, b in zip(got[:-1], got[1:]):                             # > 128 tokens only if it is one sentence
            if b - a > 128:
                assert np.searchsorted(cu, b) - np.searchsorted(cu, a) == 1


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libbert_ref.so")),
                    reason="oracle/_ref (the compiled reference) is not on this box")
def test_params_parse_matches_compiled_reference(pkg, tmp_path):
    """bert_params_parse (bert.cpp:157-193) through both libraries: same fields for every well-formed command line, exit code 0
    and the 'unknown argument' line for an unknown option, exit 0 for --help.  (A missing or non-numeric value is undefined
    behaviour in the reference -- std::stoi on argv[argc] -- and is not compared.)"""
    import subprocess
    pkg.lib()
    probe = str(tmp_path / "params_probe")
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "helpers", "params_probe.cpp"),
                        "-o", probe, "-ldl"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ours = os.path.join(ROOT, "embeddings.cpp_b200", "lib", "libbert.so")
    ref = os.path.join(ROOT, "oracle", "_ref", "libbert_ref.so")
    cases = [[], ["-t", "12"], ["--threads", "3", "--port", "9000"], ["-m", "x/y.bin", "-p", "hello world"],
             ["--model", "a.bin", "--prompt", "", "-t", "1", "--port", "1"], ["-p", "-t"], ["-t", "7", "-t", "9"]]
    for args in cases:
        a = subprocess.run([probe, ours] + args, capture_output=True, text=True)
        b = subprocess.run([probe, ref] + args, capture_output=True, text=True)
        assert a.returncode == 0 and b.returncode == 0, (args, a.stderr, b.stderr)
        assert a.stdout == b.stdout and a.stdout.startswith("ok=1"), (args, a.stdout, b.stdout)
    for args, needle in ((["--bogus"], "error: unknown argument: --bogus"), (["-h"], "usage:"), (["-t", "2", "--help"], "usage:")):
        a = subprocess.run([probe, ours] + args, capture_output=True, text=True)
        b = subprocess.run([probe, ref] + args, capture_output=True, text=True)
        assert a.returncode == b.returncode == 0, args                       # both exit(0) from inside the parser
        assert a.stdout == b.stdout == ""                                    # ... before the probe prints anything
        assert needle in a.stderr and needle in b.stderr, (args, a.stderr)


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libbert_ref.so")),
                    reason="oracle/_ref (the compiled reference) is not on this box")
def test_vocab_getters_match_compiled_reference(pkg):
    """bert_vocab_id_to_token for EVERY id (and a few outside the table), bert_n_embd and bert_n_max_tokens: this library
    (host-only context) and the compiled reference on the same files (bert.cpp:111-134; duplicates and "##" pieces follow the
    reference's two-map rules)."""
    from oracle import reflib
    for shape, ft in (("tiny-d32", "f16"), ("all-MiniLM-L6-v2", "q4_0"), ("bge-small-zh-v1.5", "q4_1")):
        model = synth.model_path(shape, ft)
        ref = reflib.RefBert(model)
        m = pkg.BertModel(model, host_only=True)
        assert m.n_embd == ref.n_embd and m.n_max_tokens == ref.n_max_tokens
        n_vocab = m.hparams()["n_vocab"]
        for i in list(range(-3, n_vocab + 3)) + [2 ** 31 - 1, -2 ** 31]:
            assert m.id_to_token(i) == ref.lib.bert_vocab_id_to_token(ref.ctx, i), (shape, i)
        m.close()
        ref.close()
