"""Drop-in proof: the reference's own example programs, compiled UNMODIFIED from
/root/reference/examples against include/bert.h + include/ggml.h and linked with this repo's libbert.so and
the libggml.so timer shim (examples/Makefile `ref` target -> examples/_ref_build/, built by
__graft_entry__.build(); the binaries travel to the GPU box, the sources are never copied), plus this repo's
batching TCP server (examples/server_b200.cpp, same wire protocol as examples/server.cpp:104-116).
"""
import os
import re
import socket
import struct
import subprocess
import threading

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from oracle import fileformat
from tools import synth

EX = os.path.join(ROOT, "examples")
REF_BIN = os.path.join(EX, "_ref_build")
HAVE_REF_SRC = os.path.isdir("/root/reference/examples")
TEXTS3 = ["你好世界", "こんにちは、世界！", "hello world"]        # examples/test_batch_encode.cpp:39-43


def _make(*targets):
    subprocess.run(["make", "-C", EX, *targets], check=True, capture_output=True)


def _floats(line):
    return np.array([float(x) for x in re.findall(r"-?\d+\.\d+", line)], dtype=np.float32)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _recv_exact(sock, n):
    buf = b""
    while len(buf) < n:
        chunk = sock.recv(n - len(buf))
        assert chunk, "server closed the connection"
        buf += chunk
    return buf


class _Server:
    def __init__(self, binary, model, port, env=None):
        self.p = subprocess.Popen([binary, "-m", model, "--port", str(port)], stdout=subprocess.PIPE,
                                  stderr=subprocess.STDOUT, text=True, env=dict(os.environ, **(env or {})))
        self.port = None
        for line in self.p.stdout:                          # load banner, then the reference's "Server running" line
            m = re.search(r"Server running on port (\d+)", line)
            if m:
                self.port = int(m.group(1))
                break
        assert self.port, "server did not start"

    def stop(self):
        self.p.terminate()
        try:
            self.p.wait(timeout=20)
        except subprocess.TimeoutExpired:
            self.p.kill()


def _client(port, texts):
    """examples/sample_client.py: read int32 n_embd, then one send per text / n_embd floats back."""
    with socket.create_connection(("127.0.0.1", port), timeout=60) as s:
        n_embd = struct.unpack("<i", _recv_exact(s, 4))[0]
        out = np.zeros((len(texts), n_embd), np.float32)
        for i, t in enumerate(texts):
            s.sendall(t if isinstance(t, bytes) else t.encode())
            out[i] = np.frombuffer(_recv_exact(s, 4 * n_embd), np.float32)
    return out


# ------------------------------------------------------------------ CPU: builds + loud failure without a GPU
@pytest.mark.skipif(not HAVE_REF_SRC, reason="the reference sources only exist in the build container")
def test_reference_examples_compile_and_link_unmodified(pkg):
    _make("ref")
    for name in ("main", "server", "test_tokenizer", "test_batch_encode", "dylib"):
        assert os.access(os.path.join(REF_BIN, name), os.X_OK), name
    nm = subprocess.run(["nm", "-D", "--undefined-only", os.path.join(REF_BIN, "main")], capture_output=True, text=True).stdout
    for sym in ("bert_load_from_file", "bert_tokenize", "bert_forward", "bert_vocab_id_to_token", "ggml_time_us", "ggml_time_init"):
        assert sym in nm, sym


def test_ggml_timer_shim_exports(pkg):
    import ctypes
    lib = ctypes.CDLL(os.path.join(os.path.dirname(pkg.LIB_PATH), "libggml.so"))
    for f in ("ggml_time_ms", "ggml_time_us", "ggml_cycles", "ggml_cycles_per_ms"):
        getattr(lib, f).restype = ctypes.c_int64
    lib.ggml_time_init()
    a = lib.ggml_time_us()
    b = lib.ggml_time_us()
    assert 0 <= b - a < 1_000_000 and abs(lib.ggml_time_ms() - b // 1000) <= 1000 and lib.ggml_cycles_per_ms() > 0


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure path")
def test_server_b200_fails_loudly_without_a_gpu(pkg):
    _make("all")
    model = synth.model_path("tiny-d64", "q4_0")
    r = subprocess.run([os.path.join(EX, "bin", "server_b200"), "-m", model, "--port", "0"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "failed to load model" in r.stderr


# ------------------------------------------------------------------ GPU: the binaries give the library's embeddings
@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(os.path.join(REF_BIN, "main")), reason="examples/_ref_build did not travel")
def test_reference_main_and_batch_encode_binaries(pkg):
    model = synth.model_path("tiny-d64", "q4_0")
    m = pkg.BertModel(model)
    prompt = "hello world , the dup"
    out = subprocess.run([os.path.join(REF_BIN, "main"), "-m", model, "-p", prompt], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    lines = [l for l in out.stdout.splitlines() if l.startswith("[")]
    ids = [int(x) for x in re.findall(r"\d+", lines[0])]
    assert ids == m.tokenize(prompt)
    want = m.forward_tokens([ids])[0]
    got = _floats(lines[1])
    assert got.shape == want.shape and np.abs(got - want).max() <= 1e-4       # printed with %1.4f
    assert "eval time" in out.stdout
    out = subprocess.run([os.path.join(REF_BIN, "test_batch_encode"), "-m", model], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    rows = [_floats(l) for l in out.stdout.splitlines() if l.startswith("[")]
    want = m.encode(TEXTS3)
    assert len(rows) == 3
    for r, w in zip(rows, want):
        assert np.abs(r - w[:10]).max() <= 1e-6                               # printed with %1.9f
    m.close()


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(os.path.join(REF_BIN, "server")), reason="examples/_ref_build did not travel")
def test_reference_server_binary_single_client(pkg):
    model = synth.model_path("tiny-d64", "q4_1")
    m = pkg.BertModel(model)
    srv = _Server(os.path.join(REF_BIN, "server"), model, _free_port())
    try:
        texts = ["hello world", "你好世界", "the dup ##ing"]
        got = _client(srv.port, texts)
        assert np.abs(got - m.encode(texts)).max() <= 4e-3
    finally:
        srv.stop()
        m.close()


@pytest.mark.gpu
def test_server_b200_batches_concurrent_clients(pkg):
    _make("all")
    model = synth.model_path("tiny-d64", "q4_0")
    mf = fileformat.read_model(model)
    m = pkg.BertModel(model)
    n_clients, per = 24, 12
    texts = synth.texts_var_len(mf.vocab, n_clients * per, 3, 60, seed=11, cjk=True)
    assert all(texts)                                       # an empty send would not reach the server
    want = m.encode(texts)
    srv = _Server(os.path.join(EX, "bin", "server_b200"), model, 0, env={"BERT_SERVER_BATCH_US": "2000"})
    try:
        got = [None] * n_clients
        errs = []

        def work(c):
            try:
                got[c] = _client(srv.port, texts[c * per:(c + 1) * per])
            except Exception as e:                                              # noqa: BLE001
                errs.append(e)
        th = [threading.Thread(target=work, args=(c,)) for c in range(n_clients)]
        [t.start() for t in th]
        [t.join() for t in th]
        assert not errs, errs
        # batch composition decides between the latency path (<= 256 tokens) and the fused-LayerNorm path: rounding only
        assert np.abs(np.concatenate(got) - want).max() <= 4e-3
    finally:
        srv.p.terminate()
        tail = srv.p.stdout.read()
        srv.p.wait(timeout=20)
    mt = re.search(r"served (\d+) requests in (\d+) batches", tail)
    assert mt and int(mt.group(1)) == n_clients * per and int(mt.group(2)) < n_clients * per, tail
    m.close()
