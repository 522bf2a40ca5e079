"""bertx_convert_safetensors / bertx_load_from_safetensors against the reference's own tools: the golden
hashes in tests/golden/hf_tiny/expected.json are of the files models/convert-to-ggml.py and models/quantize.cpp (compiled
as they are) wrote for the checkpoint in that directory (tests/golden/make_hf_golden.py)."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

HF = os.path.join(ROOT, "tests", "golden", "hf_tiny")
EXPECTED = json.load(open(os.path.join(HF, "expected.json")))


def _sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


@pytest.mark.parametrize("name,ftype", [("f32", 0), ("f16", 1)])
def test_float_conversion_is_byte_identical_to_the_reference_script(tmp_path, name, ftype):
    """ftype 0 / 1 need no GPU: header, vocabulary (`line[:-1]`), tensor order, fp16 rounding."""
    pkg = ge.load_package()
    out = str(tmp_path / f"ggml-model-{name}.bin")
    pkg.convert_safetensors(HF, out, ftype)
    assert os.path.getsize(out) == EXPECTED["bytes"][name]
    assert _sha(out) == EXPECTED["sha256"][name]


def test_conversion_reports_bad_input(tmp_path):
    pkg = ge.load_package()
    with pytest.raises(RuntimeError):
        pkg.convert_safetensors(str(tmp_path), str(tmp_path / "x.bin"), 1)          # no such checkpoint
    with pytest.raises(RuntimeError):
        pkg.convert_safetensors(HF, str(tmp_path / "x.bin"), 7)                     # no such type
    bad = tmp_path / "ckpt"
    bad.mkdir()
    for f in ("config.json", "vocab.txt"):
        (bad / f).write_bytes(open(os.path.join(HF, f), "rb").read())
    (bad / "model.safetensors").write_bytes(open(os.path.join(HF, "model.safetensors"), "rb").read()[:4096])   # truncated
    with pytest.raises(RuntimeError):
        pkg.convert_safetensors(str(bad), str(tmp_path / "x.bin"), 0)


@pytest.mark.gpu
@pytest.mark.parametrize("name,ftype", [("q4_0", 2), ("q4_1", 3)])
def test_q4_conversion_is_byte_identical_to_convert_then_quantize(tmp_path, name, ftype):
    pkg = ge.load_package()
    out = str(tmp_path / f"ggml-model-{name}.bin")
    pkg.convert_safetensors(HF, out, ftype)
    assert os.path.getsize(out) == EXPECTED["bytes"][name]
    assert _sha(out) == EXPECTED["sha256"][name]


@pytest.mark.gpu
@pytest.mark.parametrize("ftype", [1, 2, 3])
def test_loading_the_checkpoint_directly_equals_loading_the_converted_file(tmp_path, ftype):
    pkg = ge.load_package()
    out = str(tmp_path / "m.bin")
    pkg.convert_safetensors(HF, out, ftype)
    texts = [b"hello world", b"the a dup", "一丁 hello".encode(), b""]
    a = pkg.BertModel(out)
    want = a.encode(texts)
    ids = [a.tokenize(t) for t in texts]
    a.close()
    b = pkg.BertModel.from_safetensors(HF, ftype)
    got = b.encode(texts)
    assert [b.tokenize(t) for t in texts] == ids
    b.close()
    assert np.array_equal(got, want)                       # same bytes on the device, same kernels: bit-identical
    assert np.all(np.isfinite(got)) and np.allclose(np.linalg.norm(got, axis=1), 1.0, atol=1e-3)


def test_convert_tool_has_the_reference_scripts_command_line(tmp_path):
    """examples/convert_b200: `dir-model [ftype]` -> dir-model/ggml-model-{f32,f16}.bin with the script's bytes, default f16,
    usage text and exit code 1 without arguments, 'Invalid ftype' for a bad type (convert-to-ggml.py:10-51)."""
    import shutil
    import subprocess
    ge.load_package()                                                     # builds libbert.so if needed
    assert subprocess.run(["make", "-C", os.path.join(ROOT, "examples"), "bin/convert_b200"], capture_output=True).returncode == 0
    tool = os.path.join(ROOT, "examples", "bin", "convert_b200")
    ckpt = tmp_path / "ckpt"
    shutil.copytree(HF, ckpt)
    r = subprocess.run([tool], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage" in r.stderr
    r = subprocess.run([tool, str(ckpt), "9"], capture_output=True, text=True)
    assert r.returncode == 1 and "Invalid ftype" in r.stderr
    for args, name in (([str(ckpt), "0"], "f32"), ([str(ckpt)], "f16")):
        r = subprocess.run([tool] + args, capture_output=True, text=True)
        out = ckpt / f"ggml-model-{name}.bin"
        assert r.returncode == 0 and f"Done. Output file: {out}" in r.stdout, r.stderr
        assert _sha(str(out)) == EXPECTED["sha256"][name]
    r = subprocess.run([tool, str(tmp_path / "missing")], capture_output=True, text=True)
    assert r.returncode == 1
