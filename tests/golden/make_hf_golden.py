"""Generates tests/golden/hf_tiny: a tiny random Hugging Face BERT checkpoint and the SHA-256 of what the REFERENCE's
own tools make of it (run in the build container, where /root/reference and oracle/_ref exist):
    python models/convert-to-ggml.py dir 0|1          (the unmodified script, models/convert-to-ggml.py)
    quantize ggml-model-f16.bin out.bin 2|3           (oracle/_ref/quantize = models/quantize.cpp compiled as it is)
tests/test_convert.py checks bertx_convert_safetensors against these hashes.
    python tests/golden/make_hf_golden.py
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

import torch
from transformers import BertConfig, BertModel, BertTokenizerFast

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tools import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "hf_tiny")
REF_SCRIPT = "/root/reference/models/convert-to-ggml.py"
QUANTIZE = os.path.join(ROOT, "oracle", "_ref", "quantize")


def sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def main():
    work = tempfile.mkdtemp(prefix="hf_tiny_")
    d = os.path.join(work, "tiny-bert")
    os.makedirs(d)
    cfg = BertConfig(vocab_size=300, hidden_size=64, num_hidden_layers=2, num_attention_heads=2, intermediate_size=128,
                     max_position_embeddings=32, type_vocab_size=2)
    torch.manual_seed(20261105)
    model = BertModel(cfg)
    with torch.no_grad():                                           # BERT init is N(0, 0.02): give biases / LayerNorm some life
        for n, p_ in model.named_parameters():
            if n.endswith("bias"):
                p_.normal_(0.0, 0.05)
            elif "LayerNorm.weight" in n:
                p_.normal_(1.0, 0.1)
    model.save_pretrained(d, safe_serialization=True)
    vocab = synth.make_vocab(300)
    with open(os.path.join(d, "vocab.txt"), "wb") as f:
        for w in vocab:
            f.write(w + b"\n")
    BertTokenizerFast(vocab_file=os.path.join(d, "vocab.txt"), do_lower_case=True).save_pretrained(d)
    expected = {}
    for ft in (0, 1):
        subprocess.run([sys.executable, REF_SCRIPT, d, str(ft)], check=True, stdout=subprocess.DEVNULL)
    f32, f16 = os.path.join(d, "ggml-model-f32.bin"), os.path.join(d, "ggml-model-f16.bin")
    expected["f32"], expected["f16"] = sha(f32), sha(f16)
    for name, it in (("q4_0", 2), ("q4_1", 3)):
        out = os.path.join(d, f"ggml-model-{name}.bin")
        subprocess.run([QUANTIZE, f16, out, str(it)], check=True, stdout=subprocess.DEVNULL)
        expected[name] = sha(out)
    sizes = {k: os.path.getsize(os.path.join(d, f"ggml-model-{k}.bin")) for k in expected}
    os.makedirs(OUT, exist_ok=True)
    for fn in ("model.safetensors", "config.json", "vocab.txt"):
        shutil.copy(os.path.join(d, fn), os.path.join(OUT, fn))
    json.dump({"sha256": expected, "bytes": sizes,
               "made_by": "tests/golden/make_hf_golden.py: the reference's convert-to-ggml.py and quantize, unmodified"},
              open(os.path.join(OUT, "expected.json"), "w"), indent=1)
    print(json.dumps({"sha256": expected, "bytes": sizes}, indent=1))
    shutil.rmtree(work)


if __name__ == "__main__":
    main()
