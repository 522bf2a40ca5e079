"""Diagnostic for the open epilogue item (profiles/r1_head.md): with a -DBERT_EPI_LDX=2 build, count the accumulator words
whose EARLY read (both halves of a box loaded up front) differs from the LATE re-read at the shipped position.
    BERT_B200_LIB=embeddings.cpp_b200/lib/libbert_ldx2.so python tools/dbg_epi_ld.py"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import __graft_entry__ as ge  # noqa: E402

pkg = ge.load_package()
from tools import synth  # noqa: E402

model = synth.model_path("bge-base-en-v1.5", "q4_0")
texts = synth.texts_exact_len(synth.make_vocab(30522, 0), 512, 512, seed=0)
m = pkg.BertModel(model)
out = (ctypes.c_ulonglong * 2)()
prev = None
for it in range(3):
    e = m.encode(texts)
    rc = m.lib.bertx_debug_epi_ld(out)
    same = None if prev is None else bool(np.array_equal(e, prev))
    print(f"run {it}: rc {rc} mismatching words {out[0]} of {out[1]} compared; equal to previous run: {same}", flush=True)
    prev = e
