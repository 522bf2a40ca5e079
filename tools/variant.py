"""Build a variant of libbert.so with ONE source recompiled with extra -D flags (A/B and bisect runs):
    python tools/variant.py NAME source.cu -DFOO=1 [...]   ->  embeddings.cpp_b200/lib/variants/libbert_NAME.so
Select it at run time with BERT_B200_LIB (tools/ab.py alternates two builds on one box)."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200_build", os.path.join(ROOT, "embeddings.cpp_b200", "build.py"))
b = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b)
print(b.build_variant(sys.argv[1], sys.argv[2], sys.argv[3:], force=True))
