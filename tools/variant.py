"""Build a variant of libbert.so with ONE source recompiled with extra -D flags (for tools/ab.py):
    python tools/variant.py NAME source.cu -DFOO=1 [...]   ->  embeddings.cpp_b200/lib/libbert_NAME.so"""
import importlib.util
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("b200_build", os.path.join(ROOT, "embeddings.cpp_b200", "build.py"))
b = importlib.util.module_from_spec(spec)
spec.loader.exec_module(b)
b.build(verbose=False)
name, src, flags = sys.argv[1], sys.argv[2], sys.argv[3:]
stem = src.rsplit(".", 1)[0]
obj = os.path.join(b.OBJ, f"{stem}__{name}.o")
subprocess.run([b.NVCC] + b.FLAGS + flags + ["-Xptxas", "-v", "-c", os.path.join(b.SRC, src), "-o", obj], check=True,
               stderr=open(os.path.join(b.OBJ, f"{stem}__{name}.ptxas.log"), "w"))
objs = [obj if s.rsplit(".", 1)[0] == stem else os.path.join(b.OBJ, s.rsplit(".", 1)[0] + ".o") for s in b.SOURCES]
lib = os.path.join(b.LIB, f"libbert_{name}.so")
subprocess.run([b.NVCC, "-shared", "-cudart", "static", "-o", lib] + objs + ["-lpthread", "-ldl", "-lrt"], check=True)
print(lib)
