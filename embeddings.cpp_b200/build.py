"""Build libbert.so (the drop-in C-ABI library) in-tree with nvcc for sm_100a only.

    python embeddings.cpp_b200/build.py [--force]

Output: embeddings.cpp_b200/lib/libbert.so (git-ignored; travels to the GPU box).
cudart is linked statically so the library is self-contained for ctypes / dlopen users
(the reference's libbert.so needs a separate libggml.so).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "lib")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden,-pthread", "-I", os.path.join(HERE, "..", "include")]
if os.environ.get("BERT_DEBUG_WAIT"):
    FLAGS.append("-DBERT_DEBUG_WAIT")
SOURCES = ["gemm_tc.cu", "attention.cu", "attention_tc.cu", "rowwise.cu", "upload.cu", "engine_load.cu", "engine_run.cu",
           "api.cu", "api_ext.cu", "quantize.cu", "tokenizer.cpp", "model_file.cpp"]


def _newer(a, b):
    return not os.path.exists(b) or os.path.getmtime(a) > os.path.getmtime(b)


def build(force=False, verbose=True):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIB, exist_ok=True)
    headers = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".h", ".cuh"))]
    headers += [os.path.join(HERE, "..", "include", f) for f in ("bert.h", "bert_b200.h")]
    hdr_time = max(os.path.getmtime(h) for h in headers)
    jobs = []
    objs = []
    for s in SOURCES:
        src = os.path.join(SRC, s)
        obj = os.path.join(OBJ, s.rsplit(".", 1)[0] + ".o")
        objs.append(obj)
        if force or _newer(src, obj) or os.path.getmtime(obj) < hdr_time:
            jobs.append([NVCC] + FLAGS + ["-c", src, "-o", obj])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for out in ex.map(run, jobs):
            if verbose and out.strip():
                print(out, file=sys.stderr)
    # timer shim the reference's example programs link as `ggml` (include/ggml.h)
    shim_src, shim = os.path.join(HERE, "shim", "ggml_time.c"), os.path.join(LIB, "libggml.so")
    if force or _newer(shim_src, shim):
        run([os.environ.get("CC", "gcc"), "-O2", "-fPIC", "-shared", "-I", os.path.join(HERE, "..", "include"), "-o", shim, shim_src])
    lib = os.path.join(LIB, "libbert.so")
    if jobs or not os.path.exists(lib):
        run([NVCC, "-shared", "-cudart", "static", "-o", lib] + objs + ["-lpthread", "-ldl", "-lrt"])
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
