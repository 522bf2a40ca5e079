"""GPU bring-up driver: runs each kernel self-test / parity step in its own subprocess with a
timeout, so that a trap or hang in one step cannot take the others (or the box) down.
Writes gpurun_out/bringup.log.  Usage on the GPU box:  python tools/gpu_bringup.py [step ...]
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")

STEP_CODE = r'''
import sys, os, json, time
sys.path.insert(0, %(root)r)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
step = %(step)r

def cmp(a, b):
    cos = (a * b).sum(1) / np.linalg.norm(a, axis=1) / np.linalg.norm(b, axis=1)
    return float(cos.min()), float(np.abs(a - b).max())

if step.startswith("gemm:"):
    _, wt, M, N, K, epi = step.split(":")
    rc, err, ref = pkg.selftest_gemm(int(M), int(N), int(K), int(wt), int(epi), 1)
    print(json.dumps({"step": step, "rc": rc, "max_err": err, "max_ref": ref}))
elif step.startswith("attn:"):
    f = step.split(":")                      # attn:n_sent:max_len:n_head:d_head[:seed]  (seed bit 1024: ramped keys, 2048: mma.sync kernel)
    ns, L, nh, d = f[1:5]
    rc, err, ref = pkg.selftest_attention(int(ns), int(L), int(nh), int(d), int(f[5]) if len(f) > 5 else 1)
    print(json.dumps({"step": step, "rc": rc, "max_err": err, "max_ref": ref}))
elif step.startswith("parity:"):
    # parity:<shape>:<ftype>:<n_texts>:<max_len>
    _, shape, ft, n, L = step.split(":")
    from tools import synth
    from oracle import fileformat, forward, tokenizer, reflib
    path = synth.model_path(shape, ft)
    mf = fileformat.read_model(path)
    texts = synth.texts_var_len(mf.vocab, int(n), 2, int(L), seed=11, cjk=True)
    m = pkg.BertModel(path)
    t0 = time.time(); got = m.encode(texts); dt = time.time() - t0
    voc = tokenizer.Vocab(mf.vocab)
    toks = [tokenizer.tokenize(voc, t, mf.hparams.n_max_tokens) for t in texts]
    res = {"step": step, "encode_s": dt, "launches": m.kernel_launches(), "finite": bool(np.isfinite(got).all())}
    if reflib.available():
        ref = reflib.RefBert(path)
        want = ref.forward_tokens(toks, chunk=4)
        res["vs_ref"] = cmp(got, want)
    if mf.hparams.n_embd <= 256:
        om = forward.OracleModel(mf)
        res["vs_oracle"] = cmp(got, om.forward(toks))
        res["vs_exact"] = cmp(got, om.forward(toks, exact_activations=True))
    print(json.dumps(res))
'''

DEFAULT_STEPS = [
    # tcgen05 GEMM vs SIMT validation kernel: wtype:M:N:K:epilogue
    "gemm:1:128:64:64:0", "gemm:1:256:256:128:1", "gemm:1:300:192:256:5", "gemm:1:1000:768:768:3",
    "gemm:2:128:64:64:0", "gemm:2:300:256:256:1", "gemm:2:1000:768:768:7", "gemm:2:515:192:384:11",
    "gemm:3:128:128:128:1", "gemm:3:777:1024:512:15", "gemm:1:4096:3072:768:3", "gemm:2:4096:768:3072:5",
    "attn:5:64:2:64", "attn:7:200:4:64", "attn:3:512:12:64", "attn:9:130:12:32", "attn:4:512:2:32",
]


def run_step(step, env_extra=None, timeout=240):
    env = dict(os.environ)
    env.update(env_extra or {})
    code = STEP_CODE % {"root": ROOT, "step": step}
    t0 = time.time()
    try:
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=timeout, env=env)
        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
        rec = json.loads(lines[-1]) if lines else {"step": step, "error": "no result"}
        rec["exit"] = r.returncode
        if r.returncode != 0 or not lines:
            rec["stderr"] = r.stderr[-1500:]
    except subprocess.TimeoutExpired:
        rec = {"step": step, "error": "timeout"}
    rec["wall_s"] = round(time.time() - t0, 2)
    if env_extra:
        rec["env"] = env_extra
    return rec


def main():
    os.makedirs(OUT, exist_ok=True)
    steps = sys.argv[1:] or DEFAULT_STEPS
    with open(os.path.join(OUT, "bringup.log"), "a") as log:
        for s in steps:
            env = None
            if "@" in s:                         # step@KEY=VAL,KEY=VAL
                s, e = s.split("@", 1)
                env = dict(kv.split("=", 1) for kv in e.split(","))
            rec = run_step(s, env)
            line = json.dumps(rec)
            print(line, flush=True)
            log.write(line + "\n")
            log.flush()


if __name__ == "__main__":
    main()
