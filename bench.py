"""bench.py -- embeddings/sec of the bert.h hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

One process per GPU (torchrun for N > 1; RANK / LOCAL_RANK / WORLD_SIZE from the env).  A "step" is one pass of the
hot path over one batch of synthetic sentences; at the default workload (BASELINE.json configs[2], the configuration
the metric is quoted on) that is bge-base-en-v1.5 q4_0, 512 sentences x 512 tokens PER GPU (weak scaling: the global
batch of N x 512 sentences is generated once from one seed and sharded by rank with shard_range(); one weight replica
per GPU, no data-path collective).  The other workloads are BASELINE.json's remaining configs (+ the f16 A/B of config
3); `bge-large-q4_0-1M-s256` is config 5 as written: ONE job of 1 M sentences, strong scaling.

Prints ONE JSON line on rank 0.  `value` = whole-job embeddings/s with token ids resident in HBM (CUDA events on the
launch stream, max over ranks); `e2e` = the same through bert_encode_batch with host texts (tokenise + H2D + forward
+ D2H inside the timed region).  At N > 1 the line also carries `e2e_inprocess`: rank 0 alone, ONE bert_encode_batch
call over all N GPUs through the library's own in-process sharding (what a drop-in caller gets on a multi-GPU box).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# name: shape, weight type, sentences (per GPU for weak scaling, in total for strong), tokens per sentence (0 = variable),
# text generator, scaling
WORKLOADS = {
    "bge-base-q4_0-b512-s512": ("bge-base-en-v1.5", "q4_0", 512, 512, "exact", "weak"),          # configs[2], headline
    "bge-base-f16-b512-s512": ("bge-base-en-v1.5", "f16", 512, 512, "exact", "weak"),            # TMA-fed fp16 weights
    "minilm-q4_0-b1738-var": ("all-MiniLM-L6-v2", "q4_0", 1738, 0, "client", "weak"),            # configs[0]
    "bge-small-f16-b256-s128": ("bge-small-en-v1.5", "f16", 256, 128, "exact", "weak"),          # configs[1]
    "bge-small-zh-q4_1-b1024-var": ("bge-small-zh-v1.5", "q4_1", 1024, 0, "cjkvar", "weak"),     # configs[3]
    "bge-large-q4_0-b1024-s256": ("bge-large-en-v1.5", "q4_0", 1024, 256, "exact", "weak"),      # configs[4] shape, one batch
    "bge-large-q4_0-1M-s256": ("bge-large-en-v1.5", "q4_0", 1000000, 256, "templates", "strong"),  # configs[4] as written
    "tiny-q4_0-b64-s64": ("tiny-d64", "q4_0", 64, 64, "exact", "weak"),
}
DEFAULT_WORKLOAD = "bge-base-q4_0-b512-s512"
NCU_GEMM_DRAM_BYTES_PER_LAUNCH = 1.982e9   # measured (ncu --set full, mean over the 4 GEMMs of a layer at 262 144 tokens), see profiles/r2_rowwise_gemm_traffic.md


def shard_range(n_items, rank, world):
    """Contiguous shard [lo, hi) of n_items for `rank` of `world` (sizes differ by <= 1)."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def flops_per_sentence(H, I, n_layer, n):
    """Algorithmic FLOPs of one n-token sentence: GEMMs + attention."""
    gemm = n_layer * 2 * n * (4 * H * H + 2 * H * I)
    attn = n_layer * 4 * n * n * H
    return gemm, attn


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"tflops_burst": d["bf16_tflops"], "tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "hbm_gbs": d["hbm_gbs"], "source": "measured"}
    return {"tflops_burst": 1590.0, "tflops_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi SM clock / throttle-reason samples during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thr.join(timeout=2)
        sm = sorted(int(r[1]) for r in self.rows if len(r) >= 9 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) >= 9 and r[2].isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_texts(shape, kind, n, n_tok, seed=0):
    """The first n sentences of the workload's (seeded, deterministic) global batch."""
    from tools import synth
    vocab = synth.make_vocab(synth.SHAPES[shape][4], 0)      # regenerated (deterministic) instead of reading the model back
    if kind == "exact":
        return synth.texts_exact_len(vocab, n, n_tok, seed=seed)
    if kind == "client":
        return synth.sample_client_texts(vocab, n, seed=1738 + seed)
    if kind == "cjkvar":                                      # CJK / Latin / punctuation, 2..256 ids
        return synth.texts_var_len(vocab, n, 2, 256, seed=4 + seed, cjk=True)
    if kind == "templates":                                   # 4096 distinct templates (host RAM stays small)
        import numpy as np
        tpl = synth.texts_exact_len(vocab, 4096, n_tok, seed=5 + seed)
        pick = np.random.default_rng(5 + seed).integers(0, len(tpl), n)
        return [tpl[int(i)] for i in pick]
    raise ValueError(kind)


def model_file(shape, ftype):
    from tools import synth
    return synth.model_path(shape, ftype)


def padded_vs_valid(lens, batch=16):
    """Tokens the reference computes for these sentences (sort by length, slices of `batch`, pad to the slice's
    longest: bert.cpp:1413-1442, 837-843) over the valid tokens -- the GPU path packs and computes the valid ones only."""
    import numpy as np
    s = np.sort(np.asarray(lens))
    padded = sum(int(s[i:i + batch].max()) * len(s[i:i + batch]) for i in range(0, len(s), batch))
    return padded / max(int(s.sum()), 1)


def base_config(args, shape, ftype, n_sent, n_tok, scaling, world):
    return {"workload": args.workload, "model_shape": shape, "weights": ftype,
            "sentences_per_gpu": n_sent if scaling == "weak" else None,
            "sentences_total": n_sent * world if scaling == "weak" else n_sent, "seq_len": n_tok}


def run_reference(args, shape, ftype, n_sent, n_tok, kind, scaling, rank, world):
    """--impl reference: the reference's own ggml CPU path (oracle/_ref, compiled from /root/reference by oracle/Makefile)
    on a bounded sample of the same workload, all host threads."""
    if rank != 0:
        return
    from oracle import reflib
    sample = 2 if n_tok >= 256 else (8 if n_tok > 0 else 32)
    texts = make_texts(shape, kind, sample, n_tok)
    ref = reflib.RefBert(model_file(shape, ftype))
    toks = [ref.tokenize(t) for t in texts]
    for _ in range(args.warmup):
        ref.forward_tokens(toks[:1], chunk=1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ref.forward_tokens(toks, chunk=min(sample, 8))
    dt = time.perf_counter() - t0
    v = sample * args.steps / dt
    cores = os.cpu_count() or 1
    cfg = base_config(args, shape, ftype, n_sent, n_tok, scaling, world)
    cfg["sample_sentences_per_step"] = sample
    cfg["note"] = ("the reference cannot hold the full batch (12 GiB arena cap, bert.cpp:793-803: <= 10 sentences of 512 tokens): "
                   "each step is the first %d sentences of the same workload" % sample)
    base = {"value": v, "unit": "embeddings/s", "cores": cores, "kind": "reference",
            "sample": f"{sample} sentences x {n_tok or 'var'} tokens per step, bert_forward_batch, n_threads={cores}"}
    print(json.dumps({"metric": "embeddings/s", "value": v, "unit": "embeddings/s", "impl": "reference",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": scaling,
                      "vs_baseline": None, "dtype": "q4 x q8 int8 dot, f32 accumulate (ggml CPU)" if "q4" in ftype else ftype,
                      "data": "synthetic", "config": cfg, "cpu_baseline": base,
                      "e2e": {"value": v, "unit": "embeddings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def quiet_load(pkg, path):
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)                                         # keep the load banner off stdout
    try:
        return pkg.BertModel(path)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-inprocess", action="store_true", help="skip the e2e_inprocess leg at N > 1")
    args = ap.parse_args()
    shape, ftype, n_sent, n_tok, kind, scaling = WORKLOADS[args.workload]
    if args.steps is None:
        args.steps = 1 if scaling == "strong" else 10
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, shape, ftype, n_sent, n_tok, kind, scaling, rank, world)
        return

    os.environ["BERT_CUDA_DEVICES"] = str(local_rank)          # one replica per process
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load_package()
    if world > 1:
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        park = dist.new_group(backend="gloo")                   # host-side wait for the in-process leg (see below)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # rank 0 generates the (cached) model file first so ranks do not race on it
    if rank == 0:
        model_file(shape, ftype)
    barrier()
    path = model_file(shape, ftype)
    total = n_sent * world if scaling == "weak" else n_sent
    all_texts = make_texts(shape, kind, total, n_tok)           # the global batch, identical on every rank
    lo, hi = shard_range(total, rank, world)
    texts = all_texts[lo:hi]                                    # this rank's sentences
    m = quiet_load(pkg, path)
    hp = m.hparams()
    H, I, NL = hp["n_embd"], hp["n_intermediate"], hp["n_layer"]
    t0 = time.perf_counter()
    ids, lens = m.tokenize_batch(texts)
    tok_s = time.perf_counter() - t0
    T = int(lens.sum())
    m.stage(ids, lens)
    # a warm-up pass of a huge job is a prefix of it
    big = len(texts) > 65536
    if big:
        warm_n = 8192
        m.stage(ids[:int(lens[:warm_n].sum())], lens[:warm_n])
        m.run_staged(reps=args.warmup)
        m.stage(ids, lens)
    else:
        m.run_staged(reps=args.warmup)

    # ---- device-resident throughput (`value`)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    l0 = m.kernel_launches()
    ms, _ = m.run_staged(reps=args.steps)
    launches = m.kernel_launches() - l0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(ms)
    value = total * args.steps / (ms / 1e3)

    # ---- end to end through bert_encode_batch (host texts in, host floats out)
    for _ in range(2):
        m.encode(texts[:8192] if big else texts)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = m.encode(texts)
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = total * args.steps / e2e_s
    T_all = int(sum_over_ranks(float(T)))

    # ---- roofline of the dominant kernel (the tcgen05 GEMM), timed live with events
    if big:
        m.stage(ids[:int(lens[:8192].sum())], lens[:8192])
        prof_lens = lens[:8192]
    else:
        prof_lens = lens
    prof = m.profile_staged()
    peaks = measured_peaks()
    gemm_ms, gemm_launches = prof["gemm"]
    gemm_flops = sum(flops_per_sentence(H, I, NL, int(n))[0] for n in prof_lens)
    job_gemm_flops = sum(flops_per_sentence(H, I, NL, int(n))[0] for n in lens)
    job_attn_flops = sum(flops_per_sentence(H, I, NL, int(n))[1] for n in lens)
    achieved = gemm_flops / (gemm_ms / 1e3) / 1e12 if gemm_ms > 0 else 0.0
    total_prof_ms = sum(v[0] for v in prof.values())
    roofline = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05, q4 dequant fused)" if "q4" in ftype else "gemm_tc_kernel (tcgen05)",
                "achieved": achieved, "peak": peaks["tflops_sustained"], "unit": "TFLOP/s",
                "frac": achieved / peaks["tflops_sustained"], "traffic": None,
                "frac_of_burst_peak": achieved / peaks["tflops_burst"],
                "peak_source": peaks["source"] + " sustained bf16 (kernel timed inside a long step)",
                "launches_per_step": gemm_launches, "avg_launch_ms": gemm_ms / max(gemm_launches, 1),
                "algorithmic_flops_per_step": gemm_flops,
                "share_of_step": gemm_ms / total_prof_ms if total_prof_ms else None,
                "by_kernel_ms": {k: round(v[0], 3) for k, v in prof.items()},
                "whole_forward_tflops": (job_gemm_flops + job_attn_flops) / (ms / args.steps / 1e3) / 1e12}
    if args.workload == DEFAULT_WORKLOAD:
        roofline["traffic"] = NCU_GEMM_DRAM_BYTES_PER_LAUNCH
        roofline["traffic_note"] = "dram__bytes_read.sum + dram__bytes_write.sum, mean over the 4 GEMMs of a layer at 262144 tokens per launch: 1.56 / 1.56 / 2.01 / 2.80 GB against algorithmic 1.61 / 1.61 / 2.01 / 2.82 GB (profiles/r2_rowwise_gemm_traffic.md)"
    # the HBM-bound row kernels against the measured copy bandwidth (algorithmic bytes)
    wbytes = {"f32": 4.0, "f16": 2.0, "q4_0": 18.0 / 32, "q4_1": 20.0 / 32}[ftype]
    twin = 2 if "q4" in ftype else 1
    Tp, Bp = int(prof_lens.sum()), len(prof_lens)
    hbm = {}
    if prof["embed_ln"][0] > 0:
        b = Tp * (H * wbytes + 2 * H * twin)
        hbm["embed_ln"] = {"GB/s": b / (prof["embed_ln"][0] / 1e3) / 1e9, "bytes_per_step": int(b)}
    if prof["pool_norm"][0] > 0:
        b = Tp * H * 2 + Bp * H * 4
        hbm["pool_norm"] = {"GB/s": b / (prof["pool_norm"][0] / 1e3) / 1e9, "bytes_per_step": int(b)}
    for v in hbm.values():
        v["frac_of_hbm_peak"] = v["GB/s"] / peaks["hbm_gbs"]
    roofline["hbm_bound_kernels"] = hbm

    cfg = base_config(args, shape, ftype, n_sent, n_tok, scaling, world)
    cfg.update({"tokens_per_step": T_all, "parallelism": f"dp{world} (sentence shards by shard_range, no collective)",
                "l2": "inputs larger than L2: activations of one micro-batch are %.0f MB per buffer" % (min(T, 262144) * H * 2 / 1e6)})
    if n_tok == 0:
        cfg["padded_vs_valid_tokens"] = round(padded_vs_valid(lens), 3)
        cfg["length_min_mean_max"] = [int(lens.min()), float(lens.mean()), int(lens.max())]
    line = {"metric": "embeddings/s", "value": value, "unit": "embeddings/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f16 x f16 -> f32 (tcgen05 kind::f16; q4 weights dequantised in the mainloop)",
            "data": "synthetic", "config": cfg,
            "e2e": {"value": e2e_value, "unit": "embeddings/s", "h2d_bytes_per_step": int(T_all * 4 + (total + world) * 4),
                    "d2h_bytes_per_step": int(total * H * 4)},
            "host": {"tokenizer_tokens_per_s_per_rank": T / tok_s if tok_s > 0 else None, "host_threads": m.n_threads,
                     "gpu_busy_fraction_e2e": (ms / args.steps / 1e3) / (e2e_s / args.steps)},
            "gpu_launches": int(launches), "roofline": roofline, "clocks": clocks}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import reflib                               # bench cpu_baseline leg (checker timed, not shipped)
        if reflib.available():
            sample = 4 if n_tok >= 256 else 16
            ref = reflib.RefBert(path)
            toks = [ref.tokenize(t) for t in texts[:sample]]
            ref.forward_tokens(toks[:1], chunk=1)
            t0 = time.perf_counter()
            want = ref.forward_tokens(toks, chunk=min(sample, 8))
            dt = time.perf_counter() - t0
            cos = (out[:sample] * want).sum(1) / np.linalg.norm(out[:sample], axis=1) / np.linalg.norm(want, axis=1)
            line["cpu_baseline"] = {"value": sample / dt, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "reference",
                                    "sample": f"{sample} sentences x {n_tok or 'var'} tokens, bert_forward_batch of oracle/_ref/libbert_ref.so",
                                    "parity_min_cosine": float(cos.min()), "parity_max_abs": float(np.abs(out[:sample] - want).max())}
            ref.close()
        else:
            line["cpu_baseline"] = {"value": None, "unit": "embeddings/s", "cores": os.cpu_count(), "kind": "reference",
                                    "sample": "oracle/_ref not built on this box"}
    m.close()

    # ---- the product's own multi-GPU mode: ONE bert_encode_batch call over all N GPUs of this process (rank 0 only)
    # The other ranks wait on a gloo (host) group: a rank parked in an NCCL barrier keeps a spinning kernel on its GPU,
    # which rank 0's replica on that GPU would have to share the SMs with.
    if world > 1 and not args.no_inprocess:
        barrier()
        if rank == 0:
            os.environ["BERT_CUDA_DEVICES"] = ",".join(str(i) for i in range(world))
            mN = quiet_load(pkg, path)
            warm = all_texts[:8192 * world] if big else all_texts
            for _ in range(2):
                mN.encode(warm)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                outN = mN.encode(all_texts)
            dtN = time.perf_counter() - t0
            line["e2e_inprocess"] = {"value": total * args.steps / dtN, "unit": "embeddings/s", "n_devices": mN.n_devices(),
                                     "host_threads": mN.n_threads, "calls_per_step": 1,
                                     "finite": bool(np.isfinite(outN).all()),
                                     "ratio_to_torchrun_e2e": (total * args.steps / dtN) / e2e_value}
            mN.close()
        dist.barrier(group=park)
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
