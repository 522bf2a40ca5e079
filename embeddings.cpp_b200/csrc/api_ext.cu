// Measurement entry points (bert_b200.h): device-resident jobs timed with CUDA events on
// the launch stream, per-kernel-class profiling, and the kernel self-tests.
#include <cstdio>
#include <cstring>
#include <random>
#include <vector>

#include "../../include/bert_b200.h"
#include "engine.h"

using namespace b200;

extern "C" {

int32_t bertx_stage(struct bert_ctx * ctx, int32_t dev, int32_t n_sent, const bert_vocab_id * tokens_packed,
                    const int32_t * n_tokens) {
    Engine & e = ctx->eng;
    if (dev < 0 || dev >= (int) e.reps.size() || n_sent <= 0) return -1;
    std::lock_guard<std::mutex> lock(e.mu);
    Replica & r = e.reps[dev];
    if (!cuda_ok(cudaSetDevice(r.device), "cudaSetDevice")) return -2;
    cudaStreamSynchronize(r.stream);
    e.free_job(r);
    for (int i = 0; i < n_sent; ++i) if (n_tokens[i] <= 0 || n_tokens[i] > e.hp.n_max_tokens) return -3;
    e.plan(n_tokens, n_sent, r.job_order, r.job_mbs);
    std::vector<size_t> src_off(n_sent + 1, 0);
    for (int i = 0; i < n_sent; ++i) src_off[i + 1] = src_off[i] + (size_t) n_tokens[i];
    std::vector<int32_t> ids(src_off[n_sent]);
    std::vector<int32_t> cu;
    r.job_cu_off.clear();
    r.job_nspans.clear();
    std::vector<int32_t> spans;
    size_t t = 0;
    for (const MicroBatch & mb : r.job_mbs) {
        r.job_cu_off.push_back(cu.size());
        const size_t k0 = cu.size();
        int local = 0;
        for (int i = 0; i < mb.count; ++i) {
            const int idx = r.job_order[mb.first + i];
            cu.push_back(local);
            memcpy(ids.data() + t, tokens_packed + src_off[idx], (size_t) n_tokens[idx] * 4);
            t += n_tokens[idx]; local += n_tokens[idx];
        }
        cu.push_back(local);
        // the micro-batch's span boundaries follow its cu_seqlens, as in the staging slots of the encode path
        spans.assign((size_t) mb.count + 1, 0);
        const int ns = pack_spans(cu.data() + k0, mb.count, spans.data());
        cu.insert(cu.end(), spans.begin(), spans.begin() + ns + 1);
        r.job_nspans.push_back(ns);
    }
    if (!cuda_ok(cudaMalloc((void **) &r.job_ids, ids.size() * 4 + 256), "cudaMalloc job ids")) return -4;
    if (!cuda_ok(cudaMalloc((void **) &r.job_cu, cu.size() * 4 + 256), "cudaMalloc job cu")) return -4;
    if (!cuda_ok(cudaMalloc((void **) &r.job_out, (size_t) n_sent * e.hp.n_embd * 4 + 256), "cudaMalloc job out")) return -4;
    if (!cuda_ok(cudaMemcpy(r.job_ids, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice), "H2D job ids")) return -5;
    if (!cuda_ok(cudaMemcpy(r.job_cu, cu.data(), cu.size() * 4, cudaMemcpyHostToDevice), "H2D job cu")) return -5;
    r.job_sent = n_sent;
    return 0;
}

static bool run_job_once(Engine & e, Replica & r) {
    for (size_t k = 0; k < r.job_mbs.size(); ++k) {
        const MicroBatch & mb = r.job_mbs[k];
        if (!e.forward(r, r.job_ids + mb.tok_offset, r.job_cu + r.job_cu_off[k], mb.count, mb.tokens, mb.max_len,
                       r.job_out + (size_t) mb.first * e.hp.n_embd, r.job_nspans[k])) return false;
    }
    return true;
}

static int32_t copy_job_out(Engine & e, Replica & r, float * out) {
    const int H = e.hp.n_embd;
    std::vector<float> tmp((size_t) r.job_sent * H);
    if (!cuda_ok(cudaMemcpy(tmp.data(), r.job_out, tmp.size() * 4, cudaMemcpyDeviceToHost), "D2H job out")) return -8;
    for (int i = 0; i < r.job_sent; ++i) memcpy(out + (size_t) r.job_order[i] * H, tmp.data() + (size_t) i * H, (size_t) H * 4);
    return 0;
}

int32_t bertx_run_staged(struct bert_ctx * ctx, int32_t dev, int32_t reps, float * ms_total, float * out) {
    Engine & e = ctx->eng;
    if (dev < 0 || dev >= (int) e.reps.size() || reps <= 0) return -1;
    std::lock_guard<std::mutex> lock(e.mu);
    Replica & r = e.reps[dev];
    if (r.job_sent <= 0) return -2;
    if (!cuda_ok(cudaSetDevice(r.device), "cudaSetDevice")) return -3;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    bool ok = cuda_ok(cudaStreamSynchronize(r.stream), "pre sync");
    cudaEventRecord(a, r.stream);
    for (int i = 0; i < reps && ok; ++i) ok = run_job_once(e, r);
    cudaEventRecord(b, r.stream);
    ok = cuda_ok(cudaStreamSynchronize(r.stream), "run_staged sync") && ok;
    float ms = 0.f;
    if (ok) cudaEventElapsedTime(&ms, a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    if (ms_total) *ms_total = ms;
    if (!ok) return -7;
    if (out) return copy_job_out(e, r, out);
    return 0;
}

int32_t bertx_profile_staged(struct bert_ctx * ctx, int32_t dev, float * ms_by_kind, int64_t * launches_by_kind) {
    Engine & e = ctx->eng;
    if (dev < 0 || dev >= (int) e.reps.size()) return -1;
    std::lock_guard<std::mutex> lock(e.mu);
    Replica & r = e.reps[dev];
    if (r.job_sent <= 0) return -2;
    if (!cuda_ok(cudaSetDevice(r.device), "cudaSetDevice")) return -3;
    for (int k = 0; k < KK_COUNT; ++k) { ms_by_kind[k] = 0.f; launches_by_kind[k] = 0; }
    r.profiling = true;
    const bool ran = run_job_once(e, r);
    r.profiling = false;
    const bool ok = cuda_ok(cudaStreamSynchronize(r.stream), "profile sync") && ran;
    for (Replica::Span & s : r.spans) {
        float ms = 0.f;
        if (ok && cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) {
            ms_by_kind[s.kind] += ms;
            launches_by_kind[s.kind] += 1;
        }
        cudaEventDestroy(s.a); cudaEventDestroy(s.b);
    }
    r.spans.clear();
    return ok ? 0 : -7;
}

// ---------------------------------------------------------------------------- self-tests
}  // extern "C"

namespace {
struct DevBuf {
    void * p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    bool alloc(size_t bytes) { return cudaMalloc(&p, bytes + 256) == cudaSuccess; }
    template <typename T> T * as() { return (T *) p; }
};
uint16_t f2h_bits(float f) { __half h = __float2half_rn(f); uint16_t u; memcpy(&u, &h, 2); return u; }

// compare two fp16 device arrays; returns max |a-b| and max |b|
bool compare_h(const __half * da, const __half * db, size_t n, float * max_err, float * max_ref) {
    std::vector<__half> a(n), b(n);
    if (cudaMemcpy(a.data(), da, n * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    if (cudaMemcpy(b.data(), db, n * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    float me = 0.f, mr = 0.f;
    for (size_t i = 0; i < n; ++i) {
        const float x = __half2float(a[i]), y = __half2float(b[i]);
        const float d = fabsf(x - y);
        if (!(d <= me)) me = d;                 // NaN-propagating max
        if (fabsf(y) > mr) mr = fabsf(y);
    }
    *max_err = me; *max_ref = mr;
    return true;
}

// random test weight W[N][K] in the GEMM-ready device form (q4: random raw blocks in the FILE layout, repacked by
// the production upload kernel)
struct TestWeight {
    DevBuf wbuf, nib, qd, qm;
    DevWeight w;
    int build(int N, int K, int wtype, const float * d_bias, std::mt19937 & rng, bool ln_target = false) {
        std::normal_distribution<float> nd(0.f, 1.f);
        w.N = N; w.K = K; w.type = wtype; w.bias = d_bias; w.bn = ln_target ? gemm_pick_bn_ln(N) : gemm_pick_bn(N);
        const float wscale = 1.0f / sqrtf((float) K);
        if (wtype == W_F16) {
            std::vector<uint16_t> hw((size_t) N * K);
            for (auto & v : hw) v = f2h_bits(wscale * nd(rng));
            if (!wbuf.alloc(hw.size() * 2)) return -3;
            cudaMemcpy(wbuf.p, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice);
            w.w16 = wbuf.as<__half>();
            if (!make_tmap_f16_2d(&w.tmap, w.w16, N, K, w.bn / 2)) return -4;
            return 0;
        }
        const int bs = wtype == W_Q4_0 ? 18 : 20;
        const size_t nblk = (size_t) N * (K / 32);
        std::vector<uint8_t> raw(nblk * bs);
        for (size_t b = 0; b < nblk; ++b) {
            uint8_t * p = raw.data() + b * bs;
            const uint16_t d = f2h_bits(wscale * (0.05f + fabsf(nd(rng))) / 4.f);
            memcpy(p, &d, 2);
            int off = 2;
            if (wtype == W_Q4_1) { const uint16_t m = f2h_bits(-2.f * wscale * fabsf(nd(rng))); memcpy(p + 2, &m, 2); off = 4; }
            for (int j = 0; j < 16; ++j) p[off + j] = (uint8_t) (rng() & 0xFF);
        }
        if (!wbuf.alloc(raw.size()) || !nib.alloc(nblk * 16) || !qd.alloc(nblk * 2) || !qm.alloc(nblk * 2)) return -3;
        cudaMemcpy(wbuf.p, raw.data(), raw.size(), cudaMemcpyHostToDevice);
        if (launch_repack_q4(wbuf.as<uint8_t>(), wtype, N, K, 0, N, nib.as<uint4>(), qd.as<__half>(), qm.as<__half>(), 0) != cudaSuccess) return -5;
        w.qnib = nib.as<uint4>(); w.qd = qd.as<__half>(); w.qm = wtype == W_Q4_1 ? qm.as<__half>() : nullptr;
        memset(&w.tmap, 0, sizeof(w.tmap));
        return 0;
    }
};

// device-side N(0, scale^2) fp16 fill (hash of the element index + Box-Muller): large test operands without host RNG time
__global__ void fill_normal_h_kernel(__half * p, size_t n, uint32_t seed, float scale) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        uint64_t z = (uint64_t) i * 0x9E3779B97F4A7C15ull + ((uint64_t) seed << 32 | 0x632BE5ABu);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
        const float u1 = ((uint32_t) z + 1.0f) * (1.0f / 4294967808.0f), u2 = (float) (uint32_t) (z >> 32) * (1.0f / 4294967296.0f);
        p[i] = __float2half_rn(scale * sqrtf(-2.0f * __logf(u1)) * __cosf(6.2831853f * u2));
    }
}
}  // namespace

// reps == 0: run the tcgen05 kernel and the SIMT kernel once and compare;
// reps  > 0: time `reps` back-to-back launches of the tcgen05 kernel (CUDA events, after one warm-up)
static int32_t gemm_test_impl(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue, int32_t seed,
                              float * max_abs_err, float * max_ref, int reps, float * ms_out) {
    if (M <= 0 || N % 64 || K % 64 || wtype < 1 || wtype > 3) return -1;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fprintf(stderr, "bertx_selftest_gemm: no CUDA device\n"); return -2; }
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    std::mt19937 rng((uint32_t) seed);
    std::normal_distribution<float> nd(0.f, 1.f);
    const size_t rows = ((size_t) M + 127) / 128 * 128;
    std::vector<uint16_t> ha(rows * K, 0), hres((size_t) M * N);
    for (size_t i = 0; i < (size_t) M * K; ++i) ha[i] = f2h_bits(nd(rng));
    for (auto & v : hres) v = f2h_bits(nd(rng));
    std::vector<float> hbias(N);
    for (auto & v : hbias) v = 0.5f * nd(rng);
    DevBuf a, res, bias, o1, o2;
    if (!a.alloc(ha.size() * 2) || !res.alloc(hres.size() * 2) || !bias.alloc((size_t) N * 4) ||
        !o1.alloc((size_t) M * N * 2) || !o2.alloc((size_t) M * N * 2)) return -3;
    cudaMemcpy(a.p, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(res.p, hres.data(), hres.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(bias.p, hbias.data(), (size_t) N * 4, cudaMemcpyHostToDevice);
    cudaMemset(o1.p, 0xFF, (size_t) M * N * 2);
    TestWeight tw_;
    { const int rc = tw_.build(N, K, wtype, bias.as<float>(), rng, (epilogue & EPI_LN) != 0); if (rc) return rc; }
    DevWeight & w = tw_.w;
    CUtensorMap tma, tmo, tmr;
    if (!make_tmap_f16_2d(&tma, a.p, rows, K, 128)) return -4;
    if (!make_tmap_f16_2d(&tmr, res.p, M, N, 32)) return -4;
    if (!make_tmap_f16_2d(&tmo, o1.p, M, N, 32)) return -4;        // exactly M rows: the store must clip the tail tile
    const int fq = wtype == W_Q4_1 ? W_Q4_1 : W_Q4_0;
    GemmArgs g1{&tma, &tmo, &tmr, a.as<__half>(), &w, res.as<__half>(), o1.as<__half>(), M, epilogue, fq};
    DevBuf o1q, lng, lnb;
    CUtensorMap tmo2;
    if (epilogue & EPI_LN) {
        std::vector<float> hg(N), hb(N);
        for (auto & v : hg) v = 1.0f + 0.2f * nd(rng);
        for (auto & v : hb) v = 0.3f * nd(rng);
        if (!o1q.alloc((size_t) M * N * 2) || !lng.alloc((size_t) N * 4) || !lnb.alloc((size_t) N * 4)) return -3;
        cudaMemcpy(lng.p, hg.data(), (size_t) N * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(lnb.p, hb.data(), (size_t) N * 4, cudaMemcpyHostToDevice);
        cudaMemset(o1q.p, 0xFF, (size_t) M * N * 2);
        if (!make_tmap_f16_2d(&tmo2, o1q.p, M, N, 32)) return -4;
        g1.tmap_out2 = &tmo2; g1.ln_gamma = lng.as<float>(); g1.ln_beta = lnb.as<float>();
    }
    GemmArgs g2 = g1; g2.out = o2.as<__half>();
    if (!cuda_ok(launch_gemm_tc(g1, prop.multiProcessorCount, 0), "selftest gemm_tc launch")) return -6;
    if (!cuda_ok(cudaDeviceSynchronize(), "selftest gemm_tc")) return -6;
    if (reps > 0) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0, 0);
        for (int i = 0; i < reps; ++i) launch_gemm_tc(g1, prop.multiProcessorCount, 0);
        cudaEventRecord(e1, 0);
        const bool ok = cuda_ok(cudaDeviceSynchronize(), "bench gemm_tc");
        float ms = 0.f;
        if (ok) cudaEventElapsedTime(&ms, e0, e1);
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        *ms_out = ms;
        return ok ? 0 : -6;
    }
    if (epilogue & EPI_LN) {
        // reference for the fused LayerNorm epilogue: SIMT GEMM (+bias +residual) -> y, then the row kernel
        DevBuf y2, o2q;
        if (!y2.alloc((size_t) M * N * 2) || !o2q.alloc((size_t) M * N * 2)) return -3;
        g2.out = y2.as<__half>(); g2.epilogue = EPI_BIAS | EPI_RESIDUAL;
        if (!cuda_ok(launch_gemm_simt(g2, 0), "selftest gemm_simt launch")) return -7;
        if (!cuda_ok(launch_ln(y2.as<__half>(), lng.as<float>(), lnb.as<float>(), M, N, o2.as<__half>(),
                               (epilogue & EPI_FAKEQUANT) ? o2q.as<__half>() : nullptr, fq, 0), "selftest ln launch")) return -7;
        if (!cuda_ok(cudaDeviceSynchronize(), "selftest gemm_simt + ln")) return -7;
        float e1 = 0.f, r1 = 0.f, e2 = 0.f, r2 = 0.f;
        if (!compare_h(o1.as<__half>(), o2.as<__half>(), (size_t) M * N, &e1, &r1)) return -8;
        if ((epilogue & EPI_FAKEQUANT) && !compare_h(o1q.as<__half>(), o2q.as<__half>(), (size_t) M * N, &e2, &r2)) return -8;
        *max_abs_err = e1 > e2 ? e1 : e2; *max_ref = r1 > r2 ? r1 : r2;
        return 0;
    }
    if (!cuda_ok(launch_gemm_simt(g2, 0), "selftest gemm_simt launch")) return -7;
    if (!cuda_ok(cudaDeviceSynchronize(), "selftest gemm_simt")) return -7;
    return compare_h(o1.as<__half>(), o2.as<__half>(), (size_t) M * N, max_abs_err, max_ref) ? 0 : -8;
}

extern "C" {

int32_t bertx_selftest_gemm(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue, int32_t seed,
                            float * max_abs_err, float * max_ref) {
    return gemm_test_impl(M, N, K, wtype, epilogue, seed, max_abs_err, max_ref, 0, nullptr);
}

int32_t bertx_bench_gemm(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue, int32_t reps, float * ms_total) {
    float a = 0.f, b = 0.f;
    if (reps <= 0) return -1;
    return gemm_test_impl(M, N, K, wtype, epilogue, 1, &a, &b, reps, ms_total);
}

// Attention kernel the engine picks for this head size (engine_run.cu): 0 = attention_tc.cu (head size 64), 1 = attention_tc2.cu,
// 3 = attention_tc3.cu (head size 64); BERT_ATTN_IMPL = tc2 | tc3 overrides as it does for the engine
static int attn_impl_for(int d_head) {
    const char * e = getenv("BERT_ATTN_IMPL");
    if (d_head != 64 || (e && strcmp(e, "tc2") == 0)) return 1;
    if (e && strcmp(e, "tc3") == 0) return 3;
    return 0;
}

int32_t bertx_selftest_attention(int32_t n_sent, int32_t max_len, int32_t n_head, int32_t d_head, int32_t seed,
                                 float * max_abs_err, float * max_ref) {
    if (n_sent <= 0 || max_len <= 0 || (d_head != 32 && d_head != 64)) return -1;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fprintf(stderr, "bertx_selftest_attention: no CUDA device\n"); return -2; }
    std::mt19937 rng((uint32_t) seed);
    std::normal_distribution<float> nd(0.f, 1.f);
    const int H = n_head * d_head;
    std::vector<int32_t> cu(n_sent + 1, 0);
    for (int i = 0; i < n_sent; ++i) {
        const int len = i == 0 ? max_len : 1 + (int) (rng() % (uint32_t) max_len);   // ragged, first is the longest
        cu[i + 1] = cu[i] + len;
    }
    const size_t T = (size_t) cu[n_sent];
    std::vector<uint16_t> hq(T * 3 * H);
    for (auto & v : hq) v = f2h_bits(1.5f * nd(rng));
    if (seed & 1024) {                                             // ramp: K rows grow with the position (forces max rescales)
        for (int i = 0; i < n_sent; ++i) {
            for (int t = cu[i]; t < cu[i + 1]; ++t) {
                const float sc = 0.25f + 3.0f * (float) (t - cu[i]) / (float) (cu[i + 1] - cu[i]);
                for (int c = H; c < 2 * H; ++c) hq[(size_t) t * 3 * H + c] = f2h_bits(sc * 1.5f * nd(rng));
            }
        }
    }
    DevBuf qkv, dcu, o1, o2, dspan, dtb;
    if (!qkv.alloc(hq.size() * 2) || !dcu.alloc(cu.size() * 4) || !o1.alloc(T * H * 2) || !o2.alloc(T * H * 2)) return -3;
    cudaMemcpy(qkv.p, hq.data(), hq.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dcu.p, cu.data(), cu.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(o1.p, 0xFF, T * H * 2);
    {
        // the engine's own dispatch (attn_impl_for) unless a seed bit names a kernel: bit 11 attention_tc.cu, bit 12 attention_tc2.cu
        // (eight softmax warps), bit 13 attention_tc3.cu (the round-2 study kernel); head size 32 always runs attention_tc2.cu
        CUtensorMap tq, tkv;
        cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
        if (!make_attention_tmaps(&tq, &tkv, qkv.as<__half>(), T, H)) return -4;
        const int impl = d_head != 64 ? 1 : (seed & 2048) ? 0 : (seed & 4096) ? 1 : (seed & 8192) ? 3 : attn_impl_for(d_head);
        cudaError_t ae;
        if (impl == 3) ae = launch_attention_tc3(&tq, &tkv, dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0);
        else if (!(seed & 16384)) {
            // as the engine runs it: short sentences packed into spans (bit 14 of the seed: one work item per sentence)
            std::vector<int32_t> span(n_sent + 1);
            const int ns = pack_spans(cu.data(), n_sent, span.data());
            std::vector<int2> tb(T);
            for (int i = 0; i < n_sent; ++i) for (int t = cu[i]; t < cu[i + 1]; ++t) tb[t] = make_int2(cu[i], cu[i + 1]);
            if (!dspan.alloc(span.size() * 4) || !dtb.alloc(tb.size() * 8)) return -3;
            cudaMemcpy(dspan.p, span.data(), (size_t) (ns + 1) * 4, cudaMemcpyHostToDevice);
            cudaMemcpy(dtb.p, tb.data(), tb.size() * 8, cudaMemcpyHostToDevice);
            ae = impl == 1 ? launch_attention_tc2(&tq, &tkv, dspan.as<int32_t>(), ns, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0, dtb.as<int2>())
                           : launch_attention_tc(&tq, &tkv, dspan.as<int32_t>(), ns, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0, dtb.as<int2>());
        }
        else if (impl == 1) ae = launch_attention_tc2(&tq, &tkv, dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0);
        else ae = launch_attention_tc(&tq, &tkv, dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0);
        if (!cuda_ok(ae, "selftest attention launch")) return -6;
    }
    if (!cuda_ok(cudaDeviceSynchronize(), "selftest attention")) return -6;
    if (!cuda_ok(launch_attention_ref(qkv.as<__half>(), dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o2.as<__half>(), -1, 0), "selftest attention_ref launch")) return -7;
    if (!cuda_ok(cudaDeviceSynchronize(), "selftest attention_ref")) return -7;
    return compare_h(o1.as<__half>(), o2.as<__half>(), T * H, max_abs_err, max_ref) ? 0 : -8;
}

// Debug aid for a per-kernel check against numpy (tests/test_gpu.py::test_attention_against_numpy): the production attention
// kernel on random fp16 q/k/v for sentences of the given lengths; the inputs ([T][3H] fp16 bits) and the kernel's output
// ([T][H] fp16 bits) are copied to the caller's arrays.  flags: the seed bits of bertx_selftest_attention (kernel choice, bit 14
// = no span packing).
int32_t bertx_debug_attention(int32_t n_sent, const int32_t * lens, int32_t n_head, int32_t d_head, int32_t seed, int32_t flags,
                              uint16_t * qkv_out, uint16_t * out) {
    if (n_sent <= 0 || !lens || !qkv_out || !out || (d_head != 32 && d_head != 64)) return -1;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fprintf(stderr, "bertx_debug_attention: no CUDA device\n"); return -2; }
    std::mt19937 rng((uint32_t) seed);
    std::normal_distribution<float> nd(0.f, 1.f);
    const int H = n_head * d_head;
    std::vector<int32_t> cu(n_sent + 1, 0);
    int max_len = 0;
    for (int i = 0; i < n_sent; ++i) {
        if (lens[i] <= 0) return -1;
        cu[i + 1] = cu[i] + lens[i];
        max_len = std::max(max_len, (int) lens[i]);
    }
    const size_t T = (size_t) cu[n_sent];
    std::vector<uint16_t> hq(T * 3 * H);
    for (auto & v : hq) v = f2h_bits(1.5f * nd(rng));
    DevBuf qkv, dcu, o1, dspan, dtb;
    if (!qkv.alloc(hq.size() * 2) || !dcu.alloc(cu.size() * 4) || !o1.alloc(T * H * 2)) return -3;
    cudaMemcpy(qkv.p, hq.data(), hq.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dcu.p, cu.data(), cu.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(o1.p, 0xFF, T * H * 2);
    CUtensorMap tq, tkv;
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    if (!make_attention_tmaps(&tq, &tkv, qkv.as<__half>(), T, H)) return -4;
    const int impl = d_head != 64 ? 1 : (flags & 2048) ? 0 : (flags & 4096) ? 1 : (flags & 8192) ? 3 : attn_impl_for(d_head);
    cudaError_t ae;
    if (impl == 3) {
        ae = launch_attention_tc3(&tq, &tkv, dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0);
    } else if (!(flags & 16384)) {
        std::vector<int32_t> span(n_sent + 1);
        const int ns = pack_spans(cu.data(), n_sent, span.data());
        std::vector<int2> tb(T);
        for (int i = 0; i < n_sent; ++i) for (int t = cu[i]; t < cu[i + 1]; ++t) tb[t] = make_int2(cu[i], cu[i + 1]);
        if (!dspan.alloc(span.size() * 4) || !dtb.alloc(tb.size() * 8)) return -3;
        cudaMemcpy(dspan.p, span.data(), (size_t) (ns + 1) * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dtb.p, tb.data(), tb.size() * 8, cudaMemcpyHostToDevice);
        ae = impl == 1 ? launch_attention_tc2(&tq, &tkv, dspan.as<int32_t>(), ns, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0, dtb.as<int2>())
                       : launch_attention_tc(&tq, &tkv, dspan.as<int32_t>(), ns, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0, dtb.as<int2>());
    } else {
        ae = impl == 1 ? launch_attention_tc2(&tq, &tkv, dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0)
                       : launch_attention_tc(&tq, &tkv, dcu.as<int32_t>(), n_sent, max_len, n_head, d_head, o1.as<__half>(), -1, prop.multiProcessorCount, 0);
    }
    if (!cuda_ok(ae, "debug attention launch") || !cuda_ok(cudaDeviceSynchronize(), "debug attention")) return -6;
    memcpy(qkv_out, hq.data(), hq.size() * 2);
    if (!cuda_ok(cudaMemcpy(out, o1.p, T * H * 2, cudaMemcpyDeviceToHost), "debug attention copy")) return -6;
    return 0;
}

int32_t bertx_bench_attention(int32_t n_sent, int32_t len, int32_t n_head, int32_t d_head, int32_t reps, float * ms_total) {
    if (n_sent <= 0 || len <= 0 || reps <= 0 || (d_head != 32 && d_head != 64)) return -1;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fprintf(stderr, "bertx_bench_attention: no CUDA device\n"); return -2; }
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int H = n_head * d_head;
    const size_t T = (size_t) n_sent * len;
    std::vector<int32_t> cu(n_sent + 1);
    for (int i = 0; i <= n_sent; ++i) cu[i] = i * len;
    // 1 MiB of random fp16 tiled over the buffer is enough for timing
    std::mt19937 rng(7);
    std::normal_distribution<float> nd(0.f, 1.f);
    std::vector<uint16_t> chunk(1 << 19);
    for (auto & v : chunk) v = f2h_bits(nd(rng));
    DevBuf qkv, dcu, o1;
    const size_t nbytes = T * 3 * H * 2;
    if (!qkv.alloc(nbytes) || !dcu.alloc(cu.size() * 4) || !o1.alloc(T * H * 2)) return -3;
    for (size_t off = 0; off < nbytes; off += chunk.size() * 2) {
        const size_t n = std::min(chunk.size() * 2, nbytes - off);
        cudaMemcpy((char *) qkv.p + off, chunk.data(), n, cudaMemcpyHostToDevice);
    }
    cudaMemcpy(dcu.p, cu.data(), cu.size() * 4, cudaMemcpyHostToDevice);
    CUtensorMap tq, tkv;
    if (!make_attention_tmaps(&tq, &tkv, qkv.as<__half>(), T, H)) return -4;
    const int impl = attn_impl_for(d_head);
    DevBuf dspan, dtb;
    int ns = 0;
    if (impl != 3 && !getenv("BERT_ATTN_NOPACK")) {               // as the engine runs it: short sentences packed into spans
        std::vector<int32_t> span(n_sent + 1);
        ns = pack_spans(cu.data(), n_sent, span.data());
        std::vector<int2> tb(T);
        for (int i = 0; i < n_sent; ++i) for (int t = cu[i]; t < cu[i + 1]; ++t) tb[t] = make_int2(cu[i], cu[i + 1]);
        if (!dspan.alloc(span.size() * 4) || !dtb.alloc(tb.size() * 8)) return -3;
        cudaMemcpy(dspan.p, span.data(), (size_t) (ns + 1) * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dtb.p, tb.data(), tb.size() * 8, cudaMemcpyHostToDevice);
    }
    auto launch = [&]() {
        if (impl == 1 && ns > 0) return launch_attention_tc2(&tq, &tkv, dspan.as<int32_t>(), ns, len, n_head, d_head, o1.as<__half>(), W_Q4_0, prop.multiProcessorCount, 0, dtb.as<int2>());
        if (impl == 0 && ns > 0) return launch_attention_tc(&tq, &tkv, dspan.as<int32_t>(), ns, len, n_head, d_head, o1.as<__half>(), W_Q4_0, prop.multiProcessorCount, 0, dtb.as<int2>());
        if (impl == 3) return launch_attention_tc3(&tq, &tkv, dcu.as<int32_t>(), n_sent, len, n_head, d_head, o1.as<__half>(), W_Q4_0, prop.multiProcessorCount, 0);
        if (impl == 1) return launch_attention_tc2(&tq, &tkv, dcu.as<int32_t>(), n_sent, len, n_head, d_head, o1.as<__half>(), W_Q4_0, prop.multiProcessorCount, 0);
        return launch_attention_tc(&tq, &tkv, dcu.as<int32_t>(), n_sent, len, n_head, d_head, o1.as<__half>(), W_Q4_0, prop.multiProcessorCount, 0);
    };
    if (!cuda_ok(launch(), "bench attention launch") || !cuda_ok(cudaDeviceSynchronize(), "bench attention")) return -6;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, 0);
    for (int i = 0; i < reps; ++i) launch();
    cudaEventRecord(e1, 0);
    const bool ok = cuda_ok(cudaDeviceSynchronize(), "bench attention");
    float ms = 0.f;
    if (ok) cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *ms_total = ms;
    if (ok && impl == 3 && getenv("BERT_ATTN_PROF")) {            // wait clocks of CTA 0 (only a -DBERT_ATTN3_PROF=1 build counts)
        unsigned long long c[32];
        attn3_prof_fetch(c);
        const double n = (double) (reps + 1);
        fprintf(stderr, "attn3 prof (clocks per launch, CTA 0): MMA warp total %.0f = q_full %.0f + k_full %.0f + p_ready %.0f + o_free %.0f + v_full %.0f + rest; blocks %.0f items %.0f\n"
                        "  softmax warp total %.0f: s_full(next) %.0f s_full(first) %.0f pv_done(final) %.0f finalize %.0f\n"
                        "  producer total %.0f: q_empty %.0f k_empty %.0f v_empty %.0f\n",
                c[5] / n, c[0] / n, c[1] / n, c[2] / n, c[3] / n, c[4] / n, c[6] / n, c[7] / n,
                c[11] / n, c[8] / n, c[9] / n, c[10] / n, c[12] / n, c[19] / n, c[16] / n, c[17] / n, c[18] / n);
    }
    return ok ? 0 : -6;
}

// Debug (profiles/r2_gemm_race.md): one GEMM at ANY size -- device-generated A (and residual), the tcgen05 kernel run `runs`
// times, the SIMT validation kernel (+ the stand-alone LayerNorm kernel for EPI_LN) once; all outputs are copied to the
// caller's host arrays ([M][N] fp16 bit patterns) for element-wise analysis (tools/gemm_race.py).  out_ref == NULL skips the
// validation kernel; out_ref2 / out_tc2 receive the q8 twin of a fused-LayerNorm epilogue (NULL: not wanted).
int32_t bertx_debug_gemm(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue, int32_t seed, int32_t runs,
                         uint16_t * out_ref, uint16_t * out_tc, float * ms_tc, uint16_t * out_ref2, uint16_t * out_tc2,
                         uint16_t * a_out, float * w_out) {
    if (M <= 0 || N % 64 || K % 64 || wtype < 1 || wtype > 3 || runs <= 0 || !out_tc) return -1;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { fprintf(stderr, "bertx_debug_gemm: no CUDA device\n"); return -2; }
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    std::mt19937 rng((uint32_t) seed);
    std::normal_distribution<float> nd(0.f, 1.f);
    const size_t rows = ((size_t) M + 127) / 128 * 128;
    const size_t n_out = (size_t) M * N;
    const bool ln = (epilogue & EPI_LN) != 0, twin = ln && (epilogue & EPI_FAKEQUANT);
    DevBuf a, res, bias, o1, o2, o1q, o2q, y2, lng, lnb;
    if (!a.alloc(rows * K * 2) || !bias.alloc((size_t) N * 4) || !o1.alloc(n_out * 2) || !o2.alloc(n_out * 2)) return -3;
    if ((epilogue & EPI_RESIDUAL) && !res.alloc(n_out * 2)) return -3;
    cudaMemset(a.p, 0, rows * K * 2);
    fill_normal_h_kernel<<<2048, 256>>>(a.as<__half>(), (size_t) M * K, (uint32_t) seed, 1.0f);
    if (res.p) fill_normal_h_kernel<<<2048, 256>>>(res.as<__half>(), n_out, (uint32_t) seed + 77u, 1.0f);
    std::vector<float> hbias(N);
    for (auto & v : hbias) v = 0.5f * nd(rng);
    cudaMemcpy(bias.p, hbias.data(), (size_t) N * 4, cudaMemcpyHostToDevice);
    TestWeight tw;
    { const int rc = tw.build(N, K, wtype, bias.as<float>(), rng, (epilogue & EPI_LN) != 0); if (rc) return rc; }
    CUtensorMap tma, tmo, tmr, tmo2;
    if (!make_tmap_f16_2d(&tma, a.p, rows, K, 128)) return -4;
    if (!make_tmap_f16_2d(&tmo, o1.p, M, N, 32)) return -4;
    if (res.p && !make_tmap_f16_2d(&tmr, res.p, M, N, 32)) return -4;
    const int fq = wtype == W_Q4_1 ? W_Q4_1 : W_Q4_0;
    GemmArgs g1{&tma, &tmo, res.p ? &tmr : nullptr, a.as<__half>(), &tw.w, res.as<__half>(), o1.as<__half>(), M, epilogue, fq};
    if (ln) {
        std::vector<float> hg(N), hb(N);
        for (auto & v : hg) v = 1.0f + 0.2f * nd(rng);
        for (auto & v : hb) v = 0.3f * nd(rng);
        if (!o1q.alloc(n_out * 2) || !o2q.alloc(n_out * 2) || !y2.alloc(n_out * 2) || !lng.alloc((size_t) N * 4) || !lnb.alloc((size_t) N * 4)) return -3;
        cudaMemcpy(lng.p, hg.data(), (size_t) N * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(lnb.p, hb.data(), (size_t) N * 4, cudaMemcpyHostToDevice);
        if (!make_tmap_f16_2d(&tmo2, o1q.p, M, N, 32)) return -4;
        g1.tmap_out2 = &tmo2; g1.ln_gamma = lng.as<float>(); g1.ln_beta = lnb.as<float>();
    }
    if (a_out && cudaMemcpy(a_out, a.p, (size_t) M * K * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return -8;
    if (w_out) {                                     // the weights as the kernels decode them, f32 [N][K]
        DevBuf wf;
        if (!wf.alloc((size_t) N * K * 4)) return -3;
        if (wtype == W_F16) {
            std::vector<__half> hw((size_t) N * K);
            cudaMemcpy(hw.data(), tw.w.w16, hw.size() * 2, cudaMemcpyDeviceToHost);
            for (size_t i = 0; i < hw.size(); ++i) w_out[i] = __half2float(hw[i]);
        } else {
            if (!cuda_ok(launch_dequant_q4_f32(tw.w, wf.as<float>(), 0), "debug dequant") || !cuda_ok(cudaDeviceSynchronize(), "debug dequant")) return -7;
            if (cudaMemcpy(w_out, wf.p, (size_t) N * K * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -8;
        }
    }
    if (out_ref) {
        GemmArgs g2 = g1;
        if (ln) {
            g2.out = y2.as<__half>(); g2.epilogue = EPI_BIAS | EPI_RESIDUAL;
            if (!cuda_ok(launch_gemm_simt(g2, 0), "debug gemm_simt launch")) return -7;
            if (!cuda_ok(launch_ln(y2.as<__half>(), lng.as<float>(), lnb.as<float>(), M, N, o2.as<__half>(), twin ? o2q.as<__half>() : nullptr, fq, 0), "debug ln launch")) return -7;
        } else {
            g2.out = o2.as<__half>();
            if (!cuda_ok(launch_gemm_simt(g2, 0), "debug gemm_simt launch")) return -7;
        }
        if (!cuda_ok(cudaDeviceSynchronize(), "debug gemm_simt")) return -7;
        if (cudaMemcpy(out_ref, o2.p, n_out * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return -8;
        if (twin && out_ref2 && cudaMemcpy(out_ref2, o2q.p, n_out * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return -8;
    }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int r = 0; r < runs; ++r) {
        cudaMemset(o1.p, 0xFF, n_out * 2);
        if (twin) cudaMemset(o1q.p, 0xFF, n_out * 2);
        cudaEventRecord(e0, 0);
        if (!cuda_ok(launch_gemm_tc(g1, prop.multiProcessorCount, 0), "debug gemm_tc launch")) return -6;
        cudaEventRecord(e1, 0);
        if (!cuda_ok(cudaDeviceSynchronize(), "debug gemm_tc")) return -6;
        if (ms_tc) cudaEventElapsedTime(&ms_tc[r], e0, e1);
        if (cudaMemcpy(out_tc + (size_t) r * n_out, o1.p, n_out * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return -8;
        if (twin && out_tc2 && cudaMemcpy(out_tc2 + (size_t) r * n_out, o1q.p, n_out * 2, cudaMemcpyDeviceToHost) != cudaSuccess) return -8;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

}  // extern "C"
