// K2 (tcgen05), head size 64, round-2 kernel -- variable-length fused attention
//   O = softmax(Q K^T / sqrt(d)) V   per (sentence, head), over the sentence's own tokens only.
// Replaces the reference's mask build (bert.cpp:957-961), scores + scale + mask + soft_max
// (bert.cpp:1018-1025, ggml.c:12008-12077) and P*V + permute/cont/cpy (bert.cpp:1027-1036).
// Nothing of size L x L reaches HBM; padded queries / keys are never stored / always masked.
//
// Same work split as attention_tc.cu (item = (sentence, head, 128-query tile); persistent CTAs, two per SM; warp 0 TMA
// producer, warp 1 MMA issuer, warps 2-5 softmax with one query row per thread, P handed back through TMEM as the A
// operand of P V).  What changed, and why (profiles/r2_attention.md): the round-1 softmax warps ran their phases one
// after the other per key block (tcgen05.ld -> row max -> exponentials -> tcgen05.st) and the two warps a scheduler
// holds fell into lock-step, so the MUFU idled while both loaded and reduced and the ALU idled while both
// exponentiated (tensor pipe 36 %, MUFU 49 %, issue 50 %).  Overlapping the phases needs the NEXT block's scores early,
// and with two S buffers S_{b+1} could only be issued after block b-1 had been handed back.  So:
//   * THREE score buffers in TMEM (3 x 64 columns + 64 for O = the CTA's 256 columns; the row sums move back into
//     registers): issue order S_0 S_1 S_2 PV_0 S_3 PV_1 ..., S_{b+1} only depends on block b-2;
//   * K and V travel through separate 4-stage rings: a K slot is free as soon as its S MMA has run, so the K loads stay
//     three blocks ahead while a V slot lives until its P V;
//   * the softmax warps form their own warpgroup and take 208 registers (setmaxnreg), the control warps keep 40;
//   * software-pipelined softmax: a block is handled as two 32-key halves, the next half's tcgen05.ld is in flight
//     while the current half is exponentiated, and from the second block on the exponentials are taken against the
//     CURRENT reference max, so the block's row max (needed only for the rare rescale decision) is formed alongside
//     them instead of in front of them.  If a row leaves the 2^8 window the accumulator is rescaled and the block is
//     redone from the S tile, which is still in TMEM.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int AQ = 128;            // queries per item
constexpr int AK = 64;             // keys per block
constexpr int AD = 64;             // head size
constexpr int K_STAGES = 4, V_STAGES = 4;
constexpr int S_BUFS = 3;
#ifndef BERT_ATTN3_VLAG
#define BERT_ATTN3_VLAG 2
#endif
constexpr int V_LAG = BERT_ATTN3_VLAG;         // the producer issues the load of V_b after that of K_{b + V_LAG}
constexpr int T3_THREADS = 256;           // warpgroup 0: control (warp 0 TMA, warp 1 MMA, 2-3 idle), warpgroup 1: softmax
constexpr int kQBytes = AQ * AD * 2;           // 16 KB
constexpr int kKVBytes = AK * AD * 2;          // 8 KB per K or V block
constexpr int kOutStage = 4 * 4096;            // per softmax warp: [32 rows][128 B] staging for coalesced output rows
constexpr int kT3Smem = kQBytes + (K_STAGES + V_STAGES) * kKVBytes + kOutStage + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t TM_S = 0;       // S buffers at columns 0, 64, 128 (P aliases the first 32 columns of each)
constexpr uint32_t TM_O = 192;     // O accumulator: 64 columns
constexpr uint32_t TM_COLS = 256;
constexpr float kRescaleThreshold = 8.f;      // log2 units
#ifndef BERT_ATTN3_POLY_EVERY
#define BERT_ATTN3_POLY_EVERY 4
#endif
constexpr int kPolyEvery = BERT_ATTN3_POLY_EVERY;
// diagnostic builds (profiles/r2_attention.md): BERT_ATTN3_SPIN=1 polls the barriers without a suspend hint;
// BERT_ATTN3_ABLATE bit 0 drops the softmax arithmetic (P = raw score bits), bit 1 the softmax warps' TMEM loads / stores,
// bit 2 the K / V loads (the barriers are just arrived on), bit 3 the MMAs (commits only)
#ifndef BERT_ATTN3_SPIN
#define BERT_ATTN3_SPIN 0
#endif
#ifndef BERT_ATTN3_ABLATE
#define BERT_ATTN3_ABLATE 0
#endif
// BERT_ATTN3_PROF=1 (diagnostic build): CTA 0 accumulates the clocks its MMA warp and its first softmax warp spend in each
// barrier wait into g_attn3_prof (read and reset by attn3_prof_fetch)
#ifndef BERT_ATTN3_PROF
#define BERT_ATTN3_PROF 0
#endif
__device__ unsigned long long g_attn3_prof[32];
#if BERT_ATTN3_PROF
#define PROF_WAIT(slot, stmt) do { const unsigned t0_ = (unsigned) clock64(); stmt; prof[slot] += (unsigned) clock64() - t0_; } while (0)
#else
#define PROF_WAIT(slot, stmt) do { stmt; } while (0)
#endif
#if BERT_ATTN3_ABLATE & 2
__device__ __forceinline__ void ld_s3(uint32_t taddr, uint32_t (&r)[32]) {
#pragma unroll
    for (int j = 0; j < 32; ++j) asm volatile("mov.b32 %0, %1;" : "=r"(r[j]) : "r"(0x3f000000u + ((taddr * 2654435761u + j * 40503u) & 0xfffffu)));
}
__device__ __forceinline__ void st_p3(uint32_t, const uint32_t (&r)[32]) {
#pragma unroll
    for (int j = 0; j < 32; ++j) asm volatile("" :: "r"(r[j]));
}
#else
__device__ __forceinline__ void ld_s3(uint32_t taddr, uint32_t (&r)[32]) { tmem_ld_32x32b_x32(taddr, r); }
__device__ __forceinline__ void st_p3(uint32_t taddr, const uint32_t (&r)[32]) { tmem_st_32x32b_x32(taddr, r); }
#endif
#if BERT_ATTN3_SPIN
__device__ __forceinline__ void mbar_wait3(uint64_t * bar, uint32_t parity) {
    uint32_t ok = 0, spins = 0;
    while (true) {
#if BERT_ATTN3_SPIN == 2
        asm volatile("{\n\t.reg .pred P;\n\tmbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
#else
        asm volatile("{\n\t.reg .pred P;\n\tmbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
#endif
        if (ok) break;
        if (++spins > (1u << 26)) { asm volatile("trap;"); }
    }
}
#else
__device__ __forceinline__ void mbar_wait3(uint64_t * bar, uint32_t parity) { mbar_wait(bar, parity); }
#endif   // every kPolyEvery-th pair of a half takes the polynomial path (0: all on the MUFU)

struct Attn3Params {
    const int32_t * cu;
    __half * out;
    int n_sent, n_head, H, max_qt, fq_type;
    float scale_log2;
};

struct Item3 {
    int t0, t1, q0, head;
    __device__ __forceinline__ int len() const { return t1 - t0; }
    __device__ __forceinline__ bool valid() const { return q0 < t1 - t0; }
};

// Round-robin item walk (query tile fastest); the coordinates of the NEXT item are advanced incrementally and its sentence
// bounds -- two dependent-free global loads -- are fetched while the current item is processed.
struct ItemWalk3 {
    int next_item, total, stride;
    int qt, head, sent;
    int dq, dh, ds;
    Item3 cur, nxt;
    __device__ __forceinline__ void fetch(const Attn3Params & p) {
        if (next_item >= total) { nxt.t0 = nxt.t1 = nxt.q0 = nxt.head = 0; return; }
        nxt.head = head;
        nxt.q0 = qt * AQ;
        nxt.t0 = __ldg(p.cu + sent);
        nxt.t1 = __ldg(p.cu + sent + 1);
    }
    __device__ __forceinline__ ItemWalk3(const Attn3Params & p, int first, int stride_, int total_)
        : next_item(first), total(total_), stride(stride_) {
        qt = first % p.max_qt; const int hs = first / p.max_qt; head = hs % p.n_head; sent = hs / p.n_head;
        dq = stride % p.max_qt; const int sh = stride / p.max_qt; dh = sh % p.n_head; ds = sh / p.n_head;
        fetch(p);
    }
    __device__ __forceinline__ bool advance(const Attn3Params & p) {
        if (next_item >= total) return false;
        cur = nxt;
        next_item += stride;
        qt += dq; head += dh; sent += ds;
        if (qt >= p.max_qt) { qt -= p.max_qt; ++head; }
        if (head >= p.n_head) { head -= p.n_head; ++sent; }
        fetch(p);
        return true;
    }
};

__device__ __forceinline__ float ex2f3(float x) {
    float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
// 2^x for a pair on the FMA / ALU pipes instead of the MUFU (16 results/clk/SM): round-to-nearest split x = n + r with
// the 1.5 * 2^23 constant, degree-3 minimax polynomial for 2^r on [-0.5, 0.5] (max relative error 7.5e-5, well under the
// fp16 rounding of P), n added straight into the exponent field.  x <= ~8 here; everything below -125 (incl. the -inf of
// masked keys) is clamped and comes out as ~2^-125, which rounds to 0 in fp16.
__device__ __forceinline__ float2 ex2_poly3(float2 x) {
    const float magic = 12582912.f;
    x.x = fmaxf(x.x, -125.f); x.y = fmaxf(x.y, -125.f);
    const float2 t = __fadd2_rn(x, make_float2(magic, magic));
    const float2 n = __fadd2_rn(t, make_float2(-magic, -magic));
    const float2 r = __fadd2_rn(x, make_float2(-n.x, -n.y));
    float2 q = __ffma2_rn(r, make_float2(0.055171531f, 0.055171531f), make_float2(0.24261111f, 0.24261111f));
    q = __ffma2_rn(q, r, make_float2(0.69326103f, 0.69326103f));
    q = __ffma2_rn(q, r, make_float2(0.99992806f, 0.99992806f));
    return make_float2(__uint_as_float(__float_as_uint(q.x) + (__float_as_uint(t.x) << 23)),
                       __uint_as_float(__float_as_uint(q.y) + (__float_as_uint(t.y) << 23)));
}
__device__ __forceinline__ uint32_t pack_h23(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}
// tcgen05.wait::ld tied to the registers it makes valid, so that no use of them can be scheduled in front of it
__device__ __forceinline__ void tmem_ld_wait_regs(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                   "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}
// keys past the sentence end (kvalid = keys of this 32-key half inside the sentence) never win the max and weigh 0
__device__ __forceinline__ void mask_half(uint32_t (&r)[32], int kvalid) {
    if (kvalid < 32) {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (j >= kvalid) r[j] = 0xff800000u;
    }
}
__device__ __forceinline__ float max_half(const uint32_t (&r)[32]) {
#if BERT_ATTN3_ABLATE & 1
    return __uint_as_float(r[0]);
#endif
    float mx4[4] = {__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])};
#pragma unroll
    for (int j = 4; j < 32; j += 4) {
        mx4[0] = fmaxf(mx4[0], __uint_as_float(r[j])); mx4[1] = fmaxf(mx4[1], __uint_as_float(r[j + 1]));
        mx4[2] = fmaxf(mx4[2], __uint_as_float(r[j + 2])); mx4[3] = fmaxf(mx4[3], __uint_as_float(r[j + 3]));
    }
    return fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
}
// 32 raw scores -> 16 packed fp16 pairs of 2^(s * scale - m_ref); returns their sum
__device__ __forceinline__ float exp_half(const uint32_t (&r)[32], float scale_log2, float m_ref, uint32_t * pk) {
#if BERT_ATTN3_ABLATE & 1
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = r[2 * j] ^ r[2 * j + 1];
    return 1.f;
#endif
    const float2 nm2 = make_float2(-m_ref, -m_ref), sc2 = make_float2(scale_log2, scale_log2);
    float2 sum[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const float2 x = __ffma2_rn(make_float2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1])), sc2, nm2);
        const float2 e = (kPolyEvery > 0 && j % (kPolyEvery > 0 ? kPolyEvery : 1) == kPolyEvery - 1)
                             ? ex2_poly3(x) : make_float2(ex2f3(x.x), ex2f3(x.y));
        sum[j & 1] = __fadd2_rn(sum[j & 1], e);
        pk[j] = pack_h23(e.x, e.y);
    }
    return (sum[0].x + sum[0].y) + (sum[1].x + sum[1].y);
}

__global__ void __launch_bounds__(T3_THREADS, 2)
attention_tc3_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                     const Attn3Params p) {
    extern __shared__ uint8_t smem_raw[];
    // 1024-B alignment by an offset from the array (not an integer round trip), so that the compiler keeps the
    // shared address space and emits LDS/STS rather than generic LD/ST
    uint8_t * smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t * smem_q = smem;
    uint8_t * smem_k = smem + kQBytes;                               // K_STAGES x 8 KB
    uint8_t * smem_v = smem_k + K_STAGES * kKVBytes;                 // V_STAGES x 8 KB
    uint8_t * smem_o = smem_v + V_STAGES * kKVBytes;                 // 4 x 4 KB output staging
    uint64_t * bars = (uint64_t *) (smem_o + kOutStage);
    uint64_t * q_full = bars;                   // [1]
    uint64_t * q_empty = bars + 1;              // [1]
    uint64_t * k_full = bars + 2;               // [K_STAGES]
    uint64_t * k_empty = k_full + K_STAGES;     // [K_STAGES]
    uint64_t * v_full = k_empty + K_STAGES;     // [V_STAGES]
    uint64_t * v_empty = v_full + V_STAGES;     // [V_STAGES]
    uint64_t * s_full = v_empty + V_STAGES;     // [S_BUFS]
    uint64_t * p_ready = s_full + S_BUFS;       // [S_BUFS]
    uint64_t * pv_done = p_ready + S_BUFS;      // [S_BUFS]  P V of block g has reached the accumulator (barrier g % 3)
    uint64_t * o_free = pv_done + S_BUFS;       // [1]
    uint32_t * tmem_slot = (uint32_t *) (o_free + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total_items = p.n_sent * p.n_head * p.max_qt;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap_kv);
        mbar_init(q_full, 1); mbar_init(q_empty, 1);
        for (int s = 0; s < K_STAGES; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); }
        for (int s = 0; s < V_STAGES; ++s) { mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1); }
        for (int a = 0; a < S_BUFS; ++a) { mbar_init(&s_full[a], 1); mbar_init(&p_ready[a], 4); mbar_init(&pv_done[a], 1); }
        mbar_init(o_free, 4);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, TM_COLS); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();                                 // everything above overlapped the tail of the previous kernel

    // registers follow the work: the control warpgroup keeps 40 per thread, the softmax warpgroup takes 208 (it holds
    // three 32-score halves, the packed P and the item walk at once); 128 x 40 + 128 x 208 <= the CTA's 256 x 128
    if (warp < 4) {
      reg_dealloc<40>();
      if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            uint32_t kc = 0, vc = 0, item_n = 0;                        // blocks loaded into the K / V ring so far
            [[maybe_unused]] unsigned prof[20] = {}; [[maybe_unused]] const unsigned tstart = (unsigned) clock64();
            ItemWalk3 walk(p, blockIdx.x, gridDim.x, total_items);
            while (walk.advance(p)) {
                const Item3 it = walk.cur;
                if (!it.valid()) continue;
                const int nb = (it.len() + AK - 1) / AK;
                PROF_WAIT(16, mbar_wait3(q_empty, (item_n & 1) ^ 1));
                mbar_expect_tx(q_full, kQBytes);
                tma_load_2d(smem_q, &tmap_q, q_full, it.head * AD, it.t0 + it.q0);
                for (int b = 0; b < nb + V_LAG; ++b) {
                    if (b < nb) {
                        const uint32_t st = kc % K_STAGES;
                        PROF_WAIT(17, mbar_wait3(&k_empty[st], ((kc / K_STAGES) & 1) ^ 1));
#if BERT_ATTN3_ABLATE & 4
                        mbar_arrive(&k_full[st]);
#else
                        mbar_expect_tx(&k_full[st], kKVBytes);
                        tma_load_2d(smem_k + st * kKVBytes, &tmap_kv, &k_full[st], p.H + it.head * AD, it.t0 + b * AK);
#endif
                        ++kc;
                    }
                    if (b >= V_LAG) {
                        const uint32_t st = vc % V_STAGES;
                        PROF_WAIT(18, mbar_wait3(&v_empty[st], ((vc / V_STAGES) & 1) ^ 1));
#if BERT_ATTN3_ABLATE & 4
                        mbar_arrive(&v_full[st]);
#else
                        mbar_expect_tx(&v_full[st], kKVBytes);
                        tma_load_2d(smem_v + st * kKVBytes, &tmap_kv, &v_full[st], 2 * p.H + it.head * AD, it.t0 + (b - V_LAG) * AK);
#endif
                        ++vc;
                    }
                }
                ++item_n;
            }
#if BERT_ATTN3_PROF
            if (blockIdx.x == 0) { prof[19] = (unsigned) clock64() - tstart; for (int i = 16; i < 20; ++i) atomicAdd(&g_attn3_prof[i], (unsigned long long) prof[i]); }
#endif
        }
      } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        // per item: S_0 S_1 S_2 PV_0 S_3 PV_1 ... -- S_b goes into buffer b % 3, whose previous tenant P_{b-3} is consumed by
        // PV_{b-3}, issued (by this same thread, so executed) before it
        constexpr uint32_t idesc_s = umma_idesc_f16(AQ, AK);            // Q (K-major) x K (K-major)
        constexpr uint32_t idesc_o = umma_idesc_f16(AQ, AD, 1);         // P (TMEM) x V (MN-major)
        const uint64_t qdesc = umma_smem_desc_sw128(smem_u32(smem_q));
        uint32_t g = 0, item_n = 0;                                     // g: running key-block counter
        [[maybe_unused]] unsigned prof[20] = {}; [[maybe_unused]] const unsigned tstart = (unsigned) clock64();
        ItemWalk3 walk(p, blockIdx.x, gridDim.x, total_items);
        while (walk.advance(p)) {
            const Item3 it = walk.cur;
            if (!it.valid()) continue;
            const int nb = (it.len() + AK - 1) / AK;
            PROF_WAIT(0, mbar_wait3(q_full, item_n & 1));
            tc_fence_after_sync();
            for (int b = 0; b < nb + 2; ++b) {
                if (b < nb) {
                    const uint32_t gb = g + b, st = gb % K_STAGES, buf = gb % S_BUFS;
                    PROF_WAIT(1, mbar_wait3(&k_full[st], (gb / K_STAGES) & 1));
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t kdesc = umma_smem_desc_sw128(smem_u32(smem_k + st * kKVBytes));
                        const uint32_t d_tmem = tmem_base + TM_S + buf * 64;
#pragma unroll
                        for (int k = 0; k < AD / 16; ++k) if (!(BERT_ATTN3_ABLATE & 8)) umma_f16_ss(d_tmem, qdesc + 2 * k, kdesc + 2 * k, idesc_s, k != 0);
                        umma_commit(&s_full[buf]);
                        umma_commit(&k_empty[st]);
                        if (b == nb - 1) umma_commit(q_empty);          // Q tile may be overwritten
                    }
                    __syncwarp();
                }
                if (b >= 2) {
                    const uint32_t gp = g + b - 2, st = gp % V_STAGES, buf = gp % S_BUFS;
                    PROF_WAIT(2, mbar_wait3(&p_ready[buf], (gp / S_BUFS) & 1));
                    if (b == 2) PROF_WAIT(3, mbar_wait3(o_free, (item_n & 1) ^ 1));    // previous item's O has been read
                    PROF_WAIT(4, mbar_wait3(&v_full[st], (gp / V_STAGES) & 1));
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t vdesc = umma_smem_desc_sw128_mn(smem_u32(smem_v + st * kKVBytes));
                        const uint32_t a_tmem = tmem_base + TM_S + buf * 64;
#pragma unroll
                        for (int k = 0; k < AK / 16; ++k)                // 16 keys per step: 8 packed columns of P, 16 rows (2 KB) of V
                            if (!(BERT_ATTN3_ABLATE & 8)) umma_f16_ts(tmem_base + TM_O, a_tmem + 8 * k, vdesc + (uint64_t) (128 * k), idesc_o, (b > 2 || k > 0) ? 1u : 0u);
                        umma_commit(&v_empty[st]);
                        umma_commit(&pv_done[buf]);
                    }
                    __syncwarp();
                }
            }
            g += nb;
            ++item_n;
        }
#if BERT_ATTN3_PROF
        if (blockIdx.x == 0 && lane == 0) { prof[5] = (unsigned) clock64() - tstart; prof[6] = g; prof[7] = item_n; for (int i = 0; i < 8; ++i) atomicAdd(&g_attn3_prof[i], (unsigned long long) prof[i]); }
#endif
      }
    } else {
        reg_alloc<208>();
        // ------------------------------------------------------------ softmax + output (4 warps)
        const int q = warp & 3;                               // TMEM lane quadrant of this warp
        const uint32_t lane_base = tmem_base + ((uint32_t) (q * 32) << 16);
        uint32_t g = 0;
        [[maybe_unused]] unsigned prof[20] = {}; [[maybe_unused]] const unsigned tstart = (unsigned) clock64();
        ItemWalk3 walk(p, blockIdx.x, gridDim.x, total_items);
        while (walk.advance(p)) {
            const Item3 it = walk.cur;
            if (!it.valid()) continue;
            const int len = it.len();
            const int nb = (len + AK - 1) / AK;
            float m_ref, l;
            {
                // ra / rc take turns holding the first half of a block (one is consumed while the other is loaded for the next
                // block, so the load lands where it is used: no register copies), rb holds the second half
                uint32_t ra[32], rb[32], rc[32], pk[32];
                auto publish = [&](uint32_t s_addr, uint32_t gb) {      // P over the first half of this S buffer
                    st_p3(s_addr, pk);
                    tmem_st_wait();
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&p_ready[gb % S_BUFS]);
                };
                auto prefetch = [&](uint32_t gn, uint32_t (&dst)[32]) {  // first half of the next block's scores
                    PROF_WAIT(8, mbar_wait3(&s_full[gn % S_BUFS], (gn / S_BUFS) & 1));
                    tc_fence_after_sync();
                    ld_s3(lane_base + TM_S + (gn % S_BUFS) * 64, dst);
                };
                {   // ---- block 0: the reference max is the block's own row max (nothing to rescale yet)
                    PROF_WAIT(9, mbar_wait3(&s_full[g % S_BUFS], (g / S_BUFS) & 1));
                    tc_fence_after_sync();
                    const uint32_t s_addr = lane_base + TM_S + (g % S_BUFS) * 64;
                    ld_s3(s_addr, rc);
                    ld_s3(s_addr + 32, rb);
                    tmem_ld_wait_regs(rc);
                    tmem_ld_wait_regs(rb);
                    if (nb > 1) prefetch(g + 1, ra);
                    mask_half(rc, len); mask_half(rb, len - 32);
                    m_ref = fmaxf(max_half(rc), max_half(rb)) * p.scale_log2;
                    l = exp_half(rc, p.scale_log2, m_ref, pk);
                    l += exp_half(rb, p.scale_log2, m_ref, pk + 16);
                    publish(s_addr, g);
                    if (nb > 1) tmem_ld_wait_regs(ra);
                }
                // ---- blocks 1 ..: `cur` holds the first 32 scores of block b, `nxt` receives those of block b + 1
                auto step = [&](int b, uint32_t (&cur)[32], uint32_t (&nxt)[32]) {
                    const uint32_t gb = g + b;
                    const uint32_t s_addr = lane_base + TM_S + (gb % S_BUFS) * 64;
                    const int kvalid = len - b * AK;
                    ld_s3(s_addr + 32, rb);
                    mask_half(cur, kvalid);
                    const float mxa = max_half(cur);
                    float bsum = exp_half(cur, p.scale_log2, m_ref, pk);
                    tmem_ld_wait_regs(rb);
                    if (b + 1 < nb) prefetch(gb + 1, nxt);
                    mask_half(rb, kvalid - 32);
                    const float mxb = max_half(rb);
                    bsum += exp_half(rb, p.scale_log2, m_ref, pk + 16);
                    const float mx = fmaxf(mxa, mxb) * p.scale_log2;
                    const bool need = mx > m_ref + kRescaleThreshold;
                    if (__any_sync(0xffffffffu, need)) {
                        // rare: some row left the 2^8 window.  The accumulator is rescaled once the previous block's P V has
                        // landed, and the block is redone from the S tile, which is still in TMEM.  S_b complete implies
                        // P V of block b-3 complete, so barrier (gb-1) % 3 is at most one phase behind.
                        float factor = 1.f;
                        if (need) { factor = ex2f3(m_ref - mx); m_ref = mx; }
                        l *= factor;
                        mbar_wait3(&pv_done[(gb - 1) % S_BUFS], ((gb - 1) / S_BUFS) & 1);
                        tc_fence_after_sync();
#pragma unroll 1
                        for (int hh = 0; hh < 2; ++hh) {
                            tmem_ld_32x32b_x32(lane_base + TM_O + hh * 32, rb);
                            tmem_ld_wait_regs(rb);
#pragma unroll
                            for (int j = 0; j < 32; ++j) rb[j] = __float_as_uint(__uint_as_float(rb[j]) * factor);
                            tmem_st_32x32b_x32(lane_base + TM_O + hh * 32, rb);
                        }
                        bsum = 0.f;
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            tmem_ld_32x32b_x32(s_addr + hh * 32, rb);
                            tmem_ld_wait_regs(rb);
                            mask_half(rb, kvalid - hh * 32);
                            bsum += exp_half(rb, p.scale_log2, m_ref, pk + hh * 16);
                        }
                    }
                    l += bsum;
                    publish(s_addr, gb);
                    if (b + 1 < nb) tmem_ld_wait_regs(nxt);
                };
                for (int b = 1; b < nb; b += 2) {
                    step(b, ra, rc);
                    if (b + 1 < nb) step(b + 1, rc, ra);
                }
            }
            g += nb;
            // ---- finalize: O / l, optional q8 fake-quant per 32-wide block, fp16 store of valid rows
            [[maybe_unused]] const unsigned tfin = (unsigned) clock64();
            PROF_WAIT(10, mbar_wait3(&pv_done[(g - 1) % S_BUFS], ((g - 1) / S_BUFS) & 1));
            tc_fence_after_sync();
            uint32_t o0[32], o1[32];
            tmem_ld_32x32b_x32(lane_base + TM_O, o0);
            tmem_ld_32x32b_x32(lane_base + TM_O + 32, o1);
            tmem_ld_wait_regs(o0);
            tmem_ld_wait_regs(o1);
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(o_free);
            const float inv = 1.f / l;
            const float2 inv2 = make_float2(inv, inv);
            const bool fq = p.fq_type == W_Q4_0 || p.fq_type == W_Q4_1;
            uint8_t * stg = smem_o + q * 4096;                  // this warp's [32 rows][128 B], 16-B chunks XOR-swizzled by row
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                float2 v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const uint32_t * o = hh == 0 ? o0 : o1;
                    v[j] = __fmul2_rn(make_float2(__uint_as_float(o[2 * j]), __uint_as_float(o[2 * j + 1])), inv2);
                }
                if (fq) fake_quant32_v2(v, p.fq_type);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 u;
                    u.x = pack_h23(v[4 * c].x, v[4 * c].y); u.y = pack_h23(v[4 * c + 1].x, v[4 * c + 1].y);
                    u.z = pack_h23(v[4 * c + 2].x, v[4 * c + 2].y); u.w = pack_h23(v[4 * c + 3].x, v[4 * c + 3].y);
                    *reinterpret_cast<uint4 *>(stg + lane * 128 + (((hh * 4 + c) ^ (lane & 7)) << 4)) = u;
                }
            }
            __syncwarp();
            // coalesced copy-out: one instruction covers four whole 128-B rows; rows past the sentence end are skipped
            {
                const int rows_valid = len - it.q0 - q * 32;       // of this warp's 32 rows
                __half * obase = p.out + (size_t) (it.t0 + it.q0 + q * 32) * p.H + it.head * AD;
                const int ch = lane & 7, r0 = lane >> 3;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = r0 + 4 * i;
                    const uint4 u = *reinterpret_cast<const uint4 *>(stg + r * 128 + ((ch ^ (r & 7)) << 4));
                    if (r < rows_valid) *reinterpret_cast<uint4 *>(obase + (size_t) r * p.H + ch * 8) = u;
                }
            }
            __syncwarp();                                       // staging is reused by the next item
#if BERT_ATTN3_PROF
            prof[12] += (unsigned) clock64() - tfin;
#endif
        }
#if BERT_ATTN3_PROF
        if (blockIdx.x == 0 && warp == 4 && lane == 0) { prof[11] = (unsigned) clock64() - tstart; for (int i = 8; i < 13; ++i) atomicAdd(&g_attn3_prof[i], (unsigned long long) prof[i]); }
#endif
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, TM_COLS);
}

}  // namespace

// diagnostic: copies and clears the wait-clock counters of a BERT_ATTN3_PROF build (32 values)
void attn3_prof_fetch(unsigned long long * out) {
    cudaMemcpyFromSymbol(out, g_attn3_prof, sizeof(unsigned long long) * 32);
    unsigned long long z[32] = {};
    cudaMemcpyToSymbol(g_attn3_prof, z, sizeof(z));
}

cudaError_t launch_attention_tc3(const CUtensorMap * tm_q, const CUtensorMap * tm_kv, const int32_t * cu_seqlens,
                                 int n_sent, int max_len, int n_head, int d_head, __half * out, int fq_type,
                                 int num_sms, cudaStream_t st) {
    if (n_sent <= 0) return cudaSuccess;
    if (d_head != AD || !tm_q || !tm_kv) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(attention_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kT3Smem);
    if (e != cudaSuccess) return e;
    Attn3Params p;
    p.cu = cu_seqlens; p.out = out; p.n_sent = n_sent; p.n_head = n_head; p.H = n_head * d_head;
    p.max_qt = (max_len + AQ - 1) / AQ; p.fq_type = fq_type;
    p.scale_log2 = 1.4426950408889634f / sqrtf((float) d_head);
    const long long total = (long long) n_sent * n_head * p.max_qt;
    if (total > 0x7fffffffLL) return cudaErrorInvalidValue;
    const int grid = (int) (total < 2LL * num_sms ? total : 2LL * num_sms);
    return launch_pdl(attention_tc3_kernel, dim3(grid), dim3(T3_THREADS), kT3Smem, st, *tm_q, *tm_kv, p);
}

}  // namespace b200
