// K2 (tcgen05) -- variable-length fused attention for head size 64:
//   O = softmax(Q K^T / sqrt(d)) V   per (sentence, head), over the sentence's own tokens only.
// Replaces the reference's mask build (bert.cpp:957-961), scores + scale + mask + soft_max
// (bert.cpp:1018-1025, ggml.c:12008-12077) and P*V + permute/cont/cpy (bert.cpp:1027-1036).
// Nothing of size L x L reaches HBM; padded queries / keys are never stored / always masked.
//
// Work item = (sentence, head, 128-query tile); persistent CTAs (two per SM) walk the items
// round-robin.  Per CTA, warp-specialised:
//   warp 0     TMA producer: Q tile [128 x 64] once per item, K / V blocks [64 keys x 64] through a
//              4-stage ring, straight out of the packed qkv activations (128-B swizzle)
//   warp 1     MMA issuer (one elected lane): S_b = Q K_b^T into a double-buffered TMEM tile
//              (M=128, N=64), then O += P_b V_b with P read from TMEM (A operand) and V as an
//              MN-major smem operand -- V is never transposed
//   warps 2-5  softmax: one query row per thread (tcgen05.ld 32x32b), running max / sum in f32
//              registers, exp2 on the MUFU, P written back over S as packed fp16 (tcgen05.st).
//              The running max is only moved (and O rescaled in TMEM) when a block exceeds it by
//              more than 2^8, so the accumulator normally stays untouched in TMEM until the end.
// Issue order S_0, S_1, PV_0, S_2, PV_1, ... keeps the tensor pipe one score block ahead of the
// softmax warps; the second CTA on the SM fills the remaining bubbles.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int AQ = 128;            // queries per item
constexpr int AK = 64;             // keys per block
constexpr int AD = 64;             // head size
#ifndef BERT_ATTN_KV_STAGES
#define BERT_ATTN_KV_STAGES 4
#endif
#ifndef BERT_ATTN_Q_STAGES
#define BERT_ATTN_Q_STAGES 1
#endif
constexpr int KV_STAGES = BERT_ATTN_KV_STAGES;
constexpr int ATT_TC_THREADS = 192;
constexpr int kQBytes = AQ * AD * 2;          // 16 KB
constexpr int kKVBytes = AK * AD * 2;         // 8 KB each for K and V
constexpr int Q_STAGES = BERT_ATTN_Q_STAGES;
constexpr int kOutStage = 4 * 4096;            // per softmax warp: [32 rows][128 B] staging for coalesced output rows
constexpr int kOnesBytes = 16 * 128;            // 16 rows x 64 fp16 of 1.0: K-major B operand (N = 16) of the row-sum MMA
constexpr int kAttSmem = Q_STAGES * kQBytes + KV_STAGES * 2 * kKVBytes + kOutStage + kOnesBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t TM_S = 0;       // S buffers at columns 0 and 64 (P aliases the first 32 columns of each)
constexpr uint32_t TM_O = 128;     // O accumulator: 64 columns
[[maybe_unused]] constexpr uint32_t TM_L = 192;     // row sums (BERT_ATTN_MMASUM): 16 identical columns
constexpr uint32_t TM_COLS = 256;
constexpr float kRescaleThreshold = 8.f;      // log2 units

struct AttnTcParams {
    const int32_t * cu;            // token offsets of the work units: sentences, or spans of several short sentences (pack_spans)
    const int2 * tok_bounds;       // with spans: per token, [first, end) token of its own sentence (nullptr: units are sentences)
    __half * out;
    int n_sent, n_head, H, max_qt, fq_type;
    float scale_log2;
};

// t1 (the sentence's end offset) is kept raw: the length is only formed where it is used, so the two
// global loads of a prefetched item are not waited for at decode time
struct Item {
    int t0, t1, q0, head;
    __device__ __forceinline__ int len() const { return t1 - t0; }
    __device__ __forceinline__ bool valid() const { return q0 < t1 - t0; }
};

// Round-robin item walk (item = (sentence, head, query tile), query tile fastest).  The coordinates of the
// NEXT item are advanced incrementally (no integer division per item) and its sentence bounds -- two
// dependent-free global loads -- are fetched while the current item is processed.
struct ItemWalk {
    int next_item, total, stride;
    int qt, head, sent;            // coordinates of next_item
    int dq, dh, ds;                // stride split into (query tiles, heads, sentences)
    Item cur, nxt;
    __device__ __forceinline__ void fetch(const AttnTcParams & p) {
        if (next_item >= total) { nxt.t0 = nxt.t1 = nxt.q0 = nxt.head = 0; return; }
        nxt.head = head;
        nxt.q0 = qt * AQ;
        nxt.t0 = __ldg(p.cu + sent);
        nxt.t1 = __ldg(p.cu + sent + 1);
    }
    __device__ __forceinline__ ItemWalk(const AttnTcParams & p, int first, int stride_, int total_)
        : next_item(first), total(total_), stride(stride_) {
        qt = first % p.max_qt; const int hs = first / p.max_qt; head = hs % p.n_head; sent = hs / p.n_head;
        dq = stride % p.max_qt; const int sh = stride / p.max_qt; dh = sh % p.n_head; ds = sh / p.n_head;
        fetch(p);
    }
    // moves to the next item of this CTA; false when the walk is over
    __device__ __forceinline__ bool advance(const AttnTcParams & p) {
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

__device__ __forceinline__ float ex2f(float x) {
    float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
// 2^x for a pair on the FMA / ALU pipes instead of the MUFU (which is this kernel's bottleneck at 16
// results/clk/SM): round-to-nearest split x = n + r with the 1.5 * 2^23 constant, degree-3 minimax
// polynomial for 2^r on [-0.5, 0.5] (max relative error 7.5e-5, well under the fp16 rounding of P), n added
// straight into the exponent field.  x <= ~8 here; everything below -125 (incl. the -inf of masked keys)
// is clamped and comes out as ~2^-125, which rounds to 0 in fp16.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
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
// BERT_ATTN_MMASUM=1: the softmax row sums come out of the tensor core (P x ones tile -> 16 extra TMEM columns next to O)
// instead of 32 FADD2 per block and warp; the sum is then over the fp16-rounded P, i.e. exactly what multiplies V.
#ifndef BERT_ATTN_MMASUM
#define BERT_ATTN_MMASUM 1
#endif
// BERT_ATTN_SPEC=1 (needs BERT_ATTN_MMASUM): exponentials against the current reference max, row max checked afterwards
#ifndef BERT_ATTN_SPEC
#define BERT_ATTN_SPEC 0
#endif
#if BERT_ATTN_SPEC && !BERT_ATTN_MMASUM
#error "BERT_ATTN_SPEC needs BERT_ATTN_MMASUM"
#endif
#if BERT_ATTN_SPEC
#error "BERT_ATTN_SPEC predates the packed spans (per-row key bounds) and is no longer built"
#endif
#if !BERT_ATTN_MMASUM
#error "the MMA issuer always forms the row sums (P x ones): BERT_ATTN_MMASUM=0 is no longer built"
#endif
#ifndef BERT_ATTN_LDW
#define BERT_ATTN_LDW 0
#endif
#ifndef BERT_ATTN_POLY_EVERY
#define BERT_ATTN_POLY_EVERY 4
#endif
constexpr int kPolyEvery = BERT_ATTN_POLY_EVERY;   // every kPolyEvery-th pair of a block takes the polynomial path (0: all on the MUFU)
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

__global__ void __launch_bounds__(ATT_TC_THREADS, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                    const AttnTcParams p) {
    extern __shared__ uint8_t smem_raw[];
    // 1024-B alignment by an offset from the array (not an integer round trip), so that the compiler keeps the
    // shared address space and emits LDS/STS rather than generic LD/ST
    uint8_t * smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t * smem_q = smem;
    uint8_t * smem_k = smem + Q_STAGES * kQBytes;                    // KV_STAGES x 8 KB
    uint8_t * smem_v = smem_k + KV_STAGES * kKVBytes;                // KV_STAGES x 8 KB
    uint8_t * smem_o = smem_v + KV_STAGES * kKVBytes;                // 4 x 4 KB output staging
    uint8_t * smem_ones = smem_o + kOutStage;                        // 2 KB of fp16 1.0
    uint64_t * bars = (uint64_t *) (smem_ones + kOnesBytes);
    uint64_t * q_full = bars;                   // [Q_STAGES]
    uint64_t * q_empty = bars + Q_STAGES;       // [Q_STAGES]
    uint64_t * kv_full = q_empty + Q_STAGES;    // [KV_STAGES]
    uint64_t * kv_empty = kv_full + KV_STAGES;  // [KV_STAGES]
    uint64_t * s_full = kv_empty + KV_STAGES;   // [2]
    uint64_t * p_ready = s_full + 2;            // [2]
    uint64_t * pv_done = p_ready + 2;           // [2]  P V of block g has reached the accumulator (barrier g & 1)
    uint64_t * o_free = pv_done + 2;            // [1]
    uint32_t * tmem_slot = (uint32_t *) (o_free + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total_items = p.n_sent * p.n_head * p.max_qt;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap_kv);
        for (int a = 0; a < Q_STAGES; ++a) { mbar_init(&q_full[a], 1); mbar_init(&q_empty[a], 1); }
        for (int s = 0; s < KV_STAGES; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&s_full[a], 1); mbar_init(&p_ready[a], 4); }
        mbar_init(&pv_done[0], 1); mbar_init(&pv_done[1], 1);
        mbar_init(o_free, 4);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, TM_COLS); tmem_relinquish(); }
#if BERT_ATTN_MMASUM
    for (int i = threadIdx.x; i < kOnesBytes / 4; i += ATT_TC_THREADS) reinterpret_cast<uint32_t *>(smem_ones)[i] = 0x3C003C00u;
    fence_proxy_async_smem();
#endif
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();                                 // everything above overlapped the tail of the previous kernel

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            uint32_t kv_it = 0, item_n = 0;
            ItemWalk walk(p, blockIdx.x, gridDim.x, total_items);
            while (walk.advance(p)) {
                const Item it = walk.cur;
                if (!it.valid()) continue;
                const int len = it.len();
            const int nb = (len + AK - 1) / AK;
                const uint32_t qs = item_n % Q_STAGES, qph = (item_n / Q_STAGES) & 1;
                mbar_wait(&q_empty[qs], qph ^ 1);
                mbar_expect_tx(&q_full[qs], kQBytes);
                tma_load_2d(smem_q + qs * kQBytes, &tmap_q, &q_full[qs], it.head * AD, it.t0 + it.q0);
                for (int b = 0; b < nb; ++b, ++kv_it) {
                    const uint32_t st = kv_it % KV_STAGES, ph = (kv_it / KV_STAGES) & 1;
                    mbar_wait(&kv_empty[st], ph ^ 1);
                    mbar_expect_tx(&kv_full[st], 2 * kKVBytes);
                    tma_load_2d(smem_k + st * kKVBytes, &tmap_kv, &kv_full[st], p.H + it.head * AD, it.t0 + b * AK);
                    tma_load_2d(smem_v + st * kKVBytes, &tmap_kv, &kv_full[st], 2 * p.H + it.head * AD, it.t0 + b * AK);
                }
                ++item_n;
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        // Everything that does not change is formed once: the descriptors of the Q tile, of the first K / V stage and of the
        // ones tile, the S buffers' TMEM addresses; a stage's descriptor is then one add.  The round-2 wait-clock counters put
        // this issue loop at ~1 000 clocks per key block (profiles/r2_attention.md); hoisting cut its instructions by a quarter and
        // changed nothing measurable.  (The warp stays converged and elects the issuing lane per step: under a
        // `lane == 0` branch ptxas wraps every uniform-datapath instruction in its own ELECT / BRA.U.ANY loop.)
        {
            constexpr uint32_t idesc_s = umma_idesc_f16(AQ, AK);            // Q (K-major) x K (K-major)
            constexpr uint32_t idesc_o = umma_idesc_f16(AQ, AD, 1);         // P (TMEM) x V (MN-major)
            constexpr uint32_t idesc_l = umma_idesc_f16(AQ, 16);            // P (TMEM) x ones (K-major, N = 16)
            const uint64_t onesdesc = umma_smem_desc_sw128(smem_u32(smem_ones));
            const uint64_t qdesc0 = umma_smem_desc_sw128(smem_u32(smem_q));
            const uint64_t kdesc0 = umma_smem_desc_sw128(smem_u32(smem_k));
            const uint64_t vdesc0 = umma_smem_desc_sw128_mn(smem_u32(smem_v));
            const uint32_t tm_s = tmem_base + TM_S, tm_o = tmem_base + TM_O, tm_l = tmem_base + TM_L;
            uint32_t g = 0, item_n = 0;                                     // g: running key-block counter
            ItemWalk walk(p, blockIdx.x, gridDim.x, total_items);
            while (walk.advance(p)) {
                const Item it = walk.cur;
                if (!it.valid()) continue;
                const int nb = (it.len() + AK - 1) / AK;
                const uint32_t qs = item_n % Q_STAGES;
                const uint64_t qdesc = qdesc0 + (uint64_t) (qs * (kQBytes >> 4));
                mbar_wait(&q_full[qs], (item_n / Q_STAGES) & 1);
                for (int b = 0; b <= nb; ++b) {
                    if (b < nb) {
                        const uint32_t gb = g + b, st = gb % KV_STAGES;
                        mbar_wait(&kv_full[st], (gb / KV_STAGES) & 1);
                        tc_fence_after_sync();
                        if (elect_one()) {
                            const uint64_t kdesc = kdesc0 + (uint64_t) (st * (kKVBytes >> 4));
                            const uint32_t d_tmem = tm_s + (gb & 1) * 64;
#pragma unroll
                            for (int k = 0; k < AD / 16; ++k) umma_f16_ss(d_tmem, qdesc + 2 * k, kdesc + 2 * k, idesc_s, k != 0);
                            umma_commit(&s_full[gb & 1]);
                            if (b == nb - 1) umma_commit(&q_empty[qs]);     // Q tile may be overwritten
                        }
                        __syncwarp();
                    }
                    if (b >= 1) {
                        const uint32_t gp = g + b - 1, st = gp % KV_STAGES;
                        mbar_wait(&p_ready[gp & 1], (gp >> 1) & 1);
                        if (b == 1) mbar_wait(o_free, (item_n & 1) ^ 1);    // previous item's O has been read
                        tc_fence_after_sync();
                        if (elect_one()) {
                            const uint64_t vdesc = vdesc0 + (uint64_t) (st * (kKVBytes >> 4));
                            const uint32_t a_tmem = tm_s + (gp & 1) * 64;
                            const uint32_t acc = b > 1 ? 1u : 0u;
#pragma unroll
                            for (int k = 0; k < AK / 16; ++k)                // 16 keys per step: 8 packed columns of P, 16 rows (2 KB) of V
                                umma_f16_ts(tm_o, a_tmem + 8 * k, vdesc + (uint64_t) (128 * k), idesc_o, k > 0 ? 1u : acc);
#pragma unroll
                            for (int k = 0; k < AK / 16; ++k)                // row sums: P x ones
                                umma_f16_ts(tm_l, a_tmem + 8 * k, onesdesc + 2 * k, idesc_l, k > 0 ? 1u : acc);
                            umma_commit(&kv_empty[st]);
                            umma_commit(&pv_done[gp & 1]);
                        }
                        __syncwarp();
                    }
                }
                g += nb;
                ++item_n;
            }
        }
    } else {
        // ------------------------------------------------------------ softmax + output (4 warps)
        const int q = warp & 3;                               // TMEM lane quadrant of this warp
        const uint32_t lane_base = tmem_base + ((uint32_t) (q * 32) << 16);
        uint32_t g = 0;
        ItemWalk walk(p, blockIdx.x, gridDim.x, total_items);
        while (walk.advance(p)) {
            const Item it = walk.cur;
            if (!it.valid()) continue;
            const int len = it.len();
            const int nb = (len + AK - 1) / AK;
            float m_ref = -INFINITY, l = 0.f;
            // keys this thread's query row may see, relative to the item's first token: the whole unit, or (packed spans,
            // see attention_tc2.cu) the row's own sentence
            int key_lo = 0, key_hi = len;
            if (p.tok_bounds != nullptr) {
                const int row = it.q0 + q * 32 + lane;
                if (row < len) { const int2 bnd = __ldg(p.tok_bounds + it.t0 + row); key_lo = bnd.x - it.t0; key_hi = bnd.y - it.t0; }
            }
            for (int b = 0; b < nb; ++b) {
                const uint32_t gb = g + b;
                mbar_wait(&s_full[gb & 1], (gb >> 1) & 1);
                tc_fence_after_sync();
                const uint32_t s_addr = lane_base + TM_S + (gb & 1) * 64;
#if BERT_ATTN_SPEC
                // Speculative form: the exponentials are taken against the CURRENT reference max while the row max of
                // the block is formed alongside (no dependency between the two, so the FMNMX chain and the vote leave
                // the per-block critical path); only if some row exceeds its reference by more than 2^8 -- rare after
                // the first block -- the accumulator is rescaled and the block redone from the S tile still in TMEM.
                const int kvalid = len - b * AK;
                auto load_s = [&](float (&sv)[64]) {
                    uint32_t r0[32], r1[32];
                    tmem_ld_32x32b_x32(s_addr, r0);
                    tmem_ld_32x32b_x32(s_addr + 32, r1);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) { sv[j] = __uint_as_float(r0[j]); sv[32 + j] = __uint_as_float(r1[j]); }
                    if (kvalid < AK) {
#pragma unroll
                        for (int j = 0; j < 64; ++j) if (j >= kvalid) sv[j] = -INFINITY;
                    }
                };
                auto row_max = [&](const float (&sv)[64]) {
                    float mx4[4] = {sv[0], sv[1], sv[2], sv[3]};
#pragma unroll
                    for (int j = 4; j < 64; j += 4) {
                        mx4[0] = fmaxf(mx4[0], sv[j]); mx4[1] = fmaxf(mx4[1], sv[j + 1]);
                        mx4[2] = fmaxf(mx4[2], sv[j + 2]); mx4[3] = fmaxf(mx4[3], sv[j + 3]);
                    }
                    return fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
                };
                auto exp_pack = [&](const float (&sv)[64], float mref, uint32_t (&pkv)[32]) {
                    const float2 nm2 = make_float2(-mref, -mref), sc2 = make_float2(p.scale_log2, p.scale_log2);
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float2 x = __ffma2_rn(make_float2(sv[2 * j], sv[2 * j + 1]), sc2, nm2);
                        const float2 e = (kPolyEvery > 0 && j % (kPolyEvery > 0 ? kPolyEvery : 1) == kPolyEvery - 1)
                                             ? ex2_poly2(x) : make_float2(ex2f(x.x), ex2f(x.y));
                        pkv[j] = pack_h2(e.x, e.y);
                    }
                };
                float s[64];
                uint32_t pk[32];
                load_s(s);
                const float mx = row_max(s);
                if (b == 0) m_ref = mx;                         // first block: O is overwritten, nothing to rescale
                exp_pack(s, m_ref, pk);
                if (b >= 1 && __any_sync(0xffffffffu, mx > m_ref + kRescaleThreshold)) {
                    float factor = 1.f;
                    if (mx > m_ref + kRescaleThreshold) { factor = ex2f(m_ref - mx); m_ref = mx; }
                    mbar_wait(&pv_done[(gb - 1) & 1], ((gb - 1) >> 1) & 1);
                    tc_fence_after_sync();
#pragma unroll 1
                    for (int hh = 0; hh < 3; ++hh) {            // O (64 columns) and the row sums at TM_L = TM_O + 64
                        uint32_t o[32];
                        tmem_ld_32x32b_x32(lane_base + TM_O + hh * 32, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
                        tmem_st_32x32b_x32(lane_base + TM_O + hh * 32, o);
                    }
                    load_s(s);
                    exp_pack(s, m_ref, pk);
                }
#else
                uint32_t r0[32], r1[32];
#if BERT_ATTN_LDW
                tmem_ld_32x32b_x32x2_wait(s_addr, r0, s_addr + 32, r1);
#else
                tmem_ld_32x32b_x32(s_addr, r0);
                tmem_ld_32x32b_x32(s_addr + 32, r1);
                tmem_ld_wait();
#endif
                float s[64];
#pragma unroll
                for (int j = 0; j < 32; ++j) { s[j] = __uint_as_float(r0[j]); s[32 + j] = __uint_as_float(r1[j]); }
                const int klo = key_lo - b * AK, khi = key_hi - b * AK;     // this row's keys [klo, khi) of the block's 64
                if (klo > 0 || khi < AK) {
#pragma unroll
                    for (int j = 0; j < 64; ++j) if (j < klo || j >= khi) s[j] = -INFINITY;
                }
                float mx4[4] = {s[0], s[1], s[2], s[3]};      // four independent chains (a single one is latency-bound)
#pragma unroll
                for (int j = 4; j < 64; j += 4) {
                    mx4[0] = fmaxf(mx4[0], s[j]); mx4[1] = fmaxf(mx4[1], s[j + 1]);
                    mx4[2] = fmaxf(mx4[2], s[j + 2]); mx4[3] = fmaxf(mx4[3], s[j + 3]);
                }
                // (a row whose sentence has no key in this block sees -inf only: the floor keeps its reference finite, its
                // exponentials come out as 0 and nothing has to be rescaled until its first real block)
                const float mx = fmaxf(fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2, -1e30f);
                // move the reference max only on the first block or when it is exceeded by > 2^8
                const bool need = mx > m_ref + kRescaleThreshold;       // always true on the first block (m_ref = -inf)
                const bool rescale = need && m_ref > -1e29f;            // the accumulator only holds something after a real block
                float factor = 1.f;
                if (need) { factor = ex2f(m_ref - mx); m_ref = mx; l *= factor; }   // (l unused with BERT_ATTN_MMASUM)
                if (b >= 1 && __any_sync(0xffffffffu, rescale)) {
                    // rare: the previous block's P V must be in the accumulator before it may be rescaled.
                    // The wait is only paid here: S_b complete (s_full above) implies P V of block b-2 complete
                    // (issue order ..., PV_{b-2}, S_b, PV_{b-1}), so barrier (gb-1)&1 is at most one phase behind.
                    mbar_wait(&pv_done[(gb - 1) & 1], ((gb - 1) >> 1) & 1);
                    tc_fence_after_sync();
#pragma unroll 1
                    for (int hh = 0; hh < (BERT_ATTN_MMASUM ? 3 : 2); ++hh) {   // 32 columns at a time (third pass: the row sums at TM_L = TM_O + 64)
                        uint32_t o[32];
                        tmem_ld_32x32b_x32(lane_base + TM_O + hh * 32, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
                        tmem_st_32x32b_x32(lane_base + TM_O + hh * 32, o);
                    }
                }
                const float2 nm2 = make_float2(-m_ref, -m_ref), sc2 = make_float2(p.scale_log2, p.scale_log2);
                float2 sum[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
                uint32_t pk[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float2 x = __ffma2_rn(make_float2(s[2 * j], s[2 * j + 1]), sc2, nm2);
                    const float2 e = (kPolyEvery > 0 && j % (kPolyEvery > 0 ? kPolyEvery : 1) == kPolyEvery - 1)
                                         ? ex2_poly2(x) : make_float2(ex2f(x.x), ex2f(x.y));
#if !BERT_ATTN_MMASUM
                    sum[j & 1] = __fadd2_rn(sum[j & 1], e);
#endif
                    pk[j] = pack_h2(e.x, e.y);
                }
                l += (sum[0].x + sum[0].y) + (sum[1].x + sum[1].y);
#endif
                tmem_st_32x32b_x32(s_addr, pk);               // P over the first half of this S buffer
                tmem_st_wait();
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_ready[gb & 1]);
            }
            g += nb;
            // ---- finalize: O / l, optional q8 fake-quant per 32-wide block, fp16 store of valid rows
            mbar_wait(&pv_done[(g - 1) & 1], ((g - 1) >> 1) & 1);
            tc_fence_after_sync();
#if BERT_ATTN_MMASUM
            {
                uint32_t lsum;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\n\ttcgen05.wait::ld.sync.aligned;"
                             : "=r"(lsum) : "r"(lane_base + TM_L) : "memory");
                l = __uint_as_float(lsum);
            }
#endif
            uint32_t o0[32], o1[32];
#if BERT_ATTN_LDW
            tmem_ld_32x32b_x32x2_wait(lane_base + TM_O, o0, lane_base + TM_O + 32, o1);
#else
            tmem_ld_32x32b_x32(lane_base + TM_O, o0);
            tmem_ld_32x32b_x32(lane_base + TM_O + 32, o1);
            tmem_ld_wait();
#endif
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
                    u.x = pack_h2(v[4 * c].x, v[4 * c].y); u.y = pack_h2(v[4 * c + 1].x, v[4 * c + 1].y);
                    u.z = pack_h2(v[4 * c + 2].x, v[4 * c + 2].y); u.w = pack_h2(v[4 * c + 3].x, v[4 * c + 3].y);
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
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, TM_COLS);
}

}  // namespace

bool make_attention_tmaps(CUtensorMap * tm_q, CUtensorMap * tm_kv, const __half * qkv, uint64_t rows, int H) {
    return make_tmap_f16_2d(tm_q, qkv, rows, (uint64_t) 3 * H, AQ) && make_tmap_f16_2d(tm_kv, qkv, rows, (uint64_t) 3 * H, AK);
}

cudaError_t launch_attention_tc(const CUtensorMap * tm_q, const CUtensorMap * tm_kv, const int32_t * cu_seqlens,
                                int n_sent, int max_len, int n_head, int d_head, __half * out, int fq_type,
                                int num_sms, cudaStream_t st, const int2 * tok_bounds) {
    if (n_sent <= 0) return cudaSuccess;
    if (d_head != AD || !tm_q || !tm_kv) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttSmem);
    if (e != cudaSuccess) return e;
    AttnTcParams p;
    p.cu = cu_seqlens; p.tok_bounds = tok_bounds; p.out = out; p.n_sent = n_sent; p.n_head = n_head; p.H = n_head * d_head;
    p.max_qt = (max_len + AQ - 1) / AQ; p.fq_type = fq_type;
    p.scale_log2 = 1.4426950408889634f / sqrtf((float) d_head);
    const long long total = (long long) n_sent * n_head * p.max_qt;
    if (total > 0x7fffffffLL) return cudaErrorInvalidValue;
    const int grid = (int) (total < 2LL * num_sms ? total : 2LL * num_sms);
    return launch_pdl(attention_tc_kernel, dim3(grid), dim3(ATT_TC_THREADS), kAttSmem, st, *tm_q, *tm_kv, p);
}

}  // namespace b200
