// K2 (tcgen05), head sizes 64 AND 32 -- variable-length fused attention
//   O = softmax(Q K^T / sqrt(d)) V   per (sentence, head), over the sentence's own tokens only.
// Replaces the reference's mask build (bert.cpp:957-961), scores + scale + mask + soft_max
// (bert.cpp:1018-1025, ggml.c:12008-12077) and P*V + permute/cont/cpy (bert.cpp:1027-1036).
// Nothing of size L x L reaches HBM; padded queries / keys are never stored / always masked.
//
// Round-2 layout (profiles/r2_attention.md).  The round-1 kernel was bound by the per-block dependency chain of its
// softmax warps (two warps per scheduler, one query row x 64 keys per thread).  Here:
//   * a work item is (sentence, 64 columns of the hidden state, 128-query tile): ONE head of size 64, or TWO heads of
//     size 32 that share the item's Q / K / V tiles (the 64-column TMA boxes are the same either way);
//   * the item's key blocks are walked as "virtual blocks" v = block * NH + head-in-item; S_v alternates between two
//     TMEM buffers, so for head size 32 the two buffers hold the two heads of one key block;
//   * EIGHT softmax warps per CTA: every TMEM lane quadrant is served by two warps that split the 64 keys of a block,
//     i.e. a thread owns one query row x 32 keys.  Half the serial work per warp and four warps per scheduler (two CTAs
//     per SM) instead of two.  The two warps of a quadrant exchange their half-row maxima through shared memory
//     (one named barrier of 64 threads per block), which also orders the P store of one against the S load of the other.
//   * short sentences are PACKED: the host groups consecutive sentences into spans of at most 128 tokens (pack_spans) and a
//     work item serves a whole span -- one Q tile, at most two key blocks -- with every query row masked to the keys of its
//     own sentence (per-token sentence bounds written by the embedding kernel).  The 1 738 lines of the all-MiniLM workload
//     (22 tokens on average) are 62 k work items per layer unpacked, each with ~4 000 clocks of fixed cost.
// Per CTA: warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 softmax + output.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int AQ = 128;            // queries per item
constexpr int AK = 64;             // keys per block
constexpr int AW = 64;             // hidden-state columns per item (one head of 64 or two heads of 32)
constexpr int KV_STAGES = 4;
constexpr int SM_WARPS = 8;        // softmax warps
constexpr int ATT2_THREADS = 32 * (2 + SM_WARPS);
constexpr int kQBytes = AQ * AW * 2;           // 16 KB
constexpr int kKVBytes = AK * AW * 2;          // 8 KB each for K and V
constexpr int kOutStage = 4 * 4096;            // per lane quadrant: [32 rows][128 B] staging for coalesced output rows
constexpr int kOnesBytes = 16 * 128;           // 16 rows x 64 fp16 of 1.0: K-major B operand (N = 16) of the row-sum MMA
constexpr int kMaxBytes = 2 * 4 * 32 * 2 * 4;  // half-row maxima: [block parity][quadrant][row][half] f32
constexpr int kAtt2Smem = kQBytes + KV_STAGES * 2 * kKVBytes + kOutStage + kOnesBytes + kMaxBytes + 1024 /*align*/ + 256 /*barriers*/;
constexpr uint32_t TM_S = 0;       // S buffers at columns 0 and 64 (P aliases the first 32 columns of each)
constexpr uint32_t TM_O = 128;     // O accumulator: 64 columns (two heads of 32 side by side)
constexpr uint32_t TM_L = 192;     // row sums: 16 identical columns per head
constexpr uint32_t TM_COLS = 256;
constexpr float kRescaleThreshold = 8.f;      // log2 units

struct Attn2Params {
    const int32_t * cu;            // token offsets of the work units: sentences, or spans of several short sentences (pack_spans)
    const int2 * tok_bounds;       // with spans: per token, [first, end) token of its own sentence (nullptr: units are sentences)
    __half * out;
    int n_sent, n_grp, H, max_qt, fq_type;     // n_grp = H / 64 column groups
    float scale_log2;
};

struct Item2 {
    int t0, t1, q0, grp;
    __device__ __forceinline__ int len() const { return t1 - t0; }
    __device__ __forceinline__ bool valid() const { return q0 < t1 - t0; }
};

// Round-robin item walk (item = (sentence, column group, query tile), query tile fastest); the coordinates of the NEXT
// item are advanced incrementally and its sentence bounds are fetched while the current item is processed.
struct ItemWalk2 {
    int next_item, total, stride;
    int qt, grp, sent;
    int dq, dg, ds;
    Item2 cur, nxt;
    __device__ __forceinline__ void fetch(const Attn2Params & p) {
        if (next_item >= total) { nxt.t0 = nxt.t1 = nxt.q0 = nxt.grp = 0; return; }
        nxt.grp = grp;
        nxt.q0 = qt * AQ;
        nxt.t0 = __ldg(p.cu + sent);
        nxt.t1 = __ldg(p.cu + sent + 1);
    }
    __device__ __forceinline__ ItemWalk2(const Attn2Params & p, int first, int stride_, int total_)
        : next_item(first), total(total_), stride(stride_) {
        qt = first % p.max_qt; const int hs = first / p.max_qt; grp = hs % p.n_grp; sent = hs / p.n_grp;
        dq = stride % p.max_qt; const int sh = stride / p.max_qt; dg = sh % p.n_grp; ds = sh / p.n_grp;
        fetch(p);
    }
    __device__ __forceinline__ bool advance(const Attn2Params & p) {
        if (next_item >= total) return false;
        cur = nxt;
        next_item += stride;
        qt += dq; grp += dg; sent += ds;
        if (qt >= p.max_qt) { qt -= p.max_qt; ++grp; }
        if (grp >= p.n_grp) { grp -= p.n_grp; ++sent; }
        fetch(p);
        return true;
    }
};

__device__ __forceinline__ float ex2f2(float x) {
    float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
// 2^x for a pair on the FMA / ALU pipes instead of the MUFU (16 results/clk/SM): round-to-nearest split x = n + r with
// the 1.5 * 2^23 constant, degree-3 minimax polynomial for 2^r on [-0.5, 0.5] (max relative error 7.5e-5, well under the
// fp16 rounding of P), n added straight into the exponent field.  Everything below -125 (incl. the -inf of masked keys)
// is clamped and comes out as ~2^-125, which rounds to 0 in fp16.
__device__ __forceinline__ float2 ex2_poly2b(float2 x) {
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
#ifndef BERT_ATTN2_POLY_EVERY
#define BERT_ATTN2_POLY_EVERY 4
#endif
constexpr int kPolyEvery2 = BERT_ATTN2_POLY_EVERY;   // every kPolyEvery2-th pair takes the polynomial path (0: all on the MUFU)
__device__ __forceinline__ uint32_t pack_h2b(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}
// TMEM <-> registers, 16 columns
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

template <int D>
__global__ void __launch_bounds__(ATT2_THREADS, 2)
attention_tc2_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                     const Attn2Params p) {
    static_assert(D == 64 || D == 32, "head size");
    constexpr int NH = AW / D;                    // heads per item
    constexpr int KS = D / 16;                    // UMMA K steps of one head's Q K^T
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t * smem_q = smem;
    uint8_t * smem_k = smem + kQBytes;                               // KV_STAGES x 8 KB
    uint8_t * smem_v = smem_k + KV_STAGES * kKVBytes;                // KV_STAGES x 8 KB
    uint8_t * smem_o = smem_v + KV_STAGES * kKVBytes;                // 4 x 4 KB output staging
    uint8_t * smem_ones = smem_o + kOutStage;                        // 2 KB of fp16 1.0
    float * smem_mx = reinterpret_cast<float *>(smem_ones + kOnesBytes);
    uint64_t * bars = (uint64_t *) (smem_ones + kOnesBytes + kMaxBytes);
    uint64_t * q_full = bars;                   // [1]
    uint64_t * q_empty = bars + 1;              // [1]
    uint64_t * kv_full = bars + 2;              // [KV_STAGES]
    uint64_t * kv_empty = kv_full + KV_STAGES;  // [KV_STAGES]
    uint64_t * s_full = kv_empty + KV_STAGES;   // [2]
    uint64_t * p_ready = s_full + 2;            // [2]
    uint64_t * pv_done = p_ready + 2;           // [2]  P V of virtual block g has reached its accumulator (barrier g & 1)
    uint64_t * o_free = pv_done + 2;            // [1]
    uint32_t * tmem_slot = (uint32_t *) (o_free + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total_items = p.n_sent * p.n_grp * p.max_qt;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap_kv);
        mbar_init(q_full, 1); mbar_init(q_empty, 1);
        for (int s = 0; s < KV_STAGES; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&s_full[a], 1); mbar_init(&p_ready[a], SM_WARPS); mbar_init(&pv_done[a], 1); }
        mbar_init(o_free, SM_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, TM_COLS); tmem_relinquish(); }
    for (int i = threadIdx.x; i < kOnesBytes / 4; i += ATT2_THREADS) reinterpret_cast<uint32_t *>(smem_ones)[i] = 0x3C003C00u;
    fence_proxy_async_smem();
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
            ItemWalk2 walk(p, blockIdx.x, gridDim.x, total_items);
            while (walk.advance(p)) {
                const Item2 it = walk.cur;
                if (!it.valid()) continue;
                const int nb = (it.len() + AK - 1) / AK;
                mbar_wait(q_empty, (item_n & 1) ^ 1);
                mbar_expect_tx(q_full, kQBytes);
                tma_load_2d(smem_q, &tmap_q, q_full, it.grp * AW, it.t0 + it.q0);
                for (int b = 0; b < nb; ++b, ++kv_it) {
                    const uint32_t st = kv_it % KV_STAGES, ph = (kv_it / KV_STAGES) & 1;
                    mbar_wait(&kv_empty[st], ph ^ 1);
                    mbar_expect_tx(&kv_full[st], 2 * kKVBytes);
                    tma_load_2d(smem_k + st * kKVBytes, &tmap_kv, &kv_full[st], p.H + it.grp * AW, it.t0 + b * AK);
                    tma_load_2d(smem_v + st * kKVBytes, &tmap_kv, &kv_full[st], 2 * p.H + it.grp * AW, it.t0 + b * AK);
                }
                ++item_n;
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        // issue order S_0, S_1, PV_0, S_2, PV_1, ... over the item's virtual blocks keeps the tensor pipe one score
        // block ahead of the softmax warps
        constexpr uint32_t idesc_s = umma_idesc_f16(AQ, AK);            // Q (K-major) x K (K-major)
        constexpr uint32_t idesc_o = umma_idesc_f16(AQ, D, 1);          // P (TMEM) x V (MN-major), N = head size
        constexpr uint32_t idesc_l = umma_idesc_f16(AQ, 16);            // P (TMEM) x ones (K-major, N = 16)
        const uint64_t onesdesc = umma_smem_desc_sw128(smem_u32(smem_ones));
        const uint64_t qdesc = umma_smem_desc_sw128(smem_u32(smem_q));
        // (a stage's descriptor is the first stage's plus a constant: formed once, as in attention_tc.cu)
        const uint64_t kdesc0 = umma_smem_desc_sw128(smem_u32(smem_k));
        const uint64_t vdesc0 = umma_smem_desc_sw128_mn(smem_u32(smem_v));
        uint32_t gv = 0, gkv = 0, item_n = 0;
        ItemWalk2 walk(p, blockIdx.x, gridDim.x, total_items);
        while (walk.advance(p)) {
            const Item2 it = walk.cur;
            if (!it.valid()) continue;
            const int nb = (it.len() + AK - 1) / AK;
            const int nvb = nb * NH;
            mbar_wait(q_full, item_n & 1);
            tc_fence_after_sync();
            for (int v = 0; v <= nvb; ++v) {
                if (v < nvb) {
                    const int b = v / NH, hs = v % NH;
                    const uint32_t kb = gkv + b, st = kb % KV_STAGES, g = gv + v;
                    if (hs == 0) mbar_wait(&kv_full[st], (kb / KV_STAGES) & 1);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t kdesc = kdesc0 + (uint64_t) (st * (kKVBytes >> 4));
                        const uint32_t d_tmem = tmem_base + TM_S + (g & 1) * 64;
#pragma unroll
                        for (int k = 0; k < KS; ++k) {
                            const int kk = hs * KS + k;                 // this head's columns of the 64-wide tiles
                            umma_f16_ss(d_tmem, qdesc + 2 * kk, kdesc + 2 * kk, idesc_s, k != 0);
                        }
                        umma_commit(&s_full[g & 1]);
                        if (v == nvb - 1) umma_commit(q_empty);         // Q tile may be overwritten
                    }
                    __syncwarp();
                }
                if (v >= 1) {
                    const int vp = v - 1, bp = vp / NH, hp = vp % NH;
                    const uint32_t kb = gkv + bp, st = kb % KV_STAGES, g = gv + vp;
                    mbar_wait(&p_ready[g & 1], (g >> 1) & 1);
                    if (vp == 0) mbar_wait(o_free, (item_n & 1) ^ 1);   // previous item's O has been read
                    tc_fence_after_sync();
                    if (elect_one()) {
                        // V tile [64 keys][64 columns] read MN-major; the second head of a pair starts 64 B into each row
                        const uint64_t vdesc = vdesc0 + (uint64_t) (st * (kKVBytes >> 4) + hp * (D * 2 / 16));
                        const uint32_t a_tmem = tmem_base + TM_S + (g & 1) * 64;
                        const uint32_t o_tmem = tmem_base + TM_O + hp * D;
                        const uint32_t l_tmem = tmem_base + TM_L + hp * 16;
#pragma unroll
                        for (int k = 0; k < AK / 16; ++k)                // 16 keys per step: 8 packed columns of P, 16 rows of V
                            umma_f16_ts(o_tmem, a_tmem + 8 * k, vdesc + (uint64_t) (128 * k), idesc_o, (bp > 0 || k > 0) ? 1u : 0u);
#pragma unroll
                        for (int k = 0; k < AK / 16; ++k)                // row sums of exactly the fp16 P that multiplies V
                            umma_f16_ts(l_tmem, a_tmem + 8 * k, onesdesc + 2 * k, idesc_l, (bp > 0 || k > 0) ? 1u : 0u);
                        if (hp == NH - 1) umma_commit(&kv_empty[st]);
                        umma_commit(&pv_done[g & 1]);
                    }
                    __syncwarp();
                }
            }
            gv += nvb; gkv += nb;
            ++item_n;
        }
    } else {
        // ------------------------------------------------------------ softmax + output (8 warps)
        const int q = warp & 3;                               // TMEM lane quadrant of this warp
        const int hcol = (warp - 2) >> 2;                     // which 32 keys of a block / which 32 output columns
        const uint32_t lane_base = tmem_base + ((uint32_t) (q * 32) << 16);
        const uint32_t bar_id = 1 + q;                        // named barrier of the quadrant's two warps
        uint32_t g = 0;
        ItemWalk2 walk(p, blockIdx.x, gridDim.x, total_items);
        while (walk.advance(p)) {
            const Item2 it = walk.cur;
            if (!it.valid()) continue;
            const int len = it.len();
            const int nb = (len + AK - 1) / AK;
            const int nvb = nb * NH;
            float m0 = -INFINITY, m1 = -INFINITY;             // reference max per head of the item (log2 units)
            // keys this thread's query row may see, relative to the item's first token: the whole unit, or (packed spans) the
            // row's own sentence
            int key_lo = 0, key_hi = len;
            if (p.tok_bounds != nullptr) {
                const int row = it.q0 + q * 32 + lane;
                if (row < len) { const int2 bnd = __ldg(p.tok_bounds + it.t0 + row); key_lo = bnd.x - it.t0; key_hi = bnd.y - it.t0; }
            }
            for (int v = 0; v < nvb; ++v) {
                const int b = v / NH, hs = v % NH;
                const uint32_t gb = g + v;
                mbar_wait(&s_full[gb & 1], (gb >> 1) & 1);
                tc_fence_after_sync();
                const uint32_t s_addr = lane_base + TM_S + (gb & 1) * 64;
                uint32_t r[32];
                tmem_ld_32x32b_x32(s_addr + 32 * hcol, r);
                tmem_ld_wait();
                float s[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) s[j] = __uint_as_float(r[j]);
                const int klo = key_lo - b * AK - 32 * hcol, khi = key_hi - b * AK - 32 * hcol;   // this thread's keys [klo, khi) of its 32
                if (klo > 0 || khi < 32) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) if (j < klo || j >= khi) s[j] = -INFINITY;
                }
                float mx4[4] = {s[0], s[1], s[2], s[3]};
#pragma unroll
                for (int j = 4; j < 32; j += 4) {
                    mx4[0] = fmaxf(mx4[0], s[j]); mx4[1] = fmaxf(mx4[1], s[j + 1]);
                    mx4[2] = fmaxf(mx4[2], s[j + 2]); mx4[3] = fmaxf(mx4[3], s[j + 3]);
                }
                float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
                // the other half of the row lives in the quadrant's other warp
                float * slot = smem_mx + ((((gb & 1) * 4 + q) * 32 + lane) << 1);
                slot[hcol] = mx;
                named_bar_sync(bar_id, 64);                   // (also: both warps have read S before either overwrites it with P)
                // (a row whose sentence has no key in this block sees -inf only: the floor keeps its reference finite, its
                // exponentials come out as 0 and nothing has to be rescaled until its first real block)
                mx = fmaxf(fmaxf(mx, slot[hcol ^ 1]), -1e30f);
                float m_ref = (NH == 2 && hs) ? m1 : m0;
                // move the reference max only on the first block or when it is exceeded by > 2^8
                const bool need = mx > m_ref + kRescaleThreshold;     // always true on the first block (m_ref = -inf)
                const bool rescale = need && m_ref > -1e29f;          // the accumulator only holds something after a real block
                float factor = 1.f;
                if (need) { factor = ex2f2(m_ref - mx); m_ref = mx; }
                if (NH == 2 && hs) m1 = m_ref; else m0 = m_ref;
                if (b >= 1 && __any_sync(0xffffffffu, rescale)) {
                    // rare: this head's previous P V must be in the accumulator before it may be rescaled
                    const uint32_t gp = gb - NH;
                    mbar_wait(&pv_done[gp & 1], (gp >> 1) & 1);
                    tc_fence_after_sync();
                    // O (and the row sums next to it) of this head, split between the quadrant's two warps
                    const bool big = (D == 64) ? true : (hcol == 0);
                    const uint32_t addr = (D == 64) ? lane_base + TM_O + 32 * hcol
                                                    : (hcol == 0 ? lane_base + TM_O + 32 * hs : lane_base + TM_L + 16 * hs);
                    if (big) {
                        uint32_t o[32];
                        tmem_ld_32x32b_x32(addr, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
                        tmem_st_32x32b_x32(addr, o);
                    }
                    if ((D == 64 && hcol == 0) || (D == 32 && hcol == 1)) {
                        const uint32_t laddr = (D == 64) ? lane_base + TM_L : addr;
                        uint32_t o[16];
                        tmem_ld_32x32b_x16(laddr, o);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 16; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
                        tmem_st_32x32b_x16(laddr, o);
                    }
                }
                const float2 nm2 = make_float2(-m_ref, -m_ref), sc2 = make_float2(p.scale_log2, p.scale_log2);
                uint32_t pk[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 x = __ffma2_rn(make_float2(s[2 * j], s[2 * j + 1]), sc2, nm2);
                    const float2 e = (kPolyEvery2 > 0 && j % (kPolyEvery2 > 0 ? kPolyEvery2 : 1) == kPolyEvery2 - 1)
                                         ? ex2_poly2b(x) : make_float2(ex2f2(x.x), ex2f2(x.y));
                    pk[j] = pack_h2b(e.x, e.y);
                }
                tmem_st_32x32b_x16(s_addr + 16 * hcol, pk);   // P (packed fp16 pairs) over the first half of this S buffer
                tmem_st_wait();
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_ready[gb & 1]);
            }
            g += nvb;
            // ---- finalize: O / l, optional q8 fake-quant per 32-wide block, fp16 store of valid rows
            mbar_wait(&pv_done[(g - 1) & 1], ((g - 1) >> 1) & 1);
            tc_fence_after_sync();
            float l;
            {
                uint32_t lsum;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];\n\ttcgen05.wait::ld.sync.aligned;"
                             : "=r"(lsum) : "r"(lane_base + TM_L + (D == 32 ? 16 * hcol : 0)) : "memory");
                l = __uint_as_float(lsum);
            }
            uint32_t o0[32];
            tmem_ld_32x32b_x32(lane_base + TM_O + 32 * hcol, o0);     // head size 32: columns 32.. are the item's second head
            tmem_ld_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(o_free);
            const float inv = 1.f / l;
            const float2 inv2 = make_float2(inv, inv);
            const bool fq = p.fq_type == W_Q4_0 || p.fq_type == W_Q4_1;
            uint8_t * stg = smem_o + q * 4096;                  // the quadrant's [32 rows][128 B], 16-B chunks XOR-swizzled by row
            {
                float2 vv[16];
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    vv[j] = __fmul2_rn(make_float2(__uint_as_float(o0[2 * j]), __uint_as_float(o0[2 * j + 1])), inv2);
                if (fq) fake_quant32_v2(vv, p.fq_type);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 u;
                    u.x = pack_h2b(vv[4 * c].x, vv[4 * c].y); u.y = pack_h2b(vv[4 * c + 1].x, vv[4 * c + 1].y);
                    u.z = pack_h2b(vv[4 * c + 2].x, vv[4 * c + 2].y); u.w = pack_h2b(vv[4 * c + 3].x, vv[4 * c + 3].y);
                    *reinterpret_cast<uint4 *>(stg + lane * 128 + (((hcol * 4 + c) ^ (lane & 7)) << 4)) = u;
                }
            }
            named_bar_sync(bar_id, 64);
            // coalesced copy-out by the quadrant's 64 threads: one instruction covers eight whole 128-B rows; rows past the
            // sentence end are skipped
            {
                const int rows_valid = len - it.q0 - q * 32;       // of the quadrant's 32 rows
                __half * obase = p.out + (size_t) (it.t0 + it.q0 + q * 32) * p.H + it.grp * AW;
                const int t = hcol * 32 + lane;
                const int ch = t & 7, r0 = t >> 3;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int rr = r0 + 8 * i;
                    const uint4 u = *reinterpret_cast<const uint4 *>(stg + rr * 128 + ((ch ^ (rr & 7)) << 4));
                    if (rr < rows_valid) *reinterpret_cast<uint4 *>(obase + (size_t) rr * p.H + ch * 8) = u;
                }
            }
            named_bar_sync(bar_id, 64);                         // staging is reused by the next item
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, TM_COLS);
}

}  // namespace

int pack_spans(const int32_t * cu, int n_sent, int32_t * span_cu) {
    int n = 0, s = 0;
    span_cu[0] = n_sent > 0 ? cu[0] : 0;
    while (s < n_sent) {
        int e = s + 1;                                           // a sentence longer than the limit is a span of its own
        while (e < n_sent && cu[e + 1] - cu[s] <= AQ) ++e;
        span_cu[++n] = cu[e];
        s = e;
    }
    return n;
}

cudaError_t launch_attention_tc2(const CUtensorMap * tm_q, const CUtensorMap * tm_kv, const int32_t * cu_seqlens,
                                 int n_sent, int max_len, int n_head, int d_head, __half * out, int fq_type,
                                 int num_sms, cudaStream_t st, const int2 * tok_bounds) {
    if (n_sent <= 0) return cudaSuccess;
    if ((d_head != 64 && d_head != 32) || !tm_q || !tm_kv || (n_head * d_head) % AW != 0) return cudaErrorInvalidValue;
    Attn2Params p;
    p.cu = cu_seqlens; p.tok_bounds = tok_bounds; p.out = out; p.n_sent = n_sent; p.H = n_head * d_head; p.n_grp = p.H / AW;
    p.max_qt = (max_len + AQ - 1) / AQ; p.fq_type = fq_type;
    p.scale_log2 = 1.4426950408889634f / sqrtf((float) d_head);
    const long long total = (long long) n_sent * p.n_grp * p.max_qt;
    if (total > 0x7fffffffLL) return cudaErrorInvalidValue;
    const int grid = (int) (total < 2LL * num_sms ? total : 2LL * num_sms);
    if (d_head == 64) {
        cudaError_t e = cudaFuncSetAttribute(attention_tc2_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtt2Smem);
        if (e != cudaSuccess) return e;
        return launch_pdl(attention_tc2_kernel<64>, dim3(grid), dim3(ATT2_THREADS), kAtt2Smem, st, *tm_q, *tm_kv, p);
    }
    cudaError_t e = cudaFuncSetAttribute(attention_tc2_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, kAtt2Smem);
    if (e != cudaSuccess) return e;
    return launch_pdl(attention_tc2_kernel<32>, dim3(grid), dim3(ATT2_THREADS), kAtt2Smem, st, *tm_q, *tm_kv, p);
}

}  // namespace b200
