// K1 -- Y[M][N] = X[M][K] * W[N][K]^T (+bias)(+GELU)(+residual)(+q8 fake-quant), fp16 in,
// f32 accumulate in TMEM, fp16 out.  Replaces the reference's ggml_compute_forward_mul_mat
// (ggml.c:11030-11258) + the bias / GELU / residual graph nodes around it
// (bert.cpp:993-1016, 1040-1045, 1059-1071).
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0      TMA producer: A tile (128 x 64 fp16, 128B swizzle) and, for fp16 weights,
//               the B tile (BN x 64) per k-stage
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma kind::f16 (M=128, N=BN,
//               K=16) x4 per k-stage, accumulators double-buffered in TMEM
//   warps 2-5   epilogue: tcgen05.ld -> bias/GELU/residual/fake-quant -> fp16 -> global
//   warps 6-13  (q4 weights only) dequantise the BN x 64 weight stage straight from the
//               4-bit planes in HBM/L2 into the swizzled smem B tile, so the weights
//               cross HBM once at 4.5 / 5 bits (reference block formats ggml.c:893-907,
//               dequant rule ggml.c:1524-1563)
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

constexpr int BM = 128;          // rows per CTA; a CTA pair computes a 256 x BN tile
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int kSmemLimit = 232448;                   // 227 KB opt-in limit per CTA

// RES: the epilogue adds a residual (LN implies RES); without it the residual staging goes to the B ring
// EPB (fused LayerNorm, short K): the epilogue chain (residual load -> pass 1 -> statistics exchange -> pass 2 ->
// stores) is longer than the mainloop of a tile, so smem goes to the epilogue instead of the rings: the q8 twin
// gets its own staging buffer and the residual box of the NEXT tile is fetched while this tile is normalised.
// Build-time knobs kept from the round-2 bisect of the mainloop race (profiles/r2_gemm_race.md); the defaults are the
// shipped kernel, the other values exist so that the regression experiment can be repeated:
//   BERT_NIB_HANDBACK  when the dequant warps hand a 4-bit slot back to the bulk-copy warp:
//       2 (default)  at the end of the stage, after the slot's words have gone through the dequant arithmetic and the
//                    stores' MEMBAR -- the refill can then never land before a shared load has read the slot
//       0            right after ISSUING the shared loads (round 1).  This is the race: the refill of the slot (the
//                    stage 6 ahead) can land before a load has read it, and the warp dequantises the wrong stage.
//       3 / 4        after a fence.acq_rel.cta / after a named barrier of the warp group (both also close the race)
//   BERT_EPI_LDX       1 (default) both 32-column halves of an epilogue box are loaded from TMEM up front; 0: one at a time
//   BERT_DQ_SYNC       0 (default) __syncwarp + lane 0 publishes a dequantised B stage; 1: every thread arrives itself
//   BERT_GEMM_SA / BERT_GEMM_SB   ring depths of the plain q4 instantiations (timing perturbations for the tests)
#ifndef BERT_NIB_HANDBACK
#define BERT_NIB_HANDBACK 2
#endif
#ifndef BERT_EPI_LDX
#define BERT_EPI_LDX 1
#endif
#ifndef BERT_DQ_SYNC
#define BERT_DQ_SYNC 0
#endif
//   BERT_LN_RES2       0 (default) fused-LayerNorm epilogue: a warp's second residual box is requested into the residual buffer once
//                      the first has been read out of it; 1: it travels into the warp's output staging buffer (idle until
//                      pass 2) from the top of the tile.  The wait this removes was 10.6 % of the epilogue warps' samples in
//                      the attention-output GEMM, yet the kernel is no faster (0.4513 vs 0.4483 ms at 262 144 x 768 x 768,
//                      1.0994 vs 1.0917 ms at K = 3072, two alternations): the time moves to the statistics exchange
//                      (profiles/r2_rowwise_gemm_traffic.md)
#ifndef BERT_LN_RES2
#define BERT_LN_RES2 0
#endif
template <int BN, int WT, bool LN, bool RES, bool EPB = false> struct GemmCfg {
    static constexpr bool kQ4      = WT != W_F16;
    static constexpr int kBHalf    = BN / 2;                 // B rows this CTA supplies
    static constexpr int kABytes   = BM * BK * 2;
    static constexpr int kBBytes   = kBHalf * BK * 2;
    static constexpr int kNibBytes = BN * 16;                // 4-bit payload of one k-stage of this CTA's rows
    static constexpr int kScaleBytes = BN * 2;               // fp16 d (and as much again for m)
    static constexpr int kNibSlot  = kQ4 ? (kNibBytes + 2 * kScaleBytes + 127) / 128 * 128 : 0;
    static constexpr int kTmemCols = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
    static constexpr int kAccStride = BN <= 64 ? 64 : (BN <= 128 ? 128 : 256);
    static constexpr int kOutBytes = 8 * 4096;          // 8 epilogue warps x [32 rows][128 B] staging for TMA stores
    static constexpr int kResBytes = RES ? 8 * 4096 : 0; // same geometry, residual boxes prefetched by TMA
    static constexpr int kTwinBytes = (LN && EPB) ? 8 * 4096 : 0;   // EPB: staging of the q8 twin (otherwise it shares the residual buffer)
    static constexpr int kBiasBytes = LN ? 0 : 8 * 256; // per warp: the 64 bias values of its current box
    // fused LayerNorm: per-row partial statistics of up to 4 pairs x 2 warps, double-buffered, plus the
    // CTA's 256-column slices of bias / gamma / beta
    static constexpr int kLnMaxPairs = 4;              // cluster of at most 8 CTAs (the portable limit)
    static constexpr int kLnStatBytes = LN ? 2 * kLnMaxPairs * 2 * BM * 8 : 0;
    static constexpr int kLnParamBytes = LN ? 3 * BN * 4 : 0;
    static constexpr int kBarBytes = 512;
    static constexpr int kFixed = kOutBytes + kResBytes + kTwinBytes + kBiasBytes + kLnStatBytes + kLnParamBytes + 1024 /*align slack*/ + kBarBytes;
    // fp16 weights: A and B travel together through one ring.  q4 weights: A needs TMA-latency depth,
    // the dequantised B only a short ring (it is produced from smem), the 4-bit ring hides L2 latency.
    static constexpr int kFitF16   = (kSmemLimit - kFixed) / (kABytes + kBBytes);
#ifdef BERT_GEMM_SA
    static constexpr int kSA       = kQ4 ? (LN ? (EPB ? 3 : 4) : BERT_GEMM_SA) : (kFitF16 > 6 ? 6 : kFitF16);
#else
    static constexpr int kSA       = kQ4 ? (LN ? (EPB ? 3 : 4) : 5) : (kFitF16 > 6 ? 6 : kFitF16);
#endif
#ifdef BERT_GEMM_SB
    static constexpr int kSB       = kQ4 ? (RES ? (EPB ? 2 : 3) : BERT_GEMM_SB) : kSA;
#else
    static constexpr int kSB       = kQ4 ? (RES ? (EPB ? 2 : 3) : (BN > 192 ? 5 : 6)) : kSA;
#endif
    // (the 4-bit ring depth stays a multiple of the number of dequant groups, so a slot is always
    // refilled for the same group that drained it)
    static constexpr int kSN       = kQ4 ? 6 : 0;
    static constexpr int kSmemBytes = kSA * kABytes + kSB * kBBytes + kSN * kNibSlot + kFixed;
    static_assert(kSmemBytes <= kSmemLimit, "shared memory budget");
    static_assert(2 * kSA + 2 * kSB + 2 * kSN + 22 <= (kBarBytes - 8) / 8, "barrier region");
};

struct GemmParams {
    const uint4 * qnib; const __half * qd; const __half * qm;
    const float * bias; __half * out;
    const float * ln_gamma; const float * ln_beta;       // fused LayerNorm (EPI_LN)
    int M, N, K, epilogue, fq_type;
};

// Two floats -> packed fp16 pair (x in the low half), SATURATING to +-65504: the reference keeps f32 activations, so an
// outlier above the fp16 range must not become inf here (inf * 0 in a later masked P V product would poison a sentence)
__device__ __forceinline__ __half2 pack_h2_sat(float x, float y) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y), "f"(x));
    return *reinterpret_cast<__half2 *>(&r);
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
// round two floats to fp16 and back (what passing a value through an f16 table index does)
__device__ __forceinline__ float2 round_h2(float2 v) {
    return __half22float2(__floats2half2_rn(v.x, v.y));
}
// Reference GELU numerics: y = f16(gelu_f32(f16(x))) through a 64K-entry table (ggml.c:3561-3580, 4332),
// gelu(x) = 0.5 x (1 + tanh(sqrt(2/pi) x (1 + 0.044715 x^2))).  One MUFU.TANH per element (abs. error
// ~2^-11, below the table's own f16 output rounding for |x| < 8) and packed f32x2 arithmetic.
__device__ __forceinline__ float tanh_approx(float x) {
    float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y;
}
__device__ __forceinline__ float2 gelu2_ref_numerics(float2 v) {
    const float k1 = 0.79788456080286535588f, k2 = 0.79788456080286535588f * 0.044715f;
    const float2 f = round_h2(v);
    const float2 u = __fmul2_rn(f, __ffma2_rn(__fmul2_rn(f, f), make_float2(k2, k2), make_float2(k1, k1)));
    const float2 t = make_float2(tanh_approx(u.x), tanh_approx(u.y));
    const float2 h = __fmul2_rn(f, make_float2(0.5f, 0.5f));
    return round_h2(__ffma2_rn(h, t, h));
}
// The same function in packed half arithmetic (EPI_GELU_H2): one F2FP, five HMUL2/HFMA2 and one tanh.approx.f16x2 per
// PAIR instead of the f32x2 chain with two half round trips.  The input rounding to fp16 is the reference's; the
// intermediate fp16 roundings add ~1e-3 |x| of absolute error, an order of magnitude below the q8 step the
// result is quantised to next.
__device__ __forceinline__ float2 gelu2_half_arith(float2 v) {
    const __half2 f = __floats2half2_rn(v.x, v.y);
    const __half2 k1 = __float2half2_rn(0.79788456080286535588f), k2 = __float2half2_rn(0.79788456080286535588f * 0.044715f);
    const __half2 u = __hmul2(f, __hfma2(__hmul2(f, f), k2, k1));
    uint32_t ti; const uint32_t ui = *reinterpret_cast<const uint32_t *>(&u);
    asm("tanh.approx.f16x2 %0, %1;" : "=r"(ti) : "r"(ui));
    const __half2 t = *reinterpret_cast<const __half2 *>(&ti);
    const __half2 h = __hmul2(f, __float2half2_rn(0.5f));
    return __half22float2(__hfma2(h, t, h));
}
__device__ __forceinline__ void tma_store_2d(const void * tmap, const void * smem_src, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 :: "l"(tmap), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" :: "n"(N) : "memory");
}

constexpr int DQ_WARPS = 8;          // q4 dequant producer warps (two per SM sub-partition)
constexpr int EPI_WARPS = 8;         // epilogue warps (two per TMEM lane quadrant)

// two nibbles -> two fp16 weights.  `w` holds e(2i) at bits [4i,4i+4), e(2i+1) at [16+4i,..).
// The nibble is OR-ed into the mantissa of an fp16 constant whose ulp matches the nibble's
// bit position (1024 for mantissa bits 0-3, 64 for bits 4-7), so a word needs one shift for its
// four pairs; then (q - 8) * d  or  q * d + m  in half2.
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t mask, uint32_t bits) {
    uint32_t x;                              // (a & mask) | bits in one LOP3
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(x) : "r"(a), "r"(mask), "r"(bits));
    return x;
}
template <int WT>
__device__ __forceinline__ uint32_t dq_pair(uint32_t w, int i, __half2 d2, __half2 m2) {
    uint32_t x;
    unsigned short base;                 // fp16 bits of (magic + 8) for q4_0, magic for q4_1
    const uint32_t ws = i < 2 ? w : (w >> 8);                       // one shift serves pairs 2 and 3
    if ((i & 1) == 0) { x = and_or(ws, 0x000F000Fu, 0x64006400u); base = WT == W_Q4_0 ? 0x6408 : 0x6400; }   // 1024 + q
    else              { x = and_or(ws, 0x00F000F0u, 0x54005400u); base = WT == W_Q4_0 ? 0x5480 : 0x5400; }   //   64 + q
    __half2 h = __hsub2(*reinterpret_cast<__half2 *>(&x), __half2half2(__ushort_as_half(base)));
    h = WT == W_Q4_0 ? __hmul2(h, d2) : __hfma2(h, d2, m2);
    return *reinterpret_cast<uint32_t *>(&h);
}

// Tile walk shared by all roles.  Plain GEMM: cluster (= CTA pair) `cid` of `ncl` takes tiles cid,
// cid + ncl, ...; tile -> (m block of 256 rows, n block of BN columns), n fastest so that concurrently
// running pairs share A rows through L2.  Fused LayerNorm: a cluster holds N/BN pairs and takes whole
// 256-row stripes cid, cid + ncl, ...; pair `pair` of the cluster always computes n block `pair`.
struct TileWalk {
    int num_n, num_m, cid, ncl, pair;
    bool stripes;
    __device__ __forceinline__ int my_tiles() const {
        const int total = stripes ? num_m : num_m * num_n;
        return cid < total ? (total - cid + ncl - 1) / ncl : 0;
    }
    __device__ __forceinline__ void get(int i, int & m_blk, int & n_blk) const {
        const int t = cid + i * ncl;
        if (stripes) { m_blk = t; n_blk = pair; } else { m_blk = t / num_n; n_blk = t % num_n; }
    }
};

constexpr int DQ_GROUPS = 2;                       // dequant warp groups; group g takes k-stages g, g + 2, ...
constexpr int DQ_GROUP_WARPS = DQ_WARPS / DQ_GROUPS;
// Warp roles, aligned to warpgroups so that registers can be re-balanced with setmaxnreg:
//   warps 0-3   control: 0 TMA producer, 1 MMA issuer, 2 4-bit weight stream (q4), 3 idle
//   warps 4-11  epilogue (two warpgroups)
//   warps 12-19 q4 dequant (two warpgroups; absent for fp16 weights)
constexpr int CTRL_WARPS = 4;
template <int WT> constexpr int gemm_threads() { return 32 * (CTRL_WARPS + EPI_WARPS + (WT == W_F16 ? 0 : DQ_WARPS)); }
constexpr int kRegsCtrl = 40;                                            // q4: 640 threads launch with 96
// register pool (q4: 640 threads launch with 96): setmaxnreg.inc can only draw on what the CTA's own warps gave
// back -- control 128 x 56 + dequant 256 x 40 = 17408 -> the 256 epilogue threads may grow by 68, i.e. to 160
template <bool LN> struct RegBudget {
    static constexpr int kDequant = LN ? 56 : 64, kEpilogue = LN ? 160 : 152;
    static_assert(128 * (96 - kRegsCtrl) + 256 * (96 - kDequant) >= 256 * (kEpilogue - 96), "setmaxnreg pool");
};

// The eight epilogue warps' wait for a finished accumulator.  A warp suspended in mbarrier.try_wait is woken by every
// barrier event of the CTA (each TMA / bulk-copy stage that lands), re-polls and goes back to sleep: ncu counted 84 polls per
// wait and 21 % of the QKV GEMM's executed instructions in this one loop.  BERT_EPI_WAIT_LEADER=1 lets ONE warp watch the
// mbarrier while the other seven park on a hardware named barrier -- measured (profiles/r2_gemm_poll.md): no faster (QKV
// 0.3564 -> 0.3576 ms, FFN-in 0.543 -> 0.558 ms, bench 9 169 -> 9 090 emb/s), the polls sit in otherwise idle issue slots and
// the barrier couples the eight warps' start.  Off.
#ifndef BERT_EPI_WAIT_LEADER
#define BERT_EPI_WAIT_LEADER 0
#endif
__device__ __forceinline__ void wait_tfull(uint64_t * bar, uint32_t parity) {
#if BERT_EPI_WAIT_LEADER
    if ((threadIdx.x >> 5) == CTRL_WARPS) mbar_wait(bar, parity);
    asm volatile("bar.sync 1, %0;" :: "n"(32 * EPI_WARPS) : "memory");
#else
    mbar_wait(bar, parity);
#endif
}

// Launched with a runtime cluster dimension: 2 (one CTA pair) or, with fused LayerNorm, 2 * N / BN.
template <int BN, int WT, bool LN, bool RES, bool EPB>
__global__ void __launch_bounds__(gemm_threads<WT>(), 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
               const __grid_constant__ CUtensorMap tmap_out2, const GemmParams p) {
    static_assert(RES || !LN, "fused LayerNorm needs the residual path");
    static_assert(LN || !EPB, "EPB is a variant of the fused-LayerNorm epilogue");
    using Cfg = GemmCfg<BN, WT, LN, RES, EPB>;
    constexpr bool Q4 = Cfg::kQ4;
    constexpr int SA = Cfg::kSA, SB = Cfg::kSB, SN = Cfg::kSN;
    constexpr int NB = BN / 64;                               // 64-column output boxes per tile
    extern __shared__ uint8_t smem_raw[];
    // 1024-B alignment by an offset from the array (not an integer round trip), so that the compiler keeps the
    // shared address space and emits LDS/STS rather than generic LD/ST
    uint8_t * smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t * smem_a = smem;                                  // SA x 16 KB
    uint8_t * smem_b = smem_a + SA * Cfg::kABytes;            // SB x (BN/2)*128 B
    uint8_t * smem_out = smem_b + SB * Cfg::kBBytes;          // 8 warps x 4 KB staging for TMA stores
    uint8_t * smem_res = smem_out + Cfg::kOutBytes;           // 8 warps x 4 KB residual boxes
    uint8_t * smem_twin = smem_res + Cfg::kResBytes;          // EPB: 8 warps x 4 KB staging of the q8 twin
    uint8_t * smem_bias = smem_twin + Cfg::kTwinBytes;        // 8 warps x 64 floats
    uint8_t * smem_nib = smem_bias + Cfg::kBiasBytes;         // SN x 4-bit weight stages (q4 only)
    uint8_t * smem_lnstat = smem_nib + SN * Cfg::kNibSlot;    // LN: [2][pairs<=4][2][128] x (mean, M2)
    uint8_t * smem_lnpar = smem_lnstat + Cfg::kLnStatBytes;   // LN: bias | gamma | beta of this CTA's BN columns
    uint64_t * bars = (uint64_t *) (smem_lnpar + Cfg::kLnParamBytes);
    uint64_t * a_full = bars;                   // [SA]  leader's copy is live (A of both CTAs; + B for fp16 weights)
    uint64_t * a_empty = a_full + SA;           // [SA]  per CTA, multicast commit
    uint64_t * b_full = a_empty + SA;           // [SB]  q4: leader's copy, dequant warps of both CTAs arrive
    uint64_t * b_empty = b_full + SB;           // [SB]  q4: per CTA, multicast commit
    uint64_t * n_full = b_empty + SB;           // [SN]  q4: per CTA, bulk-copy bytes
    uint64_t * n_empty = n_full + SN;           // [SN]  q4: per CTA, the dequant group that read the slot
    uint64_t * tfull = n_empty + SN;            // [2]   per CTA, multicast commit
    uint64_t * tempty = tfull + 2;              // [2]   leader's copy is live: 16 epilogue warps arrive
    uint64_t * rfull = tempty + 2;              // [8] residual box, per epilogue warp
    uint64_t * ln_bar = rfull + 8;              // [2] LN: partial statistics of all pairs have landed (by tile parity)
    uint64_t * rfull2 = ln_bar + 2;             // [8] LN, 256-wide tiles: the warp's second residual box (lands in its output staging)
    uint32_t * tmem_slot = (uint32_t *) (rfull2 + 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();                 // rank in the cluster
    const uint32_t rank = crank & 1;                          // rank in the CTA pair, 0 = leader (issues the MMAs)
    const uint32_t leader = crank & ~1u;                      // cluster rank of this pair's leader
    const uint16_t pair_mask = (uint16_t) (3u << leader);
    TileWalk tw;
    tw.num_n = p.N / BN;
    tw.num_m = (p.M + 2 * BM - 1) / (2 * BM);
    tw.cid = (int) cluster_id_x();
    tw.ncl = (int) cluster_nctaid_x();
    tw.pair = (int) (crank >> 1);
    tw.stripes = LN;
    const int my_tiles = tw.my_tiles();
    const int num_k = p.K / BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        if (!Q4) tma_prefetch_desc(&tmap_b);
        tma_prefetch_desc(&tmap_out);
        if (RES) tma_prefetch_desc(&tmap_res);
        for (int s = 0; s < SA; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
        if (Q4) {
            for (int s = 0; s < SB; ++s) { mbar_init(&b_full[s], 2 * DQ_GROUP_WARPS * (BERT_DQ_SYNC == 1 ? 32 : 1)); mbar_init(&b_empty[s], 1); }
            for (int s = 0; s < SN; ++s) { mbar_init(&n_full[s], 1); mbar_init(&n_empty[s], DQ_GROUP_WARPS); }
        }
        for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 2 * EPI_WARPS); }
        for (int i = 0; i < 8; ++i) { mbar_init(&rfull[i], 1); mbar_init(&rfull2[i], 1); }
        if (LN) { mbar_init(&ln_bar[0], 1); mbar_init(&ln_bar[1], 1); }
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc_pair(tmem_slot, Cfg::kTmemCols); tmem_relinquish_pair(); }
    tc_fence_before_sync();
    __syncthreads();
    cluster_sync_all();                     // the peer's barriers exist before anything signals them
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();                             // everything above overlapped the tail of the previous kernel

    if (warp < CTRL_WARPS) {
      if constexpr (Q4) reg_dealloc<kRegsCtrl>();
      if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (both CTAs)
        if (lane == 0) {
            const uint32_t full_leader = mapa_u32(smem_u32(&a_full[0]), leader);
            int s = 0; uint32_t ph = 0;
            for (int i = 0; i < my_tiles; ++i) {
                int mb, nb; tw.get(i, mb, nb);
                const int m0 = mb * (2 * BM) + (int) rank * BM;
                const int nb0 = nb * BN + (int) rank * Cfg::kBHalf;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(&a_empty[s], ph ^ 1);
                    if (rank == 0) mbar_expect_tx(&a_full[s], 2 * (Q4 ? Cfg::kABytes : Cfg::kABytes + Cfg::kBBytes));
                    tma_load_2d_pair(smem_a + s * Cfg::kABytes, &tmap_a, full_leader + 8 * s, kb * BK, m0);
                    if (!Q4) tma_load_2d_pair(smem_b + s * Cfg::kBBytes, &tmap_b, full_leader + 8 * s, kb * BK, nb0);
                    if (++s == SA) { s = 0; ph ^= 1; }
                }
            }
        }
      } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (leader CTA only)
        if (rank == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(2 * BM, BN);
            int sa = 0, sb = 0; uint32_t pha = 0, phb = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const int acc = it & 1;
                const uint32_t acc_ph = (it >> 1) & 1;
                mbar_wait(&tempty[acc], acc_ph ^ 1);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + acc * Cfg::kAccStride;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(&a_full[sa], pha);
                    if (Q4) mbar_wait(&b_full[sb], phb);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t adesc = umma_smem_desc_sw128(smem_u32(smem_a + sa * Cfg::kABytes));
                        const uint64_t bdesc = umma_smem_desc_sw128(smem_u32(smem_b + sb * Cfg::kBBytes));
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) {
                            // +32 B per K=16 step inside the 128-B swizzle row (address field is >>4)
                            umma_f16_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                        }
                        umma_commit_pair(&a_empty[sa], pair_mask);                       // frees the stage in both CTAs
                        if (Q4) umma_commit_pair(&b_empty[sb], pair_mask);
                        if (kb == num_k - 1) umma_commit_pair(&tfull[acc], pair_mask);   // accumulators ready in both CTAs
                    }
                    __syncwarp();
                    if (++sa == SA) { sa = 0; pha ^= 1; }
                    if (++sb == SB) { sb = 0; phb ^= 1; }
                }
            }
        }
      } else if (warp == 2) {
        // ------------------------------------------------------------ 4-bit weight stream (both CTAs)
        // One thread bulk-copies each k-stage of this CTA's rows (nibbles | d | m planes, each
        // contiguous in HBM) into the smem ring, SN stages ahead of the dequant warps.
        if constexpr (Q4) if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            constexpr uint32_t kBytes = Cfg::kNibBytes + (WT == W_Q4_1 ? 2 : 1) * Cfg::kScaleBytes;
            for (int i = 0; i < my_tiles; ++i) {
                int mb, nb; tw.get(i, mb, nb);
                const int nb0 = nb * BN + (int) rank * Cfg::kBHalf;
                for (int kb = 0; kb < num_k; ++kb) {
                    const size_t blk0 = ((size_t) kb * p.N + nb0) * 2;      // stage kb of rows nb0.. : [kb][n][2] blocks
                    uint8_t * slot = smem_nib + s * Cfg::kNibSlot;
                    mbar_wait(&n_empty[s], ph ^ 1);
                    mbar_expect_tx(&n_full[s], kBytes);
                    bulk_load_1d(slot, p.qnib + blk0, Cfg::kNibBytes, &n_full[s]);
                    bulk_load_1d(slot + Cfg::kNibBytes, p.qd + blk0, Cfg::kScaleBytes, &n_full[s]);
                    if (WT == W_Q4_1) bulk_load_1d(slot + Cfg::kNibBytes + Cfg::kScaleBytes, p.qm + blk0, Cfg::kScaleBytes, &n_full[s]);
                    if (++s == SN) { s = 0; ph ^= 1; }
                }
            }
        }
      }
    } else if (warp < CTRL_WARPS + EPI_WARPS) {
      if constexpr (Q4) reg_alloc<RegBudget<LN>::kEpilogue>();
      if constexpr (LN) {
        // ------------------------------------------------------------ epilogue with fused LayerNorm
        // out = LN(acc + bias + residual) * gamma + beta, and its q8 fake-quantised twin (out2).
        // The cluster's pairs hold the N/BN column blocks of one 256-row stripe.  Pass 1: each thread
        // (one row, 64 * NBW columns) pulls its accumulators, adds bias + residual, rounds to fp16 (what
        // the unfused path stores as y) and keeps the values in registers; the accumulator is released.
        // The row's (mean, M2) partials go to every CTA holding the same rows through DSMEM; after the
        // cluster-scope barrier they are merged (Chan) and pass 2 normalises out of registers.
        static_assert(NB % 2 == 0, "fused LayerNorm needs an even number of boxes per tile");
        constexpr int NBW = NB / 2;                // boxes per warp per tile
        const int e = warp - CTRL_WARPS, q = warp & 3, half = e >> 2;
        const int row = q * 32 + lane;             // row inside this CTA's 128
        uint8_t * st = smem_out + e * 4096;
        uint8_t * rs = smem_res + e * 4096;
        uint8_t * tws = EPB ? smem_twin + e * 4096 : rs;      // twin staging (shares the residual buffer unless EPB)
        uint64_t * rbar = rfull + e;
        const float * s_bias = reinterpret_cast<const float *>(smem_lnpar);
        const float * s_gamma = s_bias + BN;
        const float * s_beta = s_gamma + BN;
        const uint32_t tempty_leader = mapa_u32(smem_u32(&tempty[0]), leader);
        const int npairs = tw.num_n;
        const bool fq_on = (p.epilogue & EPI_FAKEQUANT) != 0;
        // this CTA's column block is fixed (n block = pair): stage bias / gamma / beta once
        {
            const int t = threadIdx.x - 32 * CTRL_WARPS;                    // 0..255
            float * par = reinterpret_cast<float *>(smem_lnpar);
            for (int i = t; i < BN; i += 32 * EPI_WARPS) {
                par[i] = __ldg(p.bias + tw.pair * BN + i);
                par[BN + i] = __ldg(p.ln_gamma + tw.pair * BN + i);
                par[2 * BN + i] = __ldg(p.ln_beta + tw.pair * BN + i);
            }
            asm volatile("bar.sync 1, %0;" :: "n"(32 * EPI_WARPS) : "memory");
        }
        auto res_load = [&](int it_) {
            // first box of tile it_ -> residual buffer (lane 0 only)
            int mb, nb; tw.get(it_, mb, nb);
            mbar_expect_tx(rbar, 4096);
            tma_load_2d(rs, &tmap_res, rbar, nb * BN + half * 64, mb * (2 * BM) + (int) rank * BM + q * 32);
        };
        if (my_tiles > 0 && lane == 0) res_load(0);
        uint32_t rph = 0;                          // residual barrier phase
        constexpr bool RES2 = BERT_LN_RES2 != 0 && NBW == 2;
        uint64_t * rbar2 = rfull2 + e;
        uint32_t rph2 = 0;
        for (int it = 0; it < my_tiles; ++it) {
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            int mb, nb; tw.get(it, mb, nb);
            const int wrow0 = mb * (2 * BM) + (int) rank * BM + q * 32;
            if (RES2 && lane == 0) {
                // the second residual box of this tile -> the output staging buffer, which pass 1 does not use (the previous
                // tile's stores must have finished reading it); it travels while the accumulator and the first box are waited for
                tma_store_wait_read<0>();
                mbar_expect_tx(rbar2, 4096);
                tma_load_2d(st, &tmap_res, rbar2, nb * BN + (half + 2) * 64, wrow0);
            }
            wait_tfull(&tfull[acc], acc_ph);
            tc_fence_after_sync();
            uint32_t hv[NBW][32];                  // pre-LN values as fp16 pairs
#pragma unroll
            for (int bi = 0; bi < NBW; ++bi) {
                const int b = half + 2 * bi;
                uint4 rr[8];
                const uint8_t * rsrc = (RES2 && bi == 1) ? st : rs;
                if (RES2 && bi == 1) { mbar_wait(rbar2, rph2); rph2 ^= 1; }
                else { mbar_wait(rbar, rph); rph ^= 1; }
#pragma unroll
                for (int c = 0; c < 8; ++c) rr[c] = *reinterpret_cast<const uint4 *>(rsrc + lane * 128 + ((c ^ (lane & 7)) << 4));
                fence_loads_done();                           // before the buffer goes back to the TMA (see the plain epilogue)
                __syncwarp();
                if (!RES2 && bi + 1 < NBW && lane == 0) {     // this tile's next box
                    mbar_expect_tx(rbar, 4096);
                    tma_load_2d(rs, &tmap_res, rbar, nb * BN + (b + 2) * 64, wrow0);
                }
                // EPB: the residual buffer is not the twin's staging, so the next tile's first box can travel
                // while this tile is normalised and stored (RES2: as soon as this tile's first box has been read out)
                if (EPB && bi + 1 == (RES2 ? 1 : NBW) && it + 1 < my_tiles && lane == 0) res_load(it + 1);
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t) (q * 32) << 16) + acc * Cfg::kAccStride + b * 64 + hh * 32, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 bb = *reinterpret_cast<const float4 *>(s_bias + b * 64 + hh * 32 + 4 * i);
                        const __half2 * h = reinterpret_cast<const __half2 *>(&rr[hh * 4 + (i >> 1)]);
                        const float2 r0 = __half22float2(h[(i & 1) * 2]), r1 = __half22float2(h[(i & 1) * 2 + 1]);
                        float2 v0 = make_float2(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]));
                        float2 v1 = make_float2(__uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
                        v0 = __fadd2_rn(__fadd2_rn(v0, make_float2(bb.x, bb.y)), r0);
                        v1 = __fadd2_rn(__fadd2_rn(v1, make_float2(bb.z, bb.w)), r1);
                        __half2 a0 = pack_h2_sat(v0.x, v0.y), a1 = pack_h2_sat(v1.x, v1.y);
                        hv[bi][hh * 16 + 2 * i] = *reinterpret_cast<uint32_t *>(&a0);
                        hv[bi][hh * 16 + 2 * i + 1] = *reinterpret_cast<uint32_t *>(&a1);
                    }
                }
            }
            // accumulator consumed: hand it back to the MMA issuer before the statistics exchange
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(tempty_leader + 8 * acc);
            // ---- partial statistics over this thread's 64 * NBW values (two-pass, from the fp16 values)
            float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
            for (int bi = 0; bi < NBW; ++bi) {
#pragma unroll
                for (int i = 0; i < 32; ++i) s2 = __fadd2_rn(s2, __half22float2(*reinterpret_cast<const __half2 *>(&hv[bi][i])));
            }
            const float mean_p = (s2.x + s2.y) * (1.0f / (64 * NBW));
            float2 m2 = make_float2(0.f, 0.f);
            const float2 nmp = make_float2(-mean_p, -mean_p);
#pragma unroll
            for (int bi = 0; bi < NBW; ++bi) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float2 dlt = __fadd2_rn(__half22float2(*reinterpret_cast<const __half2 *>(&hv[bi][i])), nmp);
                    m2 = __ffma2_rn(dlt, dlt, m2);
                }
            }
            const float M2_p = m2.x + m2.y;
            // slot [parity][pair][half][row] in every CTA of the cluster with my pair rank
            const int par = it & 1;
            const uint32_t slot = smem_u32(smem_lnstat) + (uint32_t) ((((par * Cfg::kLnMaxPairs + tw.pair) * 2 + half) * BM + row) * 8);
            // st.async: the 8 bytes are counted on the destination CTA's barrier of this tile parity (no fence, no
            // arrive; the barrier's single arrival is the destination's own expect_tx below)
            const uint32_t lnb = smem_u32(&ln_bar[par]);
            for (int pj = 0; pj < npairs; ++pj) {
                const uint32_t dst = (uint32_t) (2 * pj) + rank;
                st_async_f32x2(mapa_u32(slot, dst), mean_p, M2_p, mapa_u32(lnb, dst));
            }
            if (e == 0 && lane == 0) mbar_expect_tx(&ln_bar[par], (uint32_t) npairs * EPI_WARPS * 32 * 8);
            mbar_wait(&ln_bar[par], (uint32_t) (it >> 1) & 1);
            float mean = 0.f;
            float2 part[2 * Cfg::kLnMaxPairs];
            const int nparts = 2 * npairs;
#pragma unroll
            for (int k = 0; k < 2 * Cfg::kLnMaxPairs; ++k) {
                if (k < nparts) {
                    part[k] = *reinterpret_cast<const float2 *>(smem_lnstat + (((par * Cfg::kLnMaxPairs + (k >> 1)) * 2 + (k & 1)) * BM + row) * 8);
                    mean += part[k].x;
                }
            }
            mean /= (float) nparts;
            float M2 = 0.f;
#pragma unroll
            for (int k = 0; k < 2 * Cfg::kLnMaxPairs; ++k) {
                if (k < nparts) { const float dm = part[k].x - mean; M2 += part[k].y + (float) (64 * NBW) * dm * dm; }
            }
            const float rstd = 1.0f / sqrtf(M2 / (float) p.N + 1e-5f);
            const float2 nmean = make_float2(-mean, -mean), rstd2 = make_float2(rstd, rstd);
            // ---- pass 2: normalise out of registers, store out (st) and its fake-quantised twin (rs)
#pragma unroll
            for (int bi = 0; bi < NBW; ++bi) {
                const int b = half + 2 * bi;
                const int c0 = nb * BN + b * 64;
                if (lane == 0) tma_store_wait_read<0>();       // previous stores are done with st / rs
                __syncwarp();
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    float2 v[16];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float4 g = *reinterpret_cast<const float4 *>(s_gamma + b * 64 + hh * 32 + 4 * i);
                        const float4 be = *reinterpret_cast<const float4 *>(s_beta + b * 64 + hh * 32 + 4 * i);
                        const float2 x0 = __half22float2(*reinterpret_cast<const __half2 *>(&hv[bi][hh * 16 + 2 * i]));
                        const float2 x1 = __half22float2(*reinterpret_cast<const __half2 *>(&hv[bi][hh * 16 + 2 * i + 1]));
                        const float2 n0 = __fmul2_rn(__fadd2_rn(x0, nmean), rstd2), n1 = __fmul2_rn(__fadd2_rn(x1, nmean), rstd2);
                        v[2 * i] = __ffma2_rn(n0, make_float2(g.x, g.y), make_float2(be.x, be.y));
                        v[2 * i + 1] = __ffma2_rn(n1, make_float2(g.z, g.w), make_float2(be.z, be.w));
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int c = hh * 4 + i;
                        uint4 u;
                        __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
                        for (int k = 0; k < 4; ++k) h[k] = pack_h2_sat(v[i * 4 + k].x, v[i * 4 + k].y);
                        *reinterpret_cast<uint4 *>(st + lane * 128 + ((c ^ (lane & 7)) << 4)) = u;
                    }
                    if (fq_on) {
                        fake_quant32_v2(v, p.fq_type);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int c = hh * 4 + i;
                            uint4 u;
                            __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
                            for (int k = 0; k < 4; ++k) h[k] = pack_h2_sat(v[i * 4 + k].x, v[i * 4 + k].y);
                            *reinterpret_cast<uint4 *>(tws + lane * 128 + ((c ^ (lane & 7)) << 4)) = u;
                        }
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tmap_out, st, c0, wrow0);
                    if (fq_on) tma_store_2d(&tmap_out2, tws, c0, wrow0);
                    tma_store_commit();
                }
            }
            // rs doubles as the twin's staging buffer: the next tile's residual may only land once the
            // last store has finished reading it
            if (!EPB && it + 1 < my_tiles && lane == 0) { tma_store_wait_read<0>(); res_load(it + 1); }
        }
        if (lane == 0) tma_store_wait_all<0>();
      } else {
        // ------------------------------------------------------------ epilogue (8 warps per CTA)
        // Two warps per TMEM lane quadrant take alternate 64-column boxes of the tile, so every
        // scheduler has two epilogue warps to interleave.  Per box: TMEM -> registers ; bias (from a
        // warp-private smem slot filled one box ahead) / GELU / + residual (own row of the box the
        // warp prefetched by TMA) / q8 fake-quant ; fp16 -> swizzled smem ; one TMA store of the
        // warp's [32 rows x 64 cols] sub-box.  Global traffic is whole 128-B rows in both
        // directions; everything is warp-local (no CTA barrier).
        const int e = warp - CTRL_WARPS;           // 0..7
        const int q = warp & 3;                    // TMEM lane quadrant this warp may access
        const int half = e >> 2;                   // which of the quadrant's two warps
        constexpr int NBW0 = (NB + 1) / 2, NBW1 = NB / 2;
        const int nbw = half == 0 ? NBW0 : NBW1;   // boxes of a tile this warp owns: half, half + 2, ...
        uint8_t * st = smem_out + e * 4096;
        uint8_t * rs = smem_res + e * 4096;
        float * bslot = reinterpret_cast<float *>(smem_bias + e * 256);
        uint64_t * rbar = rfull + e;
        const uint32_t tempty_leader = mapa_u32(smem_u32(&tempty[0]), leader);
        const bool has_res = RES;
        const bool has_bias = (p.epilogue & EPI_BIAS) != 0;
        const int total_boxes = my_tiles * nbw;
        auto box_coords = [&](int j, int & c0, int & row0, int & b) {
            int mb, nb; tw.get(j / nbw, mb, nb);
            b = half + 2 * (j % nbw);
            c0 = nb * BN + b * 64;
            row0 = mb * (2 * BM) + (int) rank * BM + q * 32;
        };
        float2 bnext = make_float2(0.f, 0.f);
        if (total_boxes > 0) {
            int c0, row0, b; box_coords(0, c0, row0, b);
            if (has_res && lane == 0) {
                mbar_expect_tx(rbar, 4096);
                tma_load_2d(rs, &tmap_res, rbar, c0, row0);
            }
            if (has_bias) bnext = __ldg(reinterpret_cast<const float2 *>(p.bias + c0) + lane);
        }
        int j = 0;
        for (int it = 0; it < my_tiles; ++it) {
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            wait_tfull(&tfull[acc], acc_ph);
            tc_fence_after_sync();
#pragma unroll 1
            for (int bi = 0; bi < nbw; ++bi, ++j) {
                int c0, wrow0, b; box_coords(j, c0, wrow0, b);
                // this box's bias -> smem slot (read back as broadcast LDS.128); next box's bias -> registers
                reinterpret_cast<float2 *>(bslot)[lane] = bnext;
                uint4 rr[8];
                if (has_res) {
                    mbar_wait(rbar, j & 1);
#pragma unroll
                    for (int c = 0; c < 8; ++c) rr[c] = *reinterpret_cast<const uint4 *>(rs + lane * 128 + ((c ^ (lane & 7)) << 4));
                    // the shared loads must have COMPLETED (not just issued) before the buffer is handed back to the TMA
                    // (a register "sink" does not do it: ptxas removes a mov whose result is dead)
                    fence_loads_done();
                }
                __syncwarp();
                if (j + 1 < total_boxes) {
                    int c1, r1, b1; box_coords(j + 1, c1, r1, b1);
                    if (has_res && lane == 0) {                   // prefetch the next box's residual
                        mbar_expect_tx(rbar, 4096);
                        tma_load_2d(rs, &tmap_res, rbar, c1, r1);
                    }
                    if (has_bias) bnext = __ldg(reinterpret_cast<const float2 *>(p.bias + c1) + lane);
                }
                const uint32_t acc_addr = tmem_base + ((uint32_t) (q * 32) << 16) + acc * Cfg::kAccStride + b * 64;
#if BERT_EPI_LDX
                // both 32-column halves of the box in flight at once, one wait (1 % faster than one half at a time)
                uint32_t r2[2][32];
                tmem_ld_32x32b_x32(acc_addr, r2[0]);
                tmem_ld_32x32b_x32(acc_addr + 32, r2[1]);
                tmem_ld_wait();
#endif
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
#if BERT_EPI_LDX
                    const uint32_t * r = r2[hh];
#else
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(acc_addr + hh * 32, r);
                    tmem_ld_wait();
#endif
                    float2 v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
                    if (has_bias) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 bb = *reinterpret_cast<const float4 *>(bslot + hh * 32 + 4 * i);
                            v[2 * i] = __fadd2_rn(v[2 * i], make_float2(bb.x, bb.y));
                            v[2 * i + 1] = __fadd2_rn(v[2 * i + 1], make_float2(bb.z, bb.w));
                        }
                    }
                    if (p.epilogue & EPI_GELU_H2) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = gelu2_half_arith(v[i]);
                    } else if (p.epilogue & EPI_GELU) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = gelu2_ref_numerics(v[i]);
                    }
                    if (has_res) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const __half2 * h = reinterpret_cast<const __half2 *>(&rr[hh * 4 + i]);
#pragma unroll
                            for (int k = 0; k < 4; ++k) v[i * 4 + k] = __fadd2_rn(v[i * 4 + k], __half22float2(h[k]));
                        }
                    }
                    if (p.epilogue & EPI_FAKEQUANT) fake_quant32_v2(v, p.fq_type);
                    if (hh == 0) {
                        // the TMA store of this warp's previous box must be done reading the staging buffer
                        if (lane == 0) tma_store_wait_read<0>();
                        __syncwarp();
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int c = hh * 4 + i;
                        uint4 u;
                        __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
                        for (int k = 0; k < 4; ++k) h[k] = pack_h2_sat(v[i * 4 + k].x, v[i * 4 + k].y);
                        *reinterpret_cast<uint4 *>(st + lane * 128 + ((c ^ (lane & 7)) << 4)) = u;
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    tma_store_2d(&tmap_out, st, c0, wrow0);
                    tma_store_commit();
                }
            }
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(tempty_leader + 8 * acc);
        }
        if (lane == 0) tma_store_wait_all<0>();
      }
    } else {
      if constexpr (Q4) reg_dealloc<RegBudget<LN>::kDequant>();
        // ------------------------------------------------------------ q4 dequant (both CTAs)
        // Each CTA dequantises the BN/2 weight rows it contributes to the pair's B operand, reading
        // the 4-bit stage from the smem ring the bulk-copy warp fills (no global loads here, so the
        // proxy fence never waits on HBM/L2).  Two warp groups take alternate k-stages.
        if constexpr (Q4) {
            constexpr int GT = 32 * DQ_GROUP_WARPS;              // threads per group
            constexpr int NUNIT = BN;                            // 32-weight blocks per CTA per stage
            constexpr int UNITS = (NUNIT + GT - 1) / GT;         // per thread
            const int grp = (warp - CTRL_WARPS - EPI_WARPS) / DQ_GROUP_WARPS;
            const int tg = threadIdx.x - 32 * (CTRL_WARPS + EPI_WARPS) - grp * GT;         // 0..GT-1
            const uint32_t bfull_leader = mapa_u32(smem_u32(&b_full[0]), leader);
            const int total = my_tiles * num_k;
            int sn = grp % SN, sb = grp % SB;
            uint32_t phn = 0, phb = 0;
            for (int i = grp; i < total; i += DQ_GROUPS) {
                const uint8_t * slot = smem_nib + sn * Cfg::kNibSlot;
                mbar_wait(&n_full[sn], phn);
                uint4 nib[UNITS]; __half d[UNITS], m[UNITS];
#pragma unroll
                for (int u = 0; u < UNITS; ++u) {
                    const int unit = tg + GT * u;
                    if (unit < NUNIT) {
                        nib[u] = *reinterpret_cast<const uint4 *>(slot + unit * 16);
                        d[u] = *reinterpret_cast<const __half *>(slot + Cfg::kNibBytes + unit * 2);
                        m[u] = WT == W_Q4_1 ? *reinterpret_cast<const __half *>(slot + Cfg::kNibBytes + Cfg::kScaleBytes + unit * 2) : __half();
                    }
                }
                // Handing the 4-bit slot back: NOT here.  The loads above are only issued at this point; arriving on n_empty now
                // lets the bulk-copy warp refill the slot (with the stage SN ahead) before a load has read it -- seen on B200
                // as 16-column groups of a tile carrying the weights of the stage 6 ahead (profiles/r2_gemm_race.md).
#if BERT_NIB_HANDBACK == 3
                fence_loads_done();
#elif BERT_NIB_HANDBACK == 4
                named_bar_sync(2 + grp, GT);
#endif
#if BERT_NIB_HANDBACK != 2
                __syncwarp();
                if (lane == 0) mbar_arrive(&n_empty[sn]);
#endif
                // dequantise into registers first (one 32-weight unit at a time: the register budget of these warps is small):
                // only the stores of the first batch sit between "slot free" and "slot full"
                constexpr int UB = 1;
                uint8_t * bt = smem_b + sb * Cfg::kBBytes;
#pragma unroll
                for (int u0 = 0; u0 < UNITS; u0 += UB) {
                    uint4 o[UB][4];
#pragma unroll
                    for (int v = 0; v < UB; ++v) {
                        const int u = u0 + v;
                        const __half2 d2 = __half2half2(d[u]), m2 = __half2half2(m[u]);
                        const uint32_t w[4] = {nib[u].x, nib[u].y, nib[u].z, nib[u].w};
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            o[v][c].x = dq_pair<WT>(w[c], 0, d2, m2);
                            o[v][c].y = dq_pair<WT>(w[c], 1, d2, m2);
                            o[v][c].z = dq_pair<WT>(w[c], 2, d2, m2);
                            o[v][c].w = dq_pair<WT>(w[c], 3, d2, m2);
                        }
                    }
                    if (u0 == 0) mbar_wait(&b_empty[sb], phb ^ 1);
#pragma unroll
                    for (int v = 0; v < UB; ++v) {
                        const int unit = tg + GT * (u0 + v);
                        if (unit >= NUNIT) continue;
                        const int r = unit >> 1, half_k = unit & 1;          // row in this CTA's half, which 32-block
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            const int chunk = half_k * 4 + c;
                            *reinterpret_cast<uint4 *>(bt + r * 128 + ((chunk ^ (r & 7)) << 4)) = o[v][c];
                        }
                    }
                }
                fence_proxy_async_smem();
#if BERT_DQ_SYNC == 1
                mbar_arrive_cluster(bfull_leader + 8 * sb);          // every thread publishes its own stores
#else
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(bfull_leader + 8 * sb);
#endif
#if BERT_NIB_HANDBACK == 2
                // every word loaded from the slot has gone through the dequant arithmetic into the stores above, and the
                // proxy fence's MEMBAR has completed: the slot is certainly read, hand it back
                if (lane == 0) mbar_arrive(&n_empty[sn]);
#endif
                sn += DQ_GROUPS; if (sn >= SN) { sn -= SN; phn ^= 1; }
                sb += DQ_GROUPS; if (sb >= SB) { sb -= SB; phb ^= 1; }
            }
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    cluster_sync_all();                     // nobody leaves while the peer may still touch this CTA
    if (warp == 1) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
}

// Persistent grid: as many clusters as the device can keep resident (cached per device and cluster size).
template <int BN, int WT, bool LN, bool RES, bool EPB>
static int max_clusters(int device, int num_sms, int csize) {
    static int cached[64][9] = {};
    if (device >= 0 && device < 64 && cached[device][csize] > 0) return cached[device][csize];
    using Cfg = GemmCfg<BN, WT, LN, RES, EPB>;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) (num_sms / csize * csize));
    cfg.blockDim = dim3(gemm_threads<WT>());
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned) csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_kernel<BN, WT, LN, RES, EPB>, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = csize == 2 ? num_sms / 2 : 1;
    }
    if (n > num_sms / csize) n = num_sms / csize;
    if (device >= 0 && device < 64) cached[device][csize] = n;
    return n;
}

template <int BN, int WT, bool LN, bool RES, bool EPB = false>
static cudaError_t launch_one(const GemmArgs & g, int num_sms, cudaStream_t st) {
    using Cfg = GemmCfg<BN, WT, LN, RES, EPB>;
    static_assert(Cfg::kSA >= 3, "pipeline too shallow");
    auto kern = gemm_tc_kernel<BN, WT, LN, RES, EPB>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return e;
    int device = 0;
    cudaGetDevice(&device);
    const DevWeight & w = *g.w;
    GemmParams p{w.qnib, w.qd, w.qm, w.bias, g.out, g.ln_gamma, g.ln_beta, g.M, w.N, w.K, g.epilogue, g.fq_type};
    const int num_m = (g.M + 2 * BM - 1) / (2 * BM), num_n = w.N / BN;
    const int csize = LN ? 2 * num_n : 2;                 // CTAs per cluster
    const int units = LN ? num_m : num_m * num_n;         // stripes or tiles to hand out
    int ncl_max = max_clusters<BN, WT, LN, RES, EPB>(device, num_sms, csize);
    // BERT_GEMM_MAX_CLUSTERS (debug): fewer persistent clusters -> many tiles per cluster at small M (sanitizer runs)
    static const int ncl_cap = [] { const char * e = getenv("BERT_GEMM_MAX_CLUSTERS"); return e ? atoi(e) : 0; }();
    if (ncl_cap > 0 && ncl_max > ncl_cap) ncl_max = ncl_cap;
    const int ncl = units < ncl_max ? units : ncl_max;
    if (ncl <= 0) return cudaSuccess;
    if ((g.epilogue & EPI_RESIDUAL) && !g.tmap_res) return cudaErrorInvalidValue;
    if (LN && (!g.ln_gamma || !g.ln_beta || !g.tmap_res || !w.bias || ((g.epilogue & EPI_FAKEQUANT) && !g.tmap_out2))) return cudaErrorInvalidValue;
    const CUtensorMap * tres = g.tmap_res ? g.tmap_res : g.tmap_out;
    const CUtensorMap * tout2 = g.tmap_out2 ? g.tmap_out2 : g.tmap_out;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) (csize * ncl));
    cfg.blockDim = dim3(gemm_threads<WT>());
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned) csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kern, *g.tmap_a, w.tmap, *g.tmap_out, *tres, *tout2, p);
}

int gemm_pick_bn(int N) {
    if (N % 256 == 0) return 256;
    if (N % 192 == 0) return 192;
    if (N % 128 == 0) return 128;
    if (N % 64 == 0) return 64;
    return 0;
}

// Column-block width for a weight whose GEMM is followed by residual + LayerNorm: the fused epilogue exists for 256- and
// 128-wide blocks held by one cluster of at most 4 CTA pairs, so N = 384 (all-MiniLM, bge-small-en) takes three 128-wide
// blocks rather than gemm_pick_bn's two 192-wide ones.
int gemm_pick_bn_ln(int N) {
    static const bool on = [] { const char * e = getenv("BERT_LN_BN128"); return !e || atoi(e) != 0; }();   // 0: A/B against the unfused form
    const int bn = gemm_pick_bn(N);
    if (on && bn != 256 && bn != 128 && N % 128 == 0 && N / 128 <= 4) return 128;
    return bn;
}

bool gemm_ln_fusable(const DevWeight & w) {
    // the cluster must hold every column block of a row stripe: N / BN pairs, at most 4 (8 CTAs)
    return (w.bn == 256 || w.bn == 128) && w.N / w.bn <= 4 && w.bias != nullptr;
}

cudaError_t launch_gemm_tc(const GemmArgs & g, int num_sms, cudaStream_t st) {
    const DevWeight & w = *g.w;
    int bn = w.bn;
    if (g.bn_override > 0 && (w.type == W_Q4_0 || w.type == W_Q4_1) && w.N % g.bn_override == 0 && !(g.epilogue & EPI_LN)) bn = g.bn_override;
    if (bn == 0 || w.K % BK != 0) return cudaErrorInvalidValue;
    if (g.epilogue & EPI_LN) {
        if (!gemm_ln_fusable(w) || (g.epilogue & EPI_GELU) || !(g.epilogue & EPI_RESIDUAL) || !(g.epilogue & EPI_BIAS)) return cudaErrorInvalidValue;
        // short K (the attention-output GEMM): the epilogue chain outlasts the mainloop -> EPB variant
        // (BERT_LN_EPB: 0 never, 1 when K <= 1024 (default), 2 always)
        static const int epb_mode = [] { const char * e = getenv("BERT_LN_EPB"); return e ? atoi(e) : 1; }();
        const bool epb = epb_mode == 2 || (epb_mode == 1 && w.K <= 1024);
#define DISPATCH_LN(BN_, EPB_)                                                                  \
        switch (w.type) {                                                                       \
            case W_F16:  return launch_one<BN_, W_F16, true, true, EPB_>(g, num_sms, st);        \
            case W_Q4_0: return launch_one<BN_, W_Q4_0, true, true, EPB_>(g, num_sms, st);       \
            case W_Q4_1: return launch_one<BN_, W_Q4_1, true, true, EPB_>(g, num_sms, st);       \
            default: return cudaErrorInvalidValue;                                              \
        }
        if (bn == 256) { if (epb) { DISPATCH_LN(256, true) } else { DISPATCH_LN(256, false) } }
        else { if (epb) { DISPATCH_LN(128, true) } else { DISPATCH_LN(128, false) } }
#undef DISPATCH_LN
    }
#define DISPATCH_WT(BN_)                                                                        \
    if (g.epilogue & EPI_RESIDUAL) {                                                            \
        switch (w.type) {                                                                       \
            case W_F16:  return launch_one<BN_, W_F16, false, true>(g, num_sms, st);             \
            case W_Q4_0: return launch_one<BN_, W_Q4_0, false, true>(g, num_sms, st);            \
            case W_Q4_1: return launch_one<BN_, W_Q4_1, false, true>(g, num_sms, st);            \
            default: return cudaErrorInvalidValue;                                              \
        }                                                                                       \
    }                                                                                           \
    switch (w.type) {                                                                           \
        case W_F16:  return launch_one<BN_, W_F16, false, false>(g, num_sms, st);                \
        case W_Q4_0: return launch_one<BN_, W_Q4_0, false, false>(g, num_sms, st);               \
        case W_Q4_1: return launch_one<BN_, W_Q4_1, false, false>(g, num_sms, st);               \
        default: return cudaErrorInvalidValue;                                                  \
    }
    switch (bn) {
        case 256: DISPATCH_WT(256)
        case 192: DISPATCH_WT(192)
        case 128: DISPATCH_WT(128)
        case 64:  DISPATCH_WT(64)
    }
#undef DISPATCH_WT
    return cudaErrorInvalidValue;
}

}  // namespace b200
