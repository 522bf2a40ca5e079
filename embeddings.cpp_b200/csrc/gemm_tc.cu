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

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int UMMA_K = 16;

template <int BN> struct GemmCfg {
    static constexpr int kStages   = BN >= 192 ? 4 : 6;
    static constexpr int kABytes   = BM * BK * 2;
    static constexpr int kBBytes   = BN * BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kTmemCols = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);
    static constexpr int kAccStride = BN <= 64 ? 64 : (BN <= 128 ? 128 : 256);
    static constexpr int kOutBytes = 4 * 2 * 4096;      // 4 epilogue warps x 2 buffers x [32 rows][128 B]
    static constexpr int kSmemBytes = kStages * kStageBytes + kOutBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GemmParams {
    const uint4 * qnib; const __half * qd; const __half * qm;
    const float * bias; const __half * residual; __half * out;
    int M, N, K, epilogue, fq_type;
};

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
// Reference GELU numerics: y = f16(gelu_f32(f16(x))) through a 64K-entry table
// (ggml.c:3561-3580, 4332).  tanh-GELU rewritten as f / (1 + exp(-2u)), u = sqrt(2/pi) f (1 + 0.044715 f^2),
// with ex2.approx / rcp.approx (2 MUFU per element) and packed f32x2 arithmetic.
__device__ __forceinline__ float2 gelu2_ref_numerics(float2 v) {
    const float c1 = -2.0f * 1.4426950408889634f * 0.79788456080286535588f;
    const float c2 = c1 * 0.044715f;
    const float2 f = round_h2(v);
    const float2 p = __ffma2_rn(__fmul2_rn(f, f), make_float2(c2, c2), make_float2(c1, c1));
    const float2 z = __fmul2_rn(f, p);
    const float2 d = __fadd2_rn(make_float2(ex2_approx(z.x), ex2_approx(z.y)), make_float2(1.f, 1.f));
    return round_h2(__fmul2_rn(f, make_float2(rcp_approx(d.x), rcp_approx(d.y))));
}
// q8 fake-quant of 32 consecutive values (16 float2) held by one thread; rint via the
// 1.5*2^23 magic constant (round-half-even, |x| <= 127 here) to stay on the FMA pipe
__device__ __forceinline__ void fake_quant32_v2(float2 (&v)[16], int fq_type) {
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fmaxf(fabsf(v[i].x), fabsf(v[i].y)));
    const float d = amax / 127.f;
    const float id = amax != 0.f ? 127.f / amax : 0.f;
    const float dd = fq_type == W_Q4_0 ? __half2float(__float2half_rn(d)) : d;
    const float2 id2 = make_float2(id, id), dd2 = make_float2(dd, dd);
    const float2 mg = make_float2(12582912.f, 12582912.f), nmg = make_float2(-12582912.f, -12582912.f);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float2 r = __fadd2_rn(__ffma2_rn(v[i], id2, mg), nmg);
        v[i] = __fmul2_rn(r, dd2);
    }
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

// two nibbles -> two fp16 weights.  `w` holds e(2i) at bits [4i,4i+4), e(2i+1) at [16+4i,..).
// The nibble is OR-ed into the mantissa of an fp16 constant whose ulp matches the nibble's
// bit position (1024 for mantissa bits 0-3, 64 for bits 4-7), so a word needs one shift for its
// four pairs; then (q - 8) * d  or  q * d + m  in half2.
template <int WT>
__device__ __forceinline__ uint32_t dq_pair(uint32_t w, int i, __half2 d2, __half2 m2) {
    uint32_t x;
    unsigned short base;                 // fp16 bits of (magic + 8) for q4_0, magic for q4_1
    const uint32_t ws = i < 2 ? w : (w >> 8);                       // one shift serves pairs 2 and 3
    if ((i & 1) == 0) { x = (ws & 0x000F000Fu) | 0x64006400u; base = WT == W_Q4_0 ? 0x6408 : 0x6400; }   // 1024 + q
    else              { x = (ws & 0x00F000F0u) | 0x54005400u; base = WT == W_Q4_0 ? 0x5480 : 0x5400; }   //   64 + q
    __half2 h = __hsub2(*reinterpret_cast<__half2 *>(&x), __half2half2(__ushort_as_half(base)));
    h = WT == W_Q4_0 ? __hmul2(h, d2) : __hfma2(h, d2, m2);
    return *reinterpret_cast<uint32_t *>(&h);
}

template <int BN, int WT>
__global__ void __launch_bounds__(WT == W_F16 ? 192 : 192 + 32 * DQ_WARPS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_out, const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int S = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    uint8_t * smem_a = smem;                                  // S x 16 KB
    uint8_t * smem_b = smem + S * Cfg::kABytes;               // S x BN*128 B
    uint8_t * smem_out = smem + S * Cfg::kStageBytes;         // 4 warps x 2 x 4 KB staging for TMA stores
    uint64_t * bars = (uint64_t *) (smem + S * Cfg::kStageBytes + Cfg::kOutBytes);
    uint64_t * full = bars;                 // [S]
    uint64_t * empty = bars + S;            // [S]
    uint64_t * tfull = bars + 2 * S;        // [2]
    uint64_t * tempty = bars + 2 * S + 2;   // [2]
    uint32_t * tmem_slot = (uint32_t *) (bars + 2 * S + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_n = p.N / BN;
    const int num_m = (p.M + BM - 1) / BM;
    const int num_tiles = num_m * num_n;
    const int num_k = p.K / BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        if (WT == W_F16) tma_prefetch_desc(&tmap_b);
        tma_prefetch_desc(&tmap_out);
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], WT == W_F16 ? 1 : 1 + DQ_WARPS);
            mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, Cfg::kTmemCols); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
                for (int kb = 0; kb < num_k; ++kb) {
                    mbar_wait(&empty[s], ph ^ 1);
                    mbar_expect_tx(&full[s], WT == W_F16 ? Cfg::kStageBytes : Cfg::kABytes);
                    tma_load_2d(smem_a + s * Cfg::kABytes, &tmap_a, &full[s], kb * BK, m0);
                    if (WT == W_F16) tma_load_2d(smem_b + s * Cfg::kBBytes, &tmap_b, &full[s], kb * BK, n0);
                    if (++s == S) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        constexpr uint32_t idesc = umma_idesc_f16(BM, BN);
        int s = 0; uint32_t ph = 0; int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            mbar_wait(&tempty[acc], acc_ph ^ 1);
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem_base + acc * Cfg::kAccStride;
            for (int kb = 0; kb < num_k; ++kb) {
                mbar_wait(&full[s], ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint64_t adesc = umma_smem_desc_sw128(smem_u32(smem_a + s * Cfg::kABytes));
                    const uint64_t bdesc = umma_smem_desc_sw128(smem_u32(smem_b + s * Cfg::kBBytes));
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        // +32 B per K=16 step inside the 128-B swizzle row (address field is >>4)
                        umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    }
                    umma_commit(&empty[s]);                      // frees the smem stage
                    if (kb == num_k - 1) umma_commit(&tfull[acc]);   // accumulator ready
                }
                __syncwarp();
                if (++s == S) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp < 6) {
        // ------------------------------------------------------------ epilogue (4 warps)
        // Per 64-column box: (residual: coalesced LDG -> swizzled smem) ; TMEM -> registers ;
        // bias / GELU / + residual (own row from smem) / q8 fake-quant ; fp16 -> swizzled smem ;
        // one TMA store of the warp's [32 rows x 64 cols] sub-box.  Global traffic is whole
        // 128-B rows in both directions; everything is warp-local (no CTA barrier).
        const int q = warp & 3;                    // TMEM lane quadrant this warp may access
        uint8_t * stage_base = smem_out + (warp - 2) * 2 * 4096;
        int it = 0;
        uint32_t box_ctr = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            const int wrow0 = m0 + q * 32;                  // first row of this warp's sub-tile
            mbar_wait(&tfull[acc], acc_ph);
            tc_fence_after_sync();
#pragma unroll 1
            for (int b = 0; b < BN / 64; ++b, ++box_ctr) {
                uint8_t * st = stage_base + (box_ctr & 1) * 4096;
                const int c0 = n0 + b * 64;
                // the TMA store that last read this buffer (two boxes ago) must be done reading
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
                if (p.epilogue & EPI_RESIDUAL) {
                    uint4 rr[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int id = lane + 32 * i, r = id >> 3, c = id & 7;
                        rr[i] = (wrow0 + r < p.M)
                            ? __ldg(reinterpret_cast<const uint4 *>(p.residual + (size_t) (wrow0 + r) * p.N + c0) + c)
                            : make_uint4(0, 0, 0, 0);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int id = lane + 32 * i, r = id >> 3, c = id & 7;
                        *reinterpret_cast<uint4 *>(st + r * 128 + ((c ^ (r & 7)) << 4)) = rr[i];
                    }
                    __syncwarp();
                }
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t r[32];
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t) (q * 32) << 16) + acc * Cfg::kAccStride + b * 64 + hh * 32, r);
                    tmem_ld_wait();
                    float2 v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
                    if (p.epilogue & EPI_BIAS) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 bb = __ldg(reinterpret_cast<const float4 *>(p.bias + c0 + hh * 32) + i);
                            v[2 * i] = __fadd2_rn(v[2 * i], make_float2(bb.x, bb.y));
                            v[2 * i + 1] = __fadd2_rn(v[2 * i + 1], make_float2(bb.z, bb.w));
                        }
                    }
                    if (p.epilogue & EPI_GELU) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = gelu2_ref_numerics(v[i]);
                    }
                    if (p.epilogue & EPI_RESIDUAL) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int c = hh * 4 + i;
                            const uint4 u = *reinterpret_cast<const uint4 *>(st + lane * 128 + ((c ^ (lane & 7)) << 4));
                            const __half2 * h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
                            for (int e = 0; e < 4; ++e) v[i * 4 + e] = __fadd2_rn(v[i * 4 + e], __half22float2(h[e]));
                        }
                    }
                    if (p.epilogue & EPI_FAKEQUANT) fake_quant32_v2(v, p.fq_type);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int c = hh * 4 + i;
                        uint4 u;
                        __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
                        for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[i * 4 + e].x, v[i * 4 + e].y);
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
            if (lane == 0) mbar_arrive(&tempty[acc]);
        }
        if (lane == 0) tma_store_wait_all<0>();
    } else {
        // ------------------------------------------------------------ q4 dequant producers
        if (WT != W_F16) {
            constexpr int DQT = 32 * DQ_WARPS;
            const int t = threadIdx.x - 192;                     // 0..DQT-1
            constexpr int UNITS = (BN * 2 + DQT - 1) / DQT;      // 32-weight blocks per thread per stage
            int s = 0; uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int n0 = (tile % num_n) * BN;
                for (int kb = 0; kb < num_k; ++kb) {
                    // stage kb of rows n0..n0+BN-1 is contiguous: [kb][n][2] blocks
                    const size_t blk0 = ((size_t) kb * p.N + n0) * 2;
                    uint4 nib[UNITS]; __half d[UNITS], m[UNITS];
#pragma unroll
                    for (int u = 0; u < UNITS; ++u) {
                        if (t + DQT * u < BN * 2) {
                            const size_t b = blk0 + t + DQT * u;
                            nib[u] = __ldg(p.qnib + b);
                            d[u] = __ldg(p.qd + b);
                            m[u] = WT == W_Q4_1 ? __ldg(p.qm + b) : __half();
                        }
                    }
                    mbar_wait(&empty[s], ph ^ 1);
                    uint8_t * bt = smem_b + s * Cfg::kBBytes;
#pragma unroll
                    for (int u = 0; u < UNITS; ++u) {
                        const int unit = t + DQT * u;
                        if (unit >= BN * 2) continue;
                        const int r = unit >> 1, half_k = unit & 1;          // row in tile, which 32-block
                        const __half2 d2 = __half2half2(d[u]), m2 = __half2half2(m[u]);
                        const uint32_t w[4] = {nib[u].x, nib[u].y, nib[u].z, nib[u].w};
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            uint4 o;
                            o.x = dq_pair<WT>(w[c], 0, d2, m2);
                            o.y = dq_pair<WT>(w[c], 1, d2, m2);
                            o.z = dq_pair<WT>(w[c], 2, d2, m2);
                            o.w = dq_pair<WT>(w[c], 3, d2, m2);
                            const int chunk = half_k * 4 + c;
                            *reinterpret_cast<uint4 *>(bt + r * 128 + ((chunk ^ (r & 7)) << 4)) = o;
                        }
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&full[s]);
                    if (++s == S) { s = 0; ph ^= 1; }
                }
            }
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BN, int WT>
static cudaError_t launch_one(const GemmArgs & g, int num_sms, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    auto kern = gemm_tc_kernel<BN, WT>;
    static bool attr_done = false;     // per instantiation; every device needs it -> set each time cheaply
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    (void) attr_done;
    if (e != cudaSuccess) return e;
    const DevWeight & w = *g.w;
    GemmParams p{w.qnib, w.qd, w.qm, w.bias, g.residual, g.out, g.M, w.N, w.K, g.epilogue, g.fq_type};
    const int tiles = ((g.M + BM - 1) / BM) * (w.N / BN);
    const int grid = tiles < num_sms ? tiles : num_sms;
    if (grid <= 0) return cudaSuccess;
    kern<<<grid, WT == W_F16 ? 192 : 192 + 32 * DQ_WARPS, Cfg::kSmemBytes, st>>>(*g.tmap_a, w.tmap, *g.tmap_out, p);
    return cudaGetLastError();
}

int gemm_pick_bn(int N) {
    if (N % 256 == 0) return 256;
    if (N % 192 == 0) return 192;
    if (N % 128 == 0) return 128;
    if (N % 64 == 0) return 64;
    return 0;
}

cudaError_t launch_gemm_tc(const GemmArgs & g, int num_sms, cudaStream_t st) {
    const DevWeight & w = *g.w;
    const int bn = w.bn;
    if (bn == 0 || w.K % BK != 0) return cudaErrorInvalidValue;
#define DISPATCH_WT(BN_)                                                                        \
    switch (w.type) {                                                                           \
        case W_F16:  return launch_one<BN_, W_F16>(g, num_sms, st);                              \
        case W_Q4_0: return launch_one<BN_, W_Q4_0>(g, num_sms, st);                             \
        case W_Q4_1: return launch_one<BN_, W_Q4_1>(g, num_sms, st);                             \
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
