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
//   warps 6-9   (q4 weights only) dequantise the BN x 64 weight stage straight from the
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
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GemmParams {
    const uint4 * qnib; const __half * qd; const __half * qm;
    const float * bias; const __half * residual; __half * out;
    int M, N, K, epilogue, fq_type;
};

__device__ __forceinline__ float gelu_ref_numerics(float x) {
    // reference: y = f16(gelu_f32(f16(x))) through a 64K-entry table (ggml.c:3561-3580, 4332)
    const float f = __half2float(__float2half_rn(x));
    const float u = 0.79788456080286535587989211986876f * f * (1.0f + 0.044715f * f * f);
    const float t = 1.0f - 2.0f / (__expf(2.0f * u) + 1.0f);          // tanh(u)
    return __half2float(__float2half_rn(0.5f * f * (1.0f + t)));
}

// two nibbles -> two fp16 weights.  `w` holds e(2i) at bits [4i,4i+4), e(2i+1) at [16+4i,..)
template <int WT>
__device__ __forceinline__ uint32_t dq_pair(uint32_t w, int i, __half2 d2, __half2 m2) {
    uint32_t x = ((w >> (4 * i)) & 0x000F000Fu) | 0x64006400u;        // fp16 1024 + q
    __half2 h = *reinterpret_cast<__half2 *>(&x);
    if (WT == W_Q4_0) {
        h = __hmul2(__hsub2(h, __half2half2(__ushort_as_half((unsigned short) 0x6408))), d2);   // (q-8)*d
    } else {
        h = __hfma2(__hsub2(h, __half2half2(__ushort_as_half((unsigned short) 0x6400))), d2, m2); // q*d+m
    }
    return *reinterpret_cast<uint32_t *>(&h);
}

template <int BN, int WT>
__global__ void __launch_bounds__(WT == W_F16 ? 192 : 320, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    constexpr int S = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);
    uint8_t * smem_a = smem;                                  // S x 16 KB
    uint8_t * smem_b = smem + S * Cfg::kABytes;               // S x BN*128 B
    uint64_t * bars = (uint64_t *) (smem + S * Cfg::kStageBytes);
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
        for (int s = 0; s < S; ++s) {
            mbar_init(&full[s], WT == W_F16 ? 1 : 1 + 4);
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
        const int q = warp & 3;                    // TMEM lane quadrant this warp may access
        int it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
            const int m0 = (tile / num_n) * BM, n0 = (tile % num_n) * BN;
            const int acc = it & 1;
            const uint32_t acc_ph = (it >> 1) & 1;
            mbar_wait(&tfull[acc], acc_ph);
            tc_fence_after_sync();
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < p.M;
            const size_t row_off = (size_t) row * p.N;
#pragma unroll 1
            for (int j = 0; j < BN / 32; ++j) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(tmem_base + ((uint32_t) (q * 32) << 16) + acc * Cfg::kAccStride + j * 32, r);
                tmem_ld_wait();
                const int c0 = n0 + j * 32;
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
                if (p.epilogue & EPI_BIAS) {
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        const float4 b = __ldg(reinterpret_cast<const float4 *>(p.bias + c0 + i));
                        v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
                    }
                }
                if (p.epilogue & EPI_GELU) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = gelu_ref_numerics(v[i]);
                }
                if ((p.epilogue & EPI_RESIDUAL) && row_ok) {
                    const uint4 * rp = reinterpret_cast<const uint4 *>(p.residual + row_off + c0);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const uint4 u = __ldg(rp + i);
                        const __half2 * h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = __half22float2(h[e]);
                            v[i * 8 + 2 * e] += f.x; v[i * 8 + 2 * e + 1] += f.y;
                        }
                    }
                }
                if (p.epilogue & EPI_FAKEQUANT) fake_quant32(v, p.fq_type);
                if (row_ok) {
                    uint4 * op = reinterpret_cast<uint4 *>(p.out + row_off + c0);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 u;
                        __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
                        for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[i * 8 + 2 * e], v[i * 8 + 2 * e + 1]);
                        op[i] = u;
                    }
                }
            }
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
        }
    } else {
        // ------------------------------------------------------------ q4 dequant producers (4 warps)
        if (WT != W_F16) {
            const int t = threadIdx.x - 192;                     // 0..127
            constexpr int UNITS = BN * 2 / 128;                  // 32-weight blocks per thread per stage
            int s = 0; uint32_t ph = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int n0 = (tile % num_n) * BN;
                for (int kb = 0; kb < num_k; ++kb) {
                    // stage kb of rows n0..n0+BN-1 is contiguous: [kb][n][2] blocks
                    const size_t blk0 = ((size_t) kb * p.N + n0) * 2;
                    uint4 nib[UNITS]; __half d[UNITS], m[UNITS];
#pragma unroll
                    for (int u = 0; u < UNITS; ++u) {
                        const size_t b = blk0 + t + 128 * u;
                        nib[u] = __ldg(p.qnib + b);
                        d[u] = __ldg(p.qd + b);
                        m[u] = WT == W_Q4_1 ? __ldg(p.qm + b) : __half();
                    }
                    mbar_wait(&empty[s], ph ^ 1);
                    uint8_t * bt = smem_b + s * Cfg::kBBytes;
#pragma unroll
                    for (int u = 0; u < UNITS; ++u) {
                        const int unit = t + 128 * u;
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
    kern<<<grid, WT == W_F16 ? 192 : 320, Cfg::kSmemBytes, st>>>(*g.tmap_a, w.tmap, p);
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
