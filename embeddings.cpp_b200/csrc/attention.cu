// K2 -- variable-length fused attention: per (sentence, head, 64-query tile)
//   O = softmax(Q K^T / sqrt(d)) V   over the sentence's own tokens only.
// Replaces the reference's mask build (bert.cpp:957-961: 0 / -1e5 additive mask, which the
// f16 exp table turns into exactly 0 probability), scores + scale + mask + soft_max
// (bert.cpp:1018-1025, ggml.c:12008-12077) and P*V + permute/cont/cpy (bert.cpp:1027-1036).
// Nothing of size L x L ever reaches HBM; padded queries/keys are never computed.
//
// v1 data path: fp16 operands, f32 accumulate, online softmax (exp2f) in registers,
// warp-level mma.sync m16n8k16 tiles fed by ldmatrix from XOR-swizzled shared memory,
// K/V tiles double-buffered with cp.async.  (A tcgen05/TMEM version replaces the two
// MMAs later; the softmax and the varlen tiling stay.)
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int BQ = 64, BKV = 64, ATT_THREADS = 128;

__device__ __forceinline__ void cp_async16(void * smem, const void * gmem, bool valid) {
    const uint32_t s = (uint32_t) __cvta_generic_to_shared(smem);
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(s), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void * p) {
    const uint32_t s = (uint32_t) __cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(s));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void * p) {
    const uint32_t s = (uint32_t) __cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(s));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

// byte offset of 16-B chunk `c` of row `r` in a tile with D fp16 per row (swizzled)
template <int D> __device__ __forceinline__ int sw_off(int r, int c) {
    if (D == 64) return r * 128 + ((c ^ (r & 7)) << 4);
    return r * 64 + ((c ^ ((r >> 1) & 3)) << 4);          // D == 32: two rows per 128 B
}

template <int D>
__device__ __forceinline__ void load_tile(uint8_t * dst, const __half * src, int row_stride /*elements*/,
                                          int row0, int n_valid_rows, int tid) {
    constexpr int CH = D / 8;                 // 16-B chunks per row
#pragma unroll
    for (int i = tid; i < 64 * CH; i += ATT_THREADS) {
        const int r = i / CH, c = i % CH;
        const bool ok = (row0 + r) < n_valid_rows;
        const __half * g = src + (size_t) (ok ? row0 + r : 0) * row_stride + c * 8;
        cp_async16(dst + sw_off<D>(r, c), g, ok);
    }
}

template <int D>
__global__ void __launch_bounds__(ATT_THREADS)
attention_kernel(const __half * __restrict__ qkv, const int32_t * __restrict__ cu, int H, float scale_log2,
                 __half * __restrict__ out, int fq_type) {
    pdl_launch_dependents();
    pdl_wait();
    const int sent = blockIdx.z, head = blockIdx.y, qt = blockIdx.x;
    const int t0 = cu[sent], len = cu[sent + 1] - t0;
    const int q0 = qt * BQ;
    if (q0 >= len) return;
    constexpr int TILE = 64 * D * 2;
    __shared__ __align__(128) uint8_t sm[TILE * 5];          // Q, K0, V0, K1, V1
    uint8_t * sQ = sm;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ld = 3 * H;
    const __half * qbase = qkv + (size_t) t0 * ld + head * D;
    const __half * kbase = qbase + H;
    const __half * vbase = qbase + 2 * H;

    load_tile<D>(sQ, qbase, ld, q0, len, tid);
    load_tile<D>(sm + TILE, kbase, ld, 0, len, tid);
    load_tile<D>(sm + 2 * TILE, vbase, ld, 0, len, tid);
    cp_async_commit();

    const int n_kv = (len + BKV - 1) / BKV;
    constexpr int KT = D / 16;                // k-steps for QK^T
    constexpr int NT = D / 8;                 // n-tiles of O
    uint32_t qf[KT][4];
    float o[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    const int g = lane >> 2, t = lane & 3;

    for (int kv = 0; kv < n_kv; ++kv) {
        uint8_t * sK = sm + TILE * (1 + 2 * (kv & 1));
        uint8_t * sV = sK + TILE;
        if (kv + 1 < n_kv) {
            uint8_t * nK = sm + TILE * (1 + 2 * ((kv + 1) & 1));
            load_tile<D>(nK, kbase, ld, (kv + 1) * BKV, len, tid);
            load_tile<D>(nK + TILE, vbase, ld, (kv + 1) * BKV, len, tid);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (kv == 0) {
#pragma unroll
            for (int kk = 0; kk < KT; ++kk) {
                const int r = warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
                ldsm_x4(qf[kk], sQ + sw_off<D>(r, kk * 2 + (lane >> 4)));
            }
        }
        // ---- S = Q K^T : 16 x 64 per warp -> 8 n-tiles
        float s[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
        for (int kk = 0; kk < KT; ++kk) {
#pragma unroll
            for (int jp = 0; jp < 4; ++jp) {          // pairs of key n-tiles (16 keys)
                uint32_t b[4];
                const int r = jp * 16 + (lane & 7) + 8 * (lane >> 4);
                ldsm_x4(b, sK + sw_off<D>(r, kk * 2 + ((lane >> 3) & 1)));
                mma16816(s[2 * jp], qf[kk], b[0], b[1]);
                mma16816(s[2 * jp + 1], qf[kk], b[2], b[3]);
            }
        }
        // ---- mask keys beyond the sentence, online softmax
        const int kbase_idx = kv * BKV;
        float mx[2] = {mrow[0], mrow[1]};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int key = kbase_idx + j * 8 + 2 * t + (e & 1);
                const float v = key < len ? s[j][e] * scale_log2 : -INFINITY;
                s[j][e] = v;
                mx[e >> 1] = fmaxf(mx[e >> 1], v);
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
        }
        float corr[2], sum[2] = {0.f, 0.f};
#pragma unroll
        for (int h = 0; h < 2; ++h) { corr[h] = exp2f(mrow[h] - mx[h]); mrow[h] = mx[h]; }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float p = exp2f(s[j][e] - mx[e >> 1]);
                s[j][e] = p;
                sum[e >> 1] += p;
            }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) lrow[h] = lrow[h] * corr[h] + sum[h];
#pragma unroll
        for (int j = 0; j < NT; ++j) { o[j][0] *= corr[0]; o[j][1] *= corr[0]; o[j][2] *= corr[1]; o[j][3] *= corr[1]; }
        // ---- O += P V
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {              // 16 keys per step
            uint32_t a[4];
            a[0] = pack_h2(s[2 * kk][0], s[2 * kk][1]);
            a[1] = pack_h2(s[2 * kk][2], s[2 * kk][3]);
            a[2] = pack_h2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            a[3] = pack_h2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int jp = 0; jp < NT / 2; ++jp) {     // pairs of d n-tiles
                uint32_t b[4];
                const int r = kk * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
                ldsm_x4_t(b, sV + sw_off<D>(r, jp * 2 + (lane >> 4)));
                mma16816(o[2 * jp], a, b[0], b[1]);
                mma16816(o[2 * jp + 1], a, b[2], b[3]);
            }
        }
        __syncthreads();
    }
    // ---- finalize: divide by the row sums, optional q8 fake-quant per 32-wide block, store
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 1);
        lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 2);
    }
    const float inv[2] = {1.f / lrow[0], 1.f / lrow[1]};
#pragma unroll
    for (int j = 0; j < NT; ++j) { o[j][0] *= inv[0]; o[j][1] *= inv[0]; o[j][2] *= inv[1]; o[j][3] *= inv[1]; }
    if (fq_type == W_Q4_0 || fq_type == W_Q4_1) {
#pragma unroll
        for (int blk = 0; blk < D / 32; ++blk) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float amax = 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    amax = fmaxf(amax, fmaxf(fabsf(o[blk * 4 + j][2 * h]), fabsf(o[blk * 4 + j][2 * h + 1])));
                }
                amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
                amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
                const float d = amax / 127.f;
                const float id = amax != 0.f ? 127.f / amax : 0.f;
                const float dd = fq_type == W_Q4_0 ? __half2float(__float2half_rn(d)) : d;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    o[blk * 4 + j][2 * h] = rintf(o[blk * 4 + j][2 * h] * id) * dd;
                    o[blk * 4 + j][2 * h + 1] = rintf(o[blk * 4 + j][2 * h + 1] * id) * dd;
                }
            }
        }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int row = q0 + warp * 16 + g + 8 * h;
        if (row < len) {
            __half * op = out + (size_t) (t0 + row) * H + head * D;
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                *reinterpret_cast<__half2 *>(op + j * 8 + 2 * t) = __floats2half2_rn(o[j][2 * h], o[j][2 * h + 1]);
            }
        }
    }
}

}  // namespace

cudaError_t launch_attention(const __half * qkv, const int32_t * cu_seqlens, int n_sent, int max_len,
                             int n_head, int d_head, __half * out, int fq_type, cudaStream_t st) {
    if (n_sent <= 0) return cudaSuccess;
    const int H = n_head * d_head;
    const float scale_log2 = 1.4426950408889634f / sqrtf((float) d_head);
    dim3 grid((max_len + BQ - 1) / BQ, n_head, n_sent);
    if (d_head == 64) {
        return launch_pdl(attention_kernel<64>, grid, dim3(ATT_THREADS), 0, st, qkv, cu_seqlens, H, scale_log2, out, fq_type);
    } else if (d_head == 32) {
        return launch_pdl(attention_kernel<32>, grid, dim3(ATT_THREADS), 0, st, qkv, cu_seqlens, H, scale_log2, out, fq_type);
    }
    return cudaErrorInvalidValue;
}

}  // namespace b200
