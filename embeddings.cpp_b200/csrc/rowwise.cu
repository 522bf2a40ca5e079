// The HBM-bound row kernels of the encoder forward:
//   K0 embed_ln   token/position/type gather + sum + LayerNorm  (reference bert.cpp:963-984,
//                 get_rows ggml.c:11614-11696, norm ggml.c:10605-10653 with eps 1e-5)
//   K3 ln         LayerNorm of (GEMM + bias + residual)          (bert.cpp:1047-1056, 1073-1082)
//   K4 pool_norm  mean over the sentence's tokens + L2 normalise (bert.cpp:1087-1095)
// One warp owns one row; each lane owns one 32-element block (64 B fp16) so that the q8
// activation fake-quant (one scale per 32 values, ggml.c:1146-1300 / 1425-1480) is
// thread-local.  Loads/stores are 16-B vectors, statistics are f32 warp-shuffle reductions.
#include "kernels.cuh"

namespace b200 {

namespace {

constexpr int ROWS_PER_CTA = 8;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void load32_h(const __half * p, float (&v)[32]) {
    const uint4 * q = reinterpret_cast<const uint4 *>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint4 u = __ldg(q + i);
        const __half2 * h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 f = __half22float2(h[e]);
            v[i * 8 + 2 * e] = f.x; v[i * 8 + 2 * e + 1] = f.y;
        }
    }
}
__device__ __forceinline__ void store32_h(__half * p, const float (&v)[32]) {
    uint4 * q = reinterpret_cast<uint4 *>(p);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        uint4 u;
        __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
        for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[i * 8 + 2 * e], v[i * 8 + 2 * e + 1]);
        q[i] = u;
    }
}
__device__ __forceinline__ void load32_f(const float * p, float (&v)[32]) {
    const float4 * q = reinterpret_cast<const float4 *>(p);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float4 f = __ldg(q + i);
        v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w;
    }
}

// LayerNorm of a row distributed as one 32-block per active lane, then write-out.
__device__ __forceinline__ void ln_finish(float (&v)[32], bool active, int H, int blk,
                                          const float * gamma, const float * beta,
                                          __half * out_row, __half * out_fq_row, int fq_type) {
    float s = 0.f;
    if (active) {
#pragma unroll
        for (int i = 0; i < 32; ++i) s += v[i];
    }
    const float mean = warp_sum(s) / (float) H;
    float s2 = 0.f;
    if (active) {
#pragma unroll
        for (int i = 0; i < 32; ++i) { v[i] -= mean; s2 += v[i] * v[i]; }
    }
    const float var = warp_sum(s2) / (float) H;
    const float scale = 1.0f / sqrtf(var + 1e-5f);
    if (!active) return;
    float g[32], b[32];
    load32_f(gamma + blk * 32, g);
    load32_f(beta + blk * 32, b);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = v[i] * scale * g[i] + b[i];
    store32_h(out_row + blk * 32, v);
    if (out_fq_row) {
        fake_quant32(v, fq_type);
        store32_h(out_fq_row + blk * 32, v);
    }
}

// decode one 32-weight block of an embedding row (file nibble order: byte j = e[j] | e[j+16] << 4)
__device__ __forceinline__ void decode_block(const DevTable & tb, size_t row, int blk, float (&v)[32]) {
    const int nb = tb.H / 32;
    if (tb.type == W_F32) {
        load32_f(reinterpret_cast<const float *>(tb.data) + row * tb.H + blk * 32, v);
    } else if (tb.type == W_F16) {
        load32_h(reinterpret_cast<const __half *>(tb.data) + row * tb.H + blk * 32, v);
    } else {
        const uint4 u = __ldg(reinterpret_cast<const uint4 *>(tb.data) + row * nb + blk);
        const float d = __half2float(__ldg(tb.qd + row * nb + blk));
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
        if (tb.type == W_Q4_0) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint32_t byte = (w[j >> 2] >> (8 * (j & 3))) & 0xFF;
                v[j] = (float) ((int) (byte & 15) - 8) * d;
                v[j + 16] = (float) ((int) (byte >> 4) - 8) * d;
            }
        } else {
            const float m = __half2float(__ldg(tb.qm + row * nb + blk));
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint32_t byte = (w[j >> 2] >> (8 * (j & 3))) & 0xFF;
                v[j] = (float) (byte & 15) * d + m;
                v[j + 16] = (float) (byte >> 4) * d + m;
            }
        }
    }
}

__global__ void __launch_bounds__(ROWS_PER_CTA * 32)
embed_ln_kernel(const DevTable word, const float * __restrict__ postype,
                const float * __restrict__ gamma, const float * __restrict__ beta,
                const int32_t * __restrict__ ids, const int32_t * __restrict__ cu, int n_sent, int T, int H,
                __half * __restrict__ out, __half * __restrict__ out_fq, int fq_type) {
    const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
    if (row >= T) return;
    const int lane = threadIdx.x & 31;
    const bool active = lane < H / 32;
    float v[32];
    if (active) {
        int id = ids[row];
        if (id < 0 || id >= word.rows) id = 0;              // never read out of the table
        decode_block(word, (size_t) id, lane, v);
        float pt[32];
        // position of this packed token inside its sentence: largest s with cu[s] <= row
        int lo = 0, hi = n_sent;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(cu + mid) <= row) lo = mid; else hi = mid;
        }
        const int pos = row - __ldg(cu + lo);
        load32_f(postype + (size_t) pos * H + lane * 32, pt);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += pt[i];
    }
    ln_finish(v, active, H, lane, gamma, beta, out + (size_t) row * H,
              out_fq ? out_fq + (size_t) row * H : nullptr, fq_type);
}

// K3.  Lane l owns the 16-B chunks c = l + 32 i (8 fp16 each) of the row, so every warp-level
// load/store is one fully coalesced 512-B access; a q8 block (32 values) is 4 adjacent lanes.
template <int NCH>
__global__ void __launch_bounds__(ROWS_PER_CTA * 32)
ln_kernel(const __half * __restrict__ x, const float * __restrict__ gamma, const float * __restrict__ beta,
          int T, int H, __half * __restrict__ out, __half * __restrict__ out_fq, int fq_type) {
    const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
    if (row >= T) return;
    const int lane = threadIdx.x & 31;
    const int nchunk = H / 8;
    float v[NCH][8];
    const uint4 * xr = reinterpret_cast<const uint4 *>(x + (size_t) row * H);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        if (c < nchunk) {
            const uint4 u = __ldg(xr + c);
            const __half2 * h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(h[e]);
                v[i][2 * e] = f.x; v[i][2 * e + 1] = f.y;
                s += f.x + f.y;
            }
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[i][e] = 0.f;
        }
    }
    const float mean = warp_sum(s) / (float) H;
    float s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        if (lane + 32 * i < nchunk) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { v[i][e] -= mean; s2 += v[i][e] * v[i][e]; }
        }
    }
    const float var = warp_sum(s2) / (float) H;
    const float scale = 1.0f / sqrtf(var + 1e-5f);
    uint4 * orow = reinterpret_cast<uint4 *>(out + (size_t) row * H);
    uint4 * qrow = out_fq ? reinterpret_cast<uint4 *>(out_fq + (size_t) row * H) : nullptr;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        const bool ok = c < nchunk;
        if (ok) {
            const float4 g0 = __ldg(reinterpret_cast<const float4 *>(gamma) + 2 * c);
            const float4 g1 = __ldg(reinterpret_cast<const float4 *>(gamma) + 2 * c + 1);
            const float4 b0 = __ldg(reinterpret_cast<const float4 *>(beta) + 2 * c);
            const float4 b1 = __ldg(reinterpret_cast<const float4 *>(beta) + 2 * c + 1);
            const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            uint4 u;
            __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[i][e] = v[i][e] * scale * g[e] + b[e];
#pragma unroll
            for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[i][2 * e], v[i][2 * e + 1]);
            orow[c] = u;
        }
        if (qrow) {                                        // warp-uniform branch: all lanes shuffle
            float amax = 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(v[i][e]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d = amax / 127.f;
            const float id = amax != 0.f ? 127.f / amax : 0.f;
            const float dd = fq_type == W_Q4_0 ? __half2float(__float2half_rn(d)) : d;
            if (ok) {
                uint4 u;
                __half2 * h = reinterpret_cast<__half2 *>(&u);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    h[e] = __floats2half2_rn(rintf(v[i][2 * e] * id) * dd, rintf(v[i][2 * e + 1] * id) * dd);
                }
                qrow[c] = u;
            }
        }
    }
}

// one CTA per sentence; thread = (8-column chunk, token stripe)
__global__ void __launch_bounds__(256)
pool_norm_kernel(const __half * __restrict__ x, const int32_t * __restrict__ cu, int H, float * __restrict__ out) {
    extern __shared__ float sm[];                       // [stripes][H] then reduction scratch
    const int sent = blockIdx.x;
    const int t0 = cu[sent], len = cu[sent + 1] - t0;
    const int ncol = H / 8;
    const int stripes = 256 / ncol;
    const int c = threadIdx.x % ncol, sidx = threadIdx.x / ncol;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (sidx < stripes) {
        for (int t = sidx; t < len; t += stripes) {
            const uint4 u = __ldg(reinterpret_cast<const uint4 *>(x + (size_t) (t0 + t) * H) + c);
            const __half2 * h = reinterpret_cast<const __half2 *>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(h[e]);
                acc[2 * e] += f.x; acc[2 * e + 1] += f.y;
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) sm[sidx * H + c * 8 + e] = acc[e];
    }
    __syncthreads();
    // masked mean (weights 1/len on valid tokens, bert.cpp:912) and sum of squares
    const float inv_len = 1.0f / (float) len;
    float ss = 0.f;
    for (int i = threadIdx.x; i < H; i += 256) {
        float s = 0.f;
        for (int k = 0; k < stripes; ++k) s += sm[k * H + i];
        s *= inv_len;
        sm[i] = s;                                      // stripe 0 slot now holds the pooled value
        ss += s * s;
    }
    __shared__ float red[8];
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) tot += red[k];
    const float inv_norm = 1.0f / sqrtf(tot);           // no epsilon (bert.cpp:1092-1094)
    for (int i = threadIdx.x; i < H; i += 256) out[(size_t) sent * H + i] = sm[i] * inv_norm;
}

}  // namespace

cudaError_t launch_embed_ln(const DevTable & word, const float * postype, const float * gamma, const float * beta,
                            const int32_t * ids, const int32_t * cu_seqlens, int n_sent, int T, int H,
                            __half * out, __half * out_fq, int fq_type, cudaStream_t st) {
    if (T <= 0) return cudaSuccess;
    embed_ln_kernel<<<(T + ROWS_PER_CTA - 1) / ROWS_PER_CTA, ROWS_PER_CTA * 32, 0, st>>>(
        word, postype, gamma, beta, ids, cu_seqlens, n_sent, T, H, out, out_fq, fq_type);
    return cudaGetLastError();
}

cudaError_t launch_ln(const __half * x, const float * gamma, const float * beta, int T, int H,
                      __half * out, __half * out_fq, int fq_type, cudaStream_t st) {
    if (T <= 0) return cudaSuccess;
    const int grid = (T + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    const int nch = (H / 8 + 31) / 32;
    if (nch <= 1) ln_kernel<1><<<grid, ROWS_PER_CTA * 32, 0, st>>>(x, gamma, beta, T, H, out, out_fq, fq_type);
    else if (nch == 2) ln_kernel<2><<<grid, ROWS_PER_CTA * 32, 0, st>>>(x, gamma, beta, T, H, out, out_fq, fq_type);
    else if (nch == 3) ln_kernel<3><<<grid, ROWS_PER_CTA * 32, 0, st>>>(x, gamma, beta, T, H, out, out_fq, fq_type);
    else ln_kernel<4><<<grid, ROWS_PER_CTA * 32, 0, st>>>(x, gamma, beta, T, H, out, out_fq, fq_type);
    return cudaGetLastError();
}

cudaError_t launch_pool_norm(const __half * x, const int32_t * cu_seqlens, int n_sent, int H,
                             float * out, cudaStream_t st) {
    if (n_sent <= 0) return cudaSuccess;
    const int stripes = 256 / (H / 8);
    const size_t smem = (size_t) (stripes > 0 ? stripes : 1) * H * sizeof(float);
    pool_norm_kernel<<<n_sent, 256, smem, st>>>(x, cu_seqlens, H, out);
    return cudaGetLastError();
}

}  // namespace b200
