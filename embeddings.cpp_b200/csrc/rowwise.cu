// The HBM-bound row kernels of the encoder forward:
//   K0 embed_ln   token/position/type gather + sum + LayerNorm  (reference bert.cpp:963-984,
//                 get_rows ggml.c:11614-11696, norm ggml.c:10605-10653 with eps 1e-5)
//   K3 ln         LayerNorm of (GEMM + bias + residual)          (bert.cpp:1047-1056, 1073-1082)
//   K4 pool_norm  mean over the sentence's tokens + L2 normalise (bert.cpp:1087-1095)
// One warp owns one row; lane l owns the 16-B chunks l + 32 i, so every warp access is a whole
// 512-B line group; the q8 activation fake-quant (one scale per 32 values, ggml.c:1146-1300 /
// 1425-1480) is a 4-lane shuffle.  Statistics are f32 warp-shuffle reductions.
#include <cstdlib>
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

namespace {

constexpr int ROWS_PER_CTA = 8;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Row LayerNorm + write-out shared by K0 and K3.  Lane l owns the 16-B chunks c = l + 32 i (8 fp16 each)
// of the row, so every warp-level load/store is one fully coalesced 512-B access; a q8 block (32 values)
// is 4 adjacent lanes.  `v` holds the row (0 in chunks past the row end), `s` this lane's sum.
// FULL: the row is exactly NCH * 32 chunks (H = 256 NCH), so no chunk needs a bounds predicate.
template <int NCH, bool FULL = false>
__device__ __forceinline__ void ln_row_finish(float (&v)[NCH][8], float s, int lane, int H,
                                              const float * __restrict__ gamma, const float * __restrict__ beta,
                                              __half * __restrict__ out_row, __half * __restrict__ out_fq_row, int fq_type) {
    const int nchunk = H / 8;
    const float mean = warp_sum(s) / (float) H;
    float s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        if (FULL || lane + 32 * i < nchunk) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { v[i][e] -= mean; s2 += v[i][e] * v[i][e]; }
        }
    }
    const float var = warp_sum(s2) / (float) H;
    const float scale = 1.0f / sqrtf(var + 1e-5f);
    uint4 * orow = reinterpret_cast<uint4 *>(out_row);
    uint4 * qrow = reinterpret_cast<uint4 *>(out_fq_row);
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        const bool ok = FULL || c < nchunk;
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

// K3.
template <int NCH>
__global__ void __launch_bounds__(ROWS_PER_CTA * 32)
ln_kernel(const __half * __restrict__ x, const float * __restrict__ gamma, const float * __restrict__ beta,
          int T, int H, __half * __restrict__ out, __half * __restrict__ out_fq, int fq_type) {
    pdl_launch_dependents();
    pdl_wait();
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
    ln_row_finish<NCH>(v, s, lane, H, gamma, beta, out + (size_t) row * H, out_fq ? out_fq + (size_t) row * H : nullptr, fq_type);
}

// K0.  Fused word-row gather (decoded from the table's own type) + (position + token-type-0) row + LayerNorm.
// Same chunk ownership as K3: lane l takes the 8-value chunks l + 32 i, i.e. for a q4 table 8 of a block's
// 16 payload bytes (file nibble order: byte j = e[j] | e[j+16] << 4) and the block's scale.  The position of
// a packed token inside its sentence comes from a warp-cooperative 32-way search over cu_seqlens.
template <int NCH, int WT, bool FULL>
__global__ void __launch_bounds__(ROWS_PER_CTA * 32)
embed_ln_kernel(const DevTable word, const float * __restrict__ postype,
                const float * __restrict__ gamma, const float * __restrict__ beta,
                const int32_t * __restrict__ ids, const int32_t * __restrict__ cu, int n_sent, int T, int H,
                __half * __restrict__ out, __half * __restrict__ out_fq, int fq_type, int2 * __restrict__ tok_bounds) {
    pdl_launch_dependents();
    pdl_wait();
    const int row = blockIdx.x * ROWS_PER_CTA + (threadIdx.x >> 5);
    if (row >= T) return;
    const int lane = threadIdx.x & 31;
    const int nchunk = H / 8, nb = H / 32;
    // Load order = dependency depth: the token id and the sentence search (cu_seqlens is small and cache-resident) go first and
    // overlap; then the word row (needs the id) and the position row (needs the sentence start) are requested together, and
    // nothing is consumed before every request of the row is in flight -- one DRAM round trip per row instead of one per stage.
    int id = __ldg(ids + row);
    // sentence of this packed token: largest s with cu[s] <= row, 32 pivots per round
    int lo = 0, hi = n_sent;
    while (hi - lo > 1) {
        const int step = (hi - lo + 31) / 32;
        const int idx = lo + lane * step;
        const bool le = idx < hi && __ldg(cu + idx) <= row;
        const unsigned mask = __ballot_sync(0xffffffffu, le);  // a prefix of lanes (cu is sorted, lane 0 always set)
        const int kk = 31 - __clz((int) mask);
        lo += kk * step;
        hi = min(hi, lo + step);
    }
    const int sent_t0 = __ldg(cu + lo);
    const int pos = row - sent_t0;
    // by-product for the attention kernel that packs several short sentences into one work item: [first, end) token of this
    // token's sentence
    if (tok_bounds != nullptr && lane == 0) tok_bounds[row] = make_int2(sent_t0, __ldg(cu + lo + 1));
    if (id < 0 || id >= word.rows) id = 0;                  // never read out of the table
    float v[NCH][8];
    float4 p0[NCH], p1[NCH];                                // position + token-type row
    uint4 raw[NCH];                                         // word row as stored (f32: first half; see below)
    float4 raw2[WT == W_F32 ? NCH : 1];                     // f32 table: second half of the chunk
    __half qd[NCH], qm[NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        if (!FULL) {
            raw[i] = make_uint4(0u, 0u, 0u, 0u);
            p0[i] = p1[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        qd[i] = qm[i] = __ushort_as_half(0);
        if (!FULL && c >= nchunk) continue;
        const float4 * q = reinterpret_cast<const float4 *>(postype + (size_t) pos * H) + 2 * c;
        p0[i] = __ldg(q); p1[i] = __ldg(q + 1);
        if (WT == W_F32) {
            const float4 * w = reinterpret_cast<const float4 *>(reinterpret_cast<const float *>(word.data) + (size_t) id * H) + 2 * c;
            const float4 f0 = __ldg(w);
            raw[i] = make_uint4(__float_as_uint(f0.x), __float_as_uint(f0.y), __float_as_uint(f0.z), __float_as_uint(f0.w));
            raw2[WT == W_F32 ? i : 0] = __ldg(w + 1);
        } else if (WT == W_F16) {
            raw[i] = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const __half *>(word.data) + (size_t) id * H) + c);
        } else {
            const size_t blk = (size_t) id * nb + (c >> 2);
            const uint2 u = __ldg(reinterpret_cast<const uint2 *>(word.data) + blk * 2 + (c & 1));
            raw[i].x = u.x; raw[i].y = u.y;
            qd[i] = __ldg(word.qd + blk);
            if (WT == W_Q4_1) qm[i] = __ldg(word.qm + blk);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        if (!FULL) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[i][e] = 0.f;
            if (c >= nchunk) continue;
        }
        if (WT == W_F32) {
            v[i][0] = __uint_as_float(raw[i].x); v[i][1] = __uint_as_float(raw[i].y);
            v[i][2] = __uint_as_float(raw[i].z); v[i][3] = __uint_as_float(raw[i].w);
            const float4 r2 = raw2[WT == W_F32 ? i : 0];
            v[i][4] = r2.x; v[i][5] = r2.y; v[i][6] = r2.z; v[i][7] = r2.w;
        } else if (WT == W_F16) {
            const __half2 * h = reinterpret_cast<const __half2 *>(&raw[i]);
#pragma unroll
            for (int e = 0; e < 4; ++e) { const float2 f = __half22float2(h[e]); v[i][2 * e] = f.x; v[i][2 * e + 1] = f.y; }
        } else {
            const int sh = ((c & 3) >> 1) * 4;               // values 8k .. 8k+7 of the block, k = c & 3
            const float d = __half2float(qd[i]);
            const float m = __half2float(qm[i]);
            // nibble -> float without an int-to-float conversion: OR it into the mantissa of 2^23 (exact, same value as the cast)
            const uint32_t w[2] = {raw[i].x >> sh, raw[i].y >> sh};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float f = __uint_as_float(((w[j >> 2] >> (8 * (j & 3))) & 15u) | 0x4B000000u);
                v[i][j] = WT == W_Q4_0 ? (f - 8388616.f) * d : (f - 8388608.f) * d + m;
            }
        }
        v[i][0] += p0[i].x; v[i][1] += p0[i].y; v[i][2] += p0[i].z; v[i][3] += p0[i].w;
        v[i][4] += p1[i].x; v[i][5] += p1[i].y; v[i][6] += p1[i].z; v[i][7] += p1[i].w;
#pragma unroll
        for (int e = 0; e < 8; ++e) s += v[i][e];
    }
    ln_row_finish<NCH, FULL>(v, s, lane, H, gamma, beta, out + (size_t) row * H, out_fq ? out_fq + (size_t) row * H : nullptr, fq_type);
}

// one CTA per sentence; thread = (8-column chunk, token stripe).  1024 threads keep enough 16-B loads in
// flight per SM for a grid of only n_sent CTAs; the stripe sums are combined in a fixed order (deterministic).
constexpr int POOL_THREADS = 1024;
__global__ void __launch_bounds__(POOL_THREADS)
pool_norm_kernel(const __half * __restrict__ x, const int32_t * __restrict__ cu, int H, float * __restrict__ out, int cls) {
    extern __shared__ float sm[];                       // [stripes][H] then reduction scratch
    pdl_launch_dependents();
    pdl_wait();
    const int sent = blockIdx.x;
    // CLS pooling (opt-in extension; the reference only has the masked mean): the sentence is its [CLS] row
    const int t0 = cu[sent], len = cls ? 1 : cu[sent + 1] - t0;
    const int ncol = H / 8;
    const int stripes = POOL_THREADS / ncol;
    const int c = threadIdx.x % ncol, sidx = threadIdx.x / ncol;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (sidx < stripes) {
#pragma unroll 4
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
    float pooled[2] = {0.f, 0.f};                       // H <= 1024 -> at most one column per thread ... (two kept for safety)
    int np = 0;
    for (int i = threadIdx.x; i < H; i += POOL_THREADS) {
        float s = 0.f;
        for (int k = 0; k < stripes; ++k) s += sm[k * H + i];
        s *= inv_len;
        if (np < 2) pooled[np] = s;
        ++np;
        ss += s * s;
    }
    __shared__ float red[POOL_THREADS / 32];
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < POOL_THREADS / 32; ++k) tot += red[k];
    const float inv_norm = 1.0f / sqrtf(tot);           // no epsilon (bert.cpp:1092-1094)
    np = 0;
    for (int i = threadIdx.x; i < H; i += POOL_THREADS) { out[(size_t) sent * H + i] = pooled[np < 2 ? np : 1] * inv_norm; ++np; }
}

}  // namespace

cudaError_t launch_embed_ln(const DevTable & word, const float * postype, const float * gamma, const float * beta,
                            const int32_t * ids, const int32_t * cu_seqlens, int n_sent, int T, int H,
                            __half * out, __half * out_fq, int fq_type, int2 * tok_bounds, cudaStream_t st) {
    if (T <= 0) return cudaSuccess;
    const int grid = (T + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    const int nch = (H / 8 + 31) / 32;
    const bool full = (H / 8) == 32 * nch;
#define EMBED_LAUNCH3(NCH_, WT_, FULL_) launch_pdl(embed_ln_kernel<NCH_, WT_, FULL_>, dim3(grid), dim3(ROWS_PER_CTA * 32), 0, st, \
        word, postype, gamma, beta, ids, cu_seqlens, n_sent, T, H, out, out_fq, fq_type, tok_bounds)
#define EMBED_LAUNCH2(NCH_, WT_) (full ? EMBED_LAUNCH3(NCH_, WT_, true) : EMBED_LAUNCH3(NCH_, WT_, false))
#define EMBED_LAUNCH(NCH_) (word.type == W_F32 ? EMBED_LAUNCH2(NCH_, W_F32) : word.type == W_F16 ? EMBED_LAUNCH2(NCH_, W_F16) : \
                            word.type == W_Q4_0 ? EMBED_LAUNCH2(NCH_, W_Q4_0) : EMBED_LAUNCH2(NCH_, W_Q4_1))
    if (nch <= 1) return EMBED_LAUNCH(1);
    if (nch == 2) return EMBED_LAUNCH(2);
    if (nch == 3) return EMBED_LAUNCH(3);
    return EMBED_LAUNCH(4);
#undef EMBED_LAUNCH3
#undef EMBED_LAUNCH2
#undef EMBED_LAUNCH
}

cudaError_t launch_ln(const __half * x, const float * gamma, const float * beta, int T, int H,
                      __half * out, __half * out_fq, int fq_type, cudaStream_t st) {
    if (T <= 0) return cudaSuccess;
    const int grid = (T + ROWS_PER_CTA - 1) / ROWS_PER_CTA;
    const int nch = (H / 8 + 31) / 32;
#define LN_LAUNCH(NCH_) launch_pdl(ln_kernel<NCH_>, dim3(grid), dim3(ROWS_PER_CTA * 32), 0, st, x, gamma, beta, T, H, out, out_fq, fq_type)
    if (nch <= 1) return LN_LAUNCH(1);
    if (nch == 2) return LN_LAUNCH(2);
    if (nch == 3) return LN_LAUNCH(3);
    return LN_LAUNCH(4);
#undef LN_LAUNCH
}

cudaError_t launch_pool_norm(const __half * x, const int32_t * cu_seqlens, int n_sent, int H,
                             float * out, int cls, cudaStream_t st) {
    if (n_sent <= 0) return cudaSuccess;
    if (H > 2 * POOL_THREADS) return cudaErrorInvalidValue;
    const int stripes = POOL_THREADS / (H / 8);
    const size_t smem = (size_t) (stripes > 0 ? stripes : 1) * H * sizeof(float);
    return launch_pdl(pool_norm_kernel, dim3(n_sent), dim3(POOL_THREADS), smem, st, x, cu_seqlens, H, out, cls);
}

}  // namespace b200
