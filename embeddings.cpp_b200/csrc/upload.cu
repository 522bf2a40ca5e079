// Upload-time weight transforms (run once per model load, on the device) and the TMA
// tensor-map helper.  The raw tensor bytes of the model file (reference layout:
// block_q4_0 = fp16 d + 16 nibble bytes, block_q4_1 = fp16 d + fp16 m + 16 nibble bytes,
// ggml.c:893-907; byte j holds element j in its low and element j+16 in its high nibble,
// ggml.c:1531-1541) are copied to HBM as they are and re-laid-out here; no value changes.
#include "kernels.cuh"

namespace b200 {

namespace {

__global__ void f32_to_f16_kernel(const float * __restrict__ in, __half * __restrict__ out, size_t n) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        out[i] = __float2half_rn(in[i]);
    }
}

__device__ __forceinline__ void read_block(const uint8_t * raw, int type, size_t blk,
                                           uint8_t (&qs)[16], uint16_t & d, uint16_t & m) {
    const int bs = type == W_Q4_0 ? 18 : 20;
    const uint8_t * b = raw + blk * bs;
    d = (uint16_t) b[0] | ((uint16_t) b[1] << 8);
    m = 0;
    int off = 2;
    if (type == W_Q4_1) { m = (uint16_t) b[2] | ((uint16_t) b[3] << 8); off = 4; }
#pragma unroll
    for (int j = 0; j < 16; ++j) qs[j] = b[off + j];
}

// GEMM weight planes.  One thread per 32-weight block of W[N][K].
// Output nibble order inside a block (4 words, word k = elements 8k..8k+7):
//   element 8k+2i   at bits [4i, 4i+4),  element 8k+2i+1 at bits [16+4i, 16+4i+4)
// so that ((w >> 4i) & 0x000F000F) | 0x64006400 is the fp16 pair (1024+e, 1024+e') of two
// ADJACENT weights (see dq_pair in gemm_tc.cu).
__global__ void repack_q4_kernel(const uint8_t * __restrict__ raw, int type, int N, int K, int n_offset, int N_total,
                                 uint4 * __restrict__ qnib, __half * __restrict__ qd, __half * __restrict__ qm) {
    const size_t nblk = (size_t) N * (K / 32);
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk) return;
    const int kb32 = (int) (i % (K / 32));
    const int n = (int) (i / (K / 32));
    uint8_t qs[16]; uint16_t d, m;
    read_block(raw, type, i, qs, d, m);
    uint32_t e[32];
#pragma unroll
    for (int j = 0; j < 16; ++j) { e[j] = qs[j] & 15; e[j + 16] = qs[j] >> 4; }
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t x = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) x |= (e[8 * k + 2 * q] << (4 * q)) | (e[8 * k + 2 * q + 1] << (16 + 4 * q));
        w[k] = x;
    }
    const size_t dst = ((size_t) (kb32 >> 1) * N_total + n_offset + n) * 2 + (kb32 & 1);       // [K/64][N_total][2]
    qnib[dst] = make_uint4(w[0], w[1], w[2], w[3]);
    qd[dst] = __ushort_as_half(d);
    if (type == W_Q4_1) qm[dst] = __ushort_as_half(m);
}

// Embedding-table planes: nibbles stay in file order, [rows][H/32] x 16 B (+ d, m planes).
__global__ void split_q4_table_kernel(const uint8_t * __restrict__ raw, int type, size_t nblk,
                                      uint4 * __restrict__ qnib, __half * __restrict__ qd, __half * __restrict__ qm) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nblk) return;
    uint8_t qs[16]; uint16_t d, m;
    read_block(raw, type, i, qs, d, m);
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        w[k] = (uint32_t) qs[4 * k] | ((uint32_t) qs[4 * k + 1] << 8) | ((uint32_t) qs[4 * k + 2] << 16) |
               ((uint32_t) qs[4 * k + 3] << 24);
    }
    qnib[i] = make_uint4(w[0], w[1], w[2], w[3]);
    qd[i] = __ushort_as_half(d);
    if (type == W_Q4_1) qm[i] = __ushort_as_half(m);
}

// weight element (n, k) in f32, decoded from whatever form the GEMM reads (validation path)
__device__ __forceinline__ float weight_at(const DevWeight & w, int n, int k) {
    if (w.type == W_F16) return __half2float(w.w16[(size_t) n * w.K + k]);
    const size_t blk = ((size_t) (k >> 6) * w.N + n) * 2 + ((k >> 5) & 1);
    const int e = k & 31;
    const uint4 u = w.qnib[blk];
    const uint32_t words[4] = {u.x, u.y, u.z, u.w};
    const uint32_t q = (words[e >> 3] >> (4 * ((e & 7) >> 1) + 16 * (e & 1))) & 15;
    const float d = __half2float(w.qd[blk]);
    if (w.type == W_Q4_0) return (float) ((int) q - 8) * d;
    return (float) q * d + __half2float(w.qm[blk]);
}

__global__ void dequant_f32_kernel(const DevWeight w, float * __restrict__ out) {
    const size_t n_el = (size_t) w.N * w.K;
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n_el; i += (size_t) gridDim.x * blockDim.x) {
        out[i] = weight_at(w, (int) (i / w.K), (int) (i % w.K));
    }
}

__device__ __forceinline__ float gelu_ref_numerics_v(float x) {
    const float f = __half2float(__float2half_rn(x));
    const float u = 0.79788456080286535587989211986876f * f * (1.0f + 0.044715f * f * f);
    return __half2float(__float2half_rn(0.5f * f * (1.0f + tanhf(u))));
}

// Validation GEMM: one thread per (row, 32-column block); plain f32 FMA loop, same epilogue
// semantics as gemm_tc.cu but written independently (tanhf, scalar decode).
__global__ void gemm_simt_kernel(const __half * __restrict__ a, const DevWeight w,
                                 const __half * __restrict__ residual, __half * __restrict__ out,
                                 int M, int epilogue, int fq_type) {
    const int row = blockIdx.x;                                  // rows on x: M may exceed the 65535 limit of y
    const int cb = blockIdx.y * blockDim.x + threadIdx.x;        // 32-column block
    if (row >= M || cb >= w.N / 32) return;
    float v[32];
    for (int j = 0; j < 32; ++j) v[j] = 0.f;
    const __half * ar = a + (size_t) row * w.K;
    for (int k = 0; k < w.K; ++k) {
        const float x = __half2float(ar[k]);
        for (int j = 0; j < 32; ++j) v[j] += x * weight_at(w, cb * 32 + j, k);
    }
    for (int j = 0; j < 32; ++j) {
        const int c = cb * 32 + j;
        if (epilogue & EPI_BIAS) v[j] += w.bias[c];
        if (epilogue & EPI_GELU) v[j] = gelu_ref_numerics_v(v[j]);
        if (epilogue & EPI_RESIDUAL) v[j] += __half2float(residual[(size_t) row * w.N + c]);
    }
    if (epilogue & EPI_FAKEQUANT) fake_quant32(v, fq_type);
    for (int j = 0; j < 32; ++j) out[(size_t) row * w.N + cb * 32 + j] = __float2half_rn(v[j]);
}

// Validation attention: one warp per (query, head, sentence); scores in shared memory.
__global__ void attention_ref_kernel(const __half * __restrict__ qkv, const int32_t * __restrict__ cu,
                                     int H, int D, float scale, __half * __restrict__ out, int fq_type, int max_len) {
    extern __shared__ float sc[];                       // [warps][max_len]
    const int sent = blockIdx.z, head = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qi = blockIdx.x * (blockDim.x >> 5) + warp;
    const int t0 = cu[sent], len = cu[sent + 1] - t0;
    if (qi >= len) return;
    float * s = sc + (size_t) warp * max_len;
    const int ld = 3 * H;
    const __half * q = qkv + (size_t) (t0 + qi) * ld + head * D;
    float mx = -INFINITY;
    for (int j = lane; j < len; j += 32) {
        const __half * k = qkv + (size_t) (t0 + j) * ld + H + head * D;
        float acc = 0.f;
        for (int e = 0; e < D; ++e) acc += __half2float(q[e]) * __half2float(k[e]);
        acc *= scale;
        s[j] = acc;
        mx = fmaxf(mx, acc);
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int j = lane; j < len; j += 32) { const float p = expf(s[j] - mx); s[j] = p; sum += p; }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncwarp();
    const float inv = 1.f / sum;
    for (int e0 = 0; e0 < D; e0 += 32) {                // one 32-wide block of the head per pass
        const int e = e0 + lane;
        float acc = 0.f;
        for (int j = 0; j < len; ++j) acc += s[j] * __half2float(qkv[(size_t) (t0 + j) * ld + 2 * H + head * D + e]);
        acc *= inv;
        if (fq_type == W_Q4_0 || fq_type == W_Q4_1) {
            float amax = fabsf(acc);
            for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
            const float d = amax / 127.f;
            const float id = amax != 0.f ? 127.f / amax : 0.f;
            const float dd = fq_type == W_Q4_0 ? __half2float(__float2half_rn(d)) : d;
            acc = rintf(acc * id) * dd;
        }
        out[(size_t) (t0 + qi) * H + head * D + e] = __float2half_rn(acc);
    }
}

}  // namespace

cudaError_t launch_f32_to_f16(const float * in, __half * out, size_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    f32_to_f16_kernel<<<1024, 256, 0, st>>>(in, out, n);
    return cudaGetLastError();
}

cudaError_t launch_repack_q4(const uint8_t * raw, int type, int N, int K, int n_offset, int N_total,
                             uint4 * qnib, __half * qd, __half * qm, cudaStream_t st) {
    const size_t nblk = (size_t) N * (K / 32);
    repack_q4_kernel<<<(unsigned) ((nblk + 127) / 128), 128, 0, st>>>(raw, type, N, K, n_offset, N_total, qnib, qd, qm);
    return cudaGetLastError();
}

cudaError_t launch_split_q4_table(const uint8_t * raw, int type, int rows, int H,
                                  uint4 * qnib, __half * qd, __half * qm, cudaStream_t st) {
    const size_t nblk = (size_t) rows * (H / 32);
    split_q4_table_kernel<<<(unsigned) ((nblk + 127) / 128), 128, 0, st>>>(raw, type, nblk, qnib, qd, qm);
    return cudaGetLastError();
}

cudaError_t launch_dequant_q4_f32(const DevWeight & w, float * out, cudaStream_t st) {
    dequant_f32_kernel<<<1024, 256, 0, st>>>(w, out);
    return cudaGetLastError();
}

cudaError_t launch_gemm_simt(const GemmArgs & g, cudaStream_t st) {
    if (g.M <= 0) return cudaSuccess;
    const DevWeight & w = *g.w;
    dim3 grid(g.M, (w.N / 32 + 63) / 64);
    gemm_simt_kernel<<<grid, 64, 0, st>>>(g.a, w, g.residual, g.out, g.M, g.epilogue, g.fq_type);
    return cudaGetLastError();
}

cudaError_t launch_attention_ref(const __half * qkv, const int32_t * cu_seqlens, int n_sent, int max_len,
                                 int n_head, int d_head, __half * out, int fq_type, cudaStream_t st) {
    if (n_sent <= 0) return cudaSuccess;
    const int H = n_head * d_head;
    const int warps = 4;
    dim3 grid((max_len + warps - 1) / warps, n_head, n_sent);
    attention_ref_kernel<<<grid, warps * 32, (size_t) warps * max_len * sizeof(float), st>>>(
        qkv, cu_seqlens, H, d_head, 1.0f / sqrtf((float) d_head), out, fq_type, max_len);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------- TMA map
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void * p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess) {
            fn = (EncodeTiledFn) p;
        }
    }
    return fn;
}

bool make_tmap_f16_2d(CUtensorMap * map, const void * base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * 2};           // bytes, dim 1
    const cuuint32_t box[2] = {64, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

}  // namespace b200
