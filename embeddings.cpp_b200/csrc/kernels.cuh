// Device-side interface of the encoder forward: plain structs + launcher prototypes.
// Every launcher enqueues on the given stream and returns the CUDA error of the launch.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <utility>

namespace b200 {

// weight types = the model file's ftype field (reference bert.cpp:499-521)
enum WType : int { W_F32 = 0, W_F16 = 1, W_Q4_0 = 2, W_Q4_1 = 3 };

// epilogue flags of the GEMM
// (EPI_GELU_H2, together with EPI_GELU: the tcgen05 kernel evaluates GELU in packed half arithmetic)
enum : int { EPI_BIAS = 1, EPI_GELU = 2, EPI_RESIDUAL = 4, EPI_FAKEQUANT = 8, EPI_LN = 16, EPI_GELU_H2 = 32 };

// A dense weight matrix W[N][K] (K contiguous = the file's ne0, reference bert.cpp:611-623)
// resident in HBM in its GEMM-ready form:
//   f32 / f16 files : `w16` fp16 [N][K] row-major (+ TMA map, box 64 x BN/2: each CTA of a pair loads half)
//   q4_0 / q4_1     : 4-bit planes tiled by 64-wide k-stage so that one stage of any row
//                     range is contiguous:
//                       qnib  [K/64][N][2] x 16 B   (nibble order: see repack_q4 in upload.cu)
//                       qd    [K/64][N][2] fp16     block scales d
//                       qm    [K/64][N][2] fp16     block mins m (q4_1 only)
//                     HBM holds 4.5 (q4_0) / 5 (q4_1) bits per weight, as in the file.
struct DevWeight {
    int type = W_F16;          // W_F16 (covers f32 files cast to f16) / W_Q4_0 / W_Q4_1
    int N = 0, K = 0;
    const __half * w16 = nullptr;
    const uint4 * qnib = nullptr;
    const __half * qd = nullptr;
    const __half * qm = nullptr;
    const float * bias = nullptr;   // [N] f32
    int bn = 0;                     // N-tile the TMA map was encoded for
    CUtensorMap tmap;               // B operand map (f16 mode)
};

struct GemmArgs {
    const CUtensorMap * tmap_a;     // host pointer to the A map (copied into kernel params)
    const CUtensorMap * tmap_out;   // output map, box {64 cols, 32 rows} (epilogue TMA stores)
    const CUtensorMap * tmap_res;   // residual map, same geometry (epilogue TMA prefetch) or null
    const __half * a;               // A [M][K] fp16 (used by the SIMT validation kernel only)
    const DevWeight * w;
    const __half * residual;        // [M][N] fp16 or null
    __half * out;                   // [M][N] fp16
    int M;                          // valid rows
    int epilogue;                   // EPI_* flags
    int fq_type;                    // W_Q4_0 -> q8_0 rule, W_Q4_1 -> q8_1 rule (EPI_FAKEQUANT)
    // EPI_LN (with EPI_BIAS | EPI_RESIDUAL): out = LayerNorm(acc + bias + residual) * gamma + beta; with
    // EPI_FAKEQUANT the q8 fake-quantised twin goes through tmap_out2.  See gemm_ln_fusable().
    const CUtensorMap * tmap_out2 = nullptr;
    const float * ln_gamma = nullptr;
    const float * ln_beta = nullptr;
    // q4 weights only (their HBM layout does not depend on the N-tile): use this N-tile instead of w->bn.
    // Small batches pick 64 so that 4x as many CTA pairs stream the weights (latency path).
    int bn_override = 0;
};

int gemm_pick_bn(int N);                 // N-tile (256/192/128/64) used for a weight with N rows
int gemm_pick_bn_ln(int N);              // the same for a GEMM followed by residual + LayerNorm (prefers the fusable widths 256 / 128)
cudaError_t launch_gemm_tc(const GemmArgs & g, int num_sms, cudaStream_t st);
bool gemm_ln_fusable(const DevWeight & w);   // can the LayerNorm after this GEMM run in its epilogue?
cudaError_t launch_gemm_simt(const GemmArgs & g, cudaStream_t st);      // validation kernel

// Embedding table as stored for the gather kernel (K0).
struct DevTable {
    int type = W_F32;               // file type of the word table
    int rows = 0, H = 0;
    const void * data = nullptr;    // f32 [rows][H] | f16 [rows][H] | q4 nibbles [rows][H/32] x 16 B (file order)
    const __half * qd = nullptr;    // [rows][H/32]
    const __half * qm = nullptr;
};

// K0: x[t] = LN(word[ids[t]] + postype[pos(t)]) * g + b  -> fp16 (and its q8 fake-quant twin)
cudaError_t launch_embed_ln(const DevTable & word, const float * postype /*[n_max][H] f32*/,
                            const float * gamma, const float * beta,
                            const int32_t * ids, const int32_t * cu_seqlens, int n_sent, int T, int H,
                            __half * out, __half * out_fq, int fq_type, int2 * tok_bounds, cudaStream_t st);

// K3: y = LN(x) * g + b, x already holds GEMM + bias + residual (fp16)
cudaError_t launch_ln(const __half * x, const float * gamma, const float * beta, int T, int H,
                      __half * out, __half * out_fq, int fq_type, cudaStream_t st);

// K2: per (sentence, head) softmax(Q K^T / sqrt(d)) V over the sentence's own tokens.
// qkv [T][3H] fp16 (Q | K | V, head h at columns h*d), out [T][H] fp16.
// All production kernels are tcgen05 code (the round-1 mma.sync kernel for head size 32 is gone): Q / K / V tiles come by TMA
// from the packed qkv buffer through the two maps made by make_attention_tmaps (box 64 x 128 for Q, 64 x 64 for K and V).
// attention_tc.cu: head size 64.
bool make_attention_tmaps(CUtensorMap * tm_q, CUtensorMap * tm_kv, const __half * qkv, uint64_t rows, int H);
cudaError_t launch_attention_tc(const CUtensorMap * tm_q, const CUtensorMap * tm_kv, const int32_t * cu_seqlens,
                                int n_sent, int max_len, int n_head, int d_head, __half * out, int fq_type,
                                int num_sms, cudaStream_t st, const int2 * tok_bounds = nullptr);
// round-2 kernel (attention_tc2.cu): head sizes 64 and 32 (two heads per work item), eight softmax warps per CTA; same maps.
// With tok_bounds (per token: [first, end) token of its sentence, written by launch_embed_ln) `cu_seqlens` may hold SPANS of
// several whole sentences of at most 128 tokens in total (pack_spans): one work item then serves all of them, every query row
// masked to the keys of its own sentence.
cudaError_t launch_attention_tc2(const CUtensorMap * tm_q, const CUtensorMap * tm_kv, const int32_t * cu_seqlens,
                                 int n_sent, int max_len, int n_head, int d_head, __half * out, int fq_type,
                                 int num_sms, cudaStream_t st, const int2 * tok_bounds = nullptr);
// round-2 kernel for head size 64 (attention_tc3.cu): three S buffers in TMEM, separate K / V rings, software-pipelined softmax
cudaError_t launch_attention_tc3(const CUtensorMap * tm_q, const CUtensorMap * tm_kv, const int32_t * cu_seqlens,
                                 int n_sent, int max_len, int n_head, int d_head, __half * out, int fq_type,
                                 int num_sms, cudaStream_t st);
void attn3_prof_fetch(unsigned long long * out);      // diagnostic counters of a -DBERT_ATTN3_PROF=1 build
// Greedy packing of consecutive sentences into spans of at most 128 tokens (a longer sentence is a span of its own):
// span_cu[0 .. n_spans] = token offsets of the span boundaries; returns n_spans.  Host code.
int pack_spans(const int32_t * cu_seqlens, int n_sent, int32_t * span_cu);
cudaError_t launch_attention_ref(const __half * qkv, const int32_t * cu_seqlens, int n_sent, int max_len,
                                 int n_head, int d_head, __half * out, int fq_type, cudaStream_t st);

// K4: masked mean over the sentence's tokens + L2 normalise -> f32 [n_sent][H]
cudaError_t launch_pool_norm(const __half * x, const int32_t * cu_seqlens, int n_sent, int H,
                             float * out, int cls, cudaStream_t st);

// Every kernel of the forward is launched through this helper: with programmatic dependent launch (PDL) the
// next kernel's launch latency and prologue (barrier init, TMEM allocation, descriptor prefetch) overlap the tail
// of the current one; each kernel executes griddepcontrol.wait before its first dependent global access.
// BERT_PDL=0 launches with plain stream order.
inline bool pdl_enabled() {
    static const bool on = [] { const char * e = getenv("BERT_PDL"); return !(e && atoi(e) == 0); }();
    return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args &&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// upload-time transforms (device side)
cudaError_t launch_f32_to_f16(const float * in, __half * out, size_t n, cudaStream_t st);
// rows [n_offset, n_offset+N) of a plane set holding N_total rows (lets Q|K|V share one GEMM)
cudaError_t launch_repack_q4(const uint8_t * raw, int type, int N, int K, int n_offset, int N_total,
                             uint4 * qnib, __half * qd, __half * qm, cudaStream_t st);
cudaError_t launch_split_q4_table(const uint8_t * raw, int type, int rows, int H,
                                  uint4 * qnib, __half * qd, __half * qm, cudaStream_t st);
cudaError_t launch_dequant_q4_f32(const DevWeight & w, float * out /*[N][K]*/, cudaStream_t st);

// TMA map helper (host): 2-D fp16 row-major [rows][cols], box {64 cols, box_rows}, 128B swizzle
bool make_tmap_f16_2d(CUtensorMap * map, const void * base, uint64_t rows, uint64_t cols, uint32_t box_rows);

// shared by kernels: q8 fake-quant of 32 consecutive values held by one thread
__device__ __forceinline__ void fake_quant32(float (&v)[32], int fq_type) {
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) amax = fmaxf(amax, fabsf(v[i]));
    // reference ggml.c:1146-1300 (q8_0) / 1425-1480 (q8_1): d = amax/127, id = 127/amax, rint
    const float d  = amax / 127.f;
    const float id = amax != 0.f ? 127.f / amax : 0.f;
    const float dd = fq_type == W_Q4_0 ? __half2float(__float2half_rn(d)) : d;
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = rintf(v[i] * id) * dd;
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

}  // namespace b200
