// Host-side engine behind the bert.h ABI: one weight replica + workspace + stream per GPU,
// token-budget micro-batching, and the fixed kernel sequence of the encoder forward.
// Replaces the reference's per-call ggml graph build + thread-team executor
// (bert.cpp:827-1147, ggml.c:17164-17670) and its arena sizing (bert.cpp:788-810).
#pragma once
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "model_file.h"
#include "tokenizer.h"

namespace b200 {

struct LayerDev {
    DevWeight qkv, o, ffi, ffo;
    const float * ln_att_g = nullptr, * ln_att_b = nullptr, * ln_out_g = nullptr, * ln_out_b = nullptr;
};

// A micro-batch: sentences [first, first+count) of an ordered job, `tokens` packed ids.
struct MicroBatch {
    int first = 0, count = 0;
    int tokens = 0, max_len = 0;
    size_t tok_offset = 0;          // offset of its first id in the job's packed id array
};

enum KernelKind { KK_EMBED = 0, KK_GEMM = 1, KK_ATTN = 2, KK_LN = 3, KK_POOL = 4, KK_OTHER = 5, KK_COUNT = 8 };

struct Replica {
    int device = -1;
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    std::vector<void *> allocs;      // every cudaMalloc of this replica
    std::vector<void *> pinned;

    // weights
    DevTable word;
    const float * postype = nullptr;          // [n_max_tokens][H] f32 = position + token_type[0]
    const float * ln_e_g = nullptr, * ln_e_b = nullptr;
    std::vector<LayerDev> layers;

    // workspace (fp16 unless noted), sized for t_cap tokens / b_cap sentences
    int t_cap = 0, b_cap = 0;
    __half * x = nullptr, * xq = nullptr, * qkv = nullptr, * ctx = nullptr, * y = nullptr,
           * x1 = nullptr, * x1q = nullptr, * mid = nullptr;
    CUtensorMap tm_xin, tm_ctx, tm_x1in, tm_mid;     // A-operand maps (box 64 x 128)
    CUtensorMap to_qkv, to_y, to_mid;                // output maps (box 64 x 32, epilogue TMA stores)
    CUtensorMap ta_q, ta_kv;                         // attention maps over qkv (box 64 x 128 / 64 x 64)
    CUtensorMap to_x, to_xq, to_x1, to_x1q;          // outputs of the GEMMs with fused LayerNorm (box 64 x 32)
    CUtensorMap tr_x, tr_x1;                         // residual maps (box 64 x 32, epilogue TMA loads)
    static const int kSlots = 2;
    int32_t * d_ids[kSlots] = {nullptr, nullptr};    // [t_cap]
    int32_t * d_cu[kSlots] = {nullptr, nullptr};     // [b_cap+1]
    float * d_out[kSlots] = {nullptr, nullptr};      // [b_cap][H] f32
    int32_t * h_ids[kSlots] = {nullptr, nullptr};    // pinned mirrors
    int32_t * h_cu[kSlots] = {nullptr, nullptr};
    float * h_out[kSlots] = {nullptr, nullptr};
    cudaEvent_t slot_done[kSlots] = {nullptr, nullptr};

    // staged job (bertx_stage / bertx_run_staged)
    std::vector<MicroBatch> job_mbs;
    std::vector<int> job_order;               // job position -> caller's sentence index
    int32_t * job_ids = nullptr;
    int32_t * job_cu = nullptr;               // concatenated per-micro-batch cu arrays
    float * job_out = nullptr;                // [n_sent][H]
    int job_sent = 0;
    std::vector<size_t> job_cu_off;

    // optional per-launch profiling
    bool profiling = false;
    struct Span { int kind; cudaEvent_t a, b; };
    std::vector<Span> spans;
};

struct Engine {
    HParams hp;
    Tokenizer tok;
    std::vector<Replica> reps;
    std::mutex mu;                             // a context is not re-entrant (bert.h note)
    std::atomic<int64_t> launches{0};
    int fq_type = -1;                          // W_Q4_0 / W_Q4_1 when activation fake-quant is on
    bool gemm_simt = false, attn_ref = false, attn_mma = false;
    bool fuse_ln_att = true;                   // also fuse the LayerNorm after the attention-output GEMM (BERT_FUSE_LN=1: ffn-out only)
    bool fuse_ln = true;                       // LayerNorm in the epilogue of the o / ffn-out GEMMs (BERT_FUSE_LN=0 disables)  // validation kernels forced by env (debug only)
    int t_cap = 0, b_cap = 0;
    bool gelu_h2 = true;                       // GELU of the FFN-in epilogue in packed half arithmetic (BERT_GELU_H2=0: f32x2 chain)
    int pooling = 0;                           // 0 = masked mean over all valid tokens (the reference, bert.cpp:1087-1089); 1 = [CLS] row (opt-in: BERT_POOLING=cls / bertx_set_pooling)
    int small_t = 256;                         // micro-batches up to this many tokens take the latency path (BERT_SMALL_T)

    ~Engine();
    bool init_devices(const ModelFile & mf);
    bool upload(Replica & r, const ModelFile & mf);
    bool alloc_workspace(Replica & r);
    void free_job(Replica & r);

    // enqueue the forward of one micro-batch whose ids / cu are already on the device
    bool forward(Replica & r, const int32_t * d_ids, const int32_t * d_cu, int n_sent, int T, int max_len,
                 float * d_out);

    // order sentences by length (descending) and cut into micro-batches
    void plan(const int32_t * n_tokens, int n, std::vector<int> & order, std::vector<MicroBatch> & mbs) const;

    // full host path: token lists -> embeddings scattered to caller pointers
    bool run_tokens(const int32_t * const * tokens, const int32_t * n_tokens, int n, float * const * out);
    bool run_on_replica(Replica & r, const std::vector<int> & order, const std::vector<MicroBatch> & mbs,
                        const std::vector<int> & mine, const int32_t * const * tokens, const int32_t * n_tokens,
                        float * const * out);
};

bool cuda_ok(cudaError_t e, const char * what);

}  // namespace b200

// the opaque context of bert.h (reference bert.cpp:100-109)
struct bert_ctx {
    b200::Engine eng;
};
