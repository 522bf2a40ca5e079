// Host-side engine behind the bert.h ABI: one weight replica + workspace + stream per GPU,
// token-budget micro-batching, and the fixed kernel sequence of the encoder forward.
// Replaces the reference's per-call ggml graph build + thread-team executor
// (bert.cpp:827-1147, ggml.c:17164-17670) and its arena sizing (bert.cpp:788-810).
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
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

// One bert_encode_batch / bert_forward_batch call in flight: the calling thread submits micro-batches and waits here.
struct JobState {
    std::mutex m;
    std::condition_variable cv;
    int pending = 0;                 // micro-batches submitted and not yet scattered to the caller
    int64_t retired_tokens = 0;      // tokens of the micro-batches that are done (back-pressure for the tokenizer threads)
    bool ok = true;
};

// A micro-batch ready to run on whichever GPU takes it: per-sentence token pointers, lengths and output rows in PACK order.
struct MbTask {
    std::shared_ptr<void> keep;      // owner of the token storage (null when the caller owns it)
    std::vector<const int32_t *> tok;
    std::vector<int32_t> len;
    std::vector<float *> out;
    int tokens = 0, max_len = 0;
    JobState * job = nullptr;
};

// Persistent host threads for the tokenizer (no thread creation per call): run(n, fn) starts fn(0..n-1) on n pool threads
// and returns at once; wait() blocks until all of them have returned.
class HostPool {
public:
    ~HostPool();
    void run(int n, std::function<void(int)> fn);
    void wait();
private:
    void body(int idx, uint64_t seen);
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, done_cv_;
    std::function<void(int)> fn_;
    uint64_t gen_ = 0;
    int want_ = 0, active_ = 0;
    bool stop_ = false;
};

enum KernelKind { KK_EMBED = 0, KK_GEMM = 1, KK_ATTN = 2, KK_LN = 3, KK_POOL = 4, KK_OTHER = 5, KK_COUNT = 8 };

struct Replica {
    int device = -1;
    int num_sms = 0;
    cudaStream_t stream = nullptr;                // compute
    cudaStream_t copy_in = nullptr, copy_out = nullptr;   // H2D of the next micro-batch / D2H of the previous one overlap the forward
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
    static const int kSlots = 3;
    int32_t * d_ids[kSlots] = {};                    // [t_cap]
    int32_t * d_cu[kSlots] = {};                     // [2 (b_cap+1)]: cu_seqlens [n+1], then the span boundaries of pack_spans [n_spans+1]
    int2 * tok_bounds = nullptr;                     // [t_cap] per token: [first, end) token of its sentence (written by embed_ln)
    float * d_out[kSlots] = {};                      // [b_cap][H] f32
    int32_t * h_ids[kSlots] = {};                    // pinned mirrors
    int32_t * h_cu[kSlots] = {};
    float * h_out[kSlots] = {};
    cudaEvent_t h2d_done[kSlots] = {}, fwd_done[kSlots] = {}, slot_done[kSlots] = {};
    int slot_ready[kSlots] = {};                     // set by the D2H-done host callback (guarded by Engine::q_mu)
    alignas(8) unsigned char cb_store[kSlots][32] = {};   // argument blocks of those callbacks

    // staged job (bertx_stage / bertx_run_staged)
    std::vector<MicroBatch> job_mbs;
    std::vector<int> job_order;               // job position -> caller's sentence index
    int32_t * job_ids = nullptr;
    int32_t * job_cu = nullptr;               // concatenated per-micro-batch cu arrays (each followed by its span boundaries)
    std::vector<int> job_nspans;              // spans per micro-batch
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
    bool gemm_simt = false, attn_ref = false, attn_v2 = false, attn_v3 = false, attn_pack = true;
    bool fuse_ln_att = true;                   // also fuse the LayerNorm after the attention-output GEMM (BERT_FUSE_LN=1: ffn-out only)
    bool fuse_ln = true;                       // LayerNorm in the epilogue of the o / ffn-out GEMMs (BERT_FUSE_LN=0 disables)  // validation kernels forced by env (debug only)
    int t_cap = 0, b_cap = 0;
    bool gelu_h2 = true;                       // GELU of the FFN-in epilogue in packed half arithmetic (BERT_GELU_H2=0: f32x2 chain)
    int pooling = 0;                           // 0 = masked mean over all valid tokens (the reference, bert.cpp:1087-1089); 1 = [CLS] row (opt-in: BERT_POOLING=cls / bertx_set_pooling)
    int small_t = 256;                         // micro-batches up to this many tokens take the latency path (BERT_SMALL_T)

    ~Engine();
    bool init_devices(const ModelFile & mf);
    void init_limits();
    bool upload(Replica & r, const ModelFile & mf);
    bool alloc_workspace(Replica & r);
    void free_job(Replica & r);

    // enqueue the forward of one micro-batch whose ids / cu are already on the device
    // n_spans > 0: d_cu[n_sent + 1 ..] holds the boundaries of n_spans packed spans (pack_spans)
    bool forward(Replica & r, const int32_t * d_ids, const int32_t * d_cu, int n_sent, int T, int max_len,
                 float * d_out, int n_spans = 0);

    // order sentences by length (descending) and cut into micro-batches
    void plan(const int32_t * n_tokens, int n, std::vector<int> & order, std::vector<MicroBatch> & mbs) const;

    // ---- host runtime (pipeline.cu): persistent per-GPU worker threads fed by one queue of micro-batches
    // bert_forward_batch: token lists -> embeddings scattered to caller pointers
    bool run_tokens(const int32_t * const * tokens, const int32_t * n_tokens, int n, float * const * out);
    // bert_encode_batch: texts -> tokenizer pool -> micro-batches -> GPUs, all stages overlapped
    bool encode_texts(int n_threads, int n, const char ** texts, float ** out);
    // cut sentences (indices `order`, already sorted by length) into tasks of at most `budget` tokens
    void cut_tasks(const std::vector<int> & order, const int32_t * const * tok, const int32_t * len, float * const * out,
                   int budget, const std::shared_ptr<void> & keep, JobState * job, std::vector<MbTask> & tasks) const;
    // micro-batch token budget for a group of `tokens` tokens over nd GPUs, and the token target of the k-th chunk of a
    // bert_encode_batch call (small first chunks start the GPUs early); both functions of the counts only
    int piece_budget(int64_t tokens, int64_t nd) const;
    int64_t chunk_target(int chunk_no, int64_t nd) const;
    bool chunk_complete(int64_t chunk_tokens, int64_t target, int64_t seen_tokens, int64_t seen_sent, int64_t total_sent) const;
    static constexpr int kTokBlock = 16;          // sentences per tokenizer work item; chunks end on block boundaries
    // the micro-batches bert_encode_batch would cut for these token counts on nd GPUs (host only; bertx_plan)
    int plan_encode(const int32_t * n_tokens, int n, int nd, int32_t * mb_of_sentence, std::vector<int> & mb_tokens) const;
    void submit(std::vector<MbTask> & tasks, JobState & job);
    void wait_job(JobState & job);
    bool run_inline(Replica & r, MbTask & t);     // latency path: one micro-batch on the calling thread
    void start_workers();
    void stop_workers();
    void gpu_worker(int ri);
    bool slot_submit(Replica & r, int s, MbTask & t, bool notify);
    void slot_retire(Replica & r, int s, MbTask & t, bool ok);

    // BERT_TRACE=1: timestamped host-side events of each bert_encode_batch call, printed to stderr when the call returns
    bool tracing = false;
    std::mutex trace_mu;
    std::vector<std::pair<double, std::string>> trace_ev;
    void trace(const char * what, long a = 0, long b = 0);

    std::vector<std::thread> workers;
    std::mutex q_mu;
    std::condition_variable q_cv;
    std::deque<MbTask> queue;
    std::vector<int64_t> busy_tokens;          // per replica: tokens of the micro-batches it holds in flight (guarded by q_mu)
    bool stopping = false;
    HostPool pool;
};

bool cuda_ok(cudaError_t e, const char * what);

}  // namespace b200

// the opaque context of bert.h (reference bert.cpp:100-109)
struct bert_ctx {
    b200::Engine eng;
};
