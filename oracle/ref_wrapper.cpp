// TEST INFRASTRUCTURE ONLY (oracle/): builds the UNMODIFIED reference engine into
// oracle/_ref/libbert_ref.so by including its translation unit where it lies under
// /root/reference (never copied into this repo), and adds ONE extra entry point.
//
// Why the extra entry point: the reference sizes its ggml arena from a 4-token probe
// (bert.cpp:771-781) times 1.15 * 512 (bert.cpp:788-791), which ignores the
// O(L^2 * n_head) attention tensors; for L > ~384 ggml's pool is exhausted and the
// library dereferences NULL.  oracle_ref_reserve() pre-sizes
// ctx->buf_compute and ctx->max_batch_n so that bert_forward_batch (bert.cpp:845-853)
// never enters bert_resize_ctx.  No arithmetic is touched: every kernel that runs is
// the reference's own ggml code.
#include "bert.cpp"   // -I/root/reference  (reference translation unit, read-only)

extern "C" {

// Make the compute arena `bytes` large and declare it good for `max_batch_n` sentences.
__attribute__((visibility("default")))
void oracle_ref_reserve(bert_ctx * ctx, int64_t bytes, int32_t max_batch_n) {
    if ((int64_t) ctx->buf_compute.size < bytes) {
        ctx->buf_compute.resize((size_t) bytes);
    }
    ctx->max_batch_n = max_batch_n;
    // keep bert_resize_ctx's bookkeeping consistent (bert.cpp:804-809)
    ctx->mem_per_input = bytes / (max_batch_n > 0 ? max_batch_n : 1);
}

// Arena bytes the reference measured per token at load (bert.cpp:1139-1141).
__attribute__((visibility("default")))
int64_t oracle_ref_mem_per_token(bert_ctx * ctx) {
    return (int64_t) ctx->mem_per_token;
}

__attribute__((visibility("default")))
int32_t oracle_ref_n_head(bert_ctx * ctx) { return ctx->model.hparams.n_head; }

__attribute__((visibility("default")))
int32_t oracle_ref_n_layer(bert_ctx * ctx) { return ctx->model.hparams.n_layer; }

}
