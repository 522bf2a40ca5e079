/*
 * bert_b200.h -- extension entry points of libbert.so beyond the reference ABI (bert.h).
 *
 * Plain C ABI (pointers and sizes only).  These exist for measurement and kernel-level
 * parity tests; a consumer of the reference never needs them.  The reference has no
 * counterpart except where cited.
 */
#ifndef BERT_B200_H
#define BERT_B200_H

#include "bert.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Parse the model file and build the tokenizer only -- no GPU is touched.  bert_tokenize,
 * bert_vocab_id_to_token, bert_n_embd, bert_n_max_tokens, bertx_tokenize_batch work on the
 * returned context; every forward / encode call fails loudly (there is no CPU path). */
BERT_API struct bert_ctx * bertx_load_host_only(const char * fname);

/* Number of GPUs this context holds a weight replica on (BERT_CUDA_DEVICES / all visible). */
BERT_API int32_t bertx_n_devices(struct bert_ctx * ctx);

/* Model header fields (reference bert.cpp:453-459), index: 0 n_vocab, 1 n_max_tokens,
 * 2 n_embd, 3 n_intermediate, 4 n_head, 5 n_layer, 6 ftype. */
BERT_API int32_t bertx_hparam(struct bert_ctx * ctx, int32_t index);

/* Tokenise n_inputs texts with the host tokenizer (threaded); ids are written packed
 * back-to-back into tokens_out (capacity cap ids), lengths into n_tokens_out.
 * Returns total ids written or -1 if cap is too small.  Same ids as bert_tokenize. */
BERT_API int64_t bertx_tokenize_batch(struct bert_ctx * ctx, int32_t n_threads, int32_t n_inputs,
                                      const char ** texts, bert_vocab_id * tokens_out, int64_t cap,
                                      int32_t * n_tokens_out);

/* The micro-batches one bert_encode_batch call would cut for sentences of these token counts on n_devices GPUs (the
 * scheduler of reference bert.cpp:1394-1443 re-designed: deterministic chunks on 16-sentence boundaries, sorted by
 * length, packed to a token budget; any GPU takes the next micro-batch).  Works on a host-only context.  Writes the
 * micro-batch index of every sentence and the token count of every micro-batch (capacity cap); returns their number. */
BERT_API int32_t bertx_plan(struct bert_ctx * ctx, int32_t n_devices, int32_t n_sent, const int32_t * n_tokens,
                            int32_t * mb_of_sentence, int32_t * mb_tokens, int32_t cap);

/* The attention work units of one micro-batch (host only, no context needed): consecutive sentences whose token counts
 * n_tokens[0..n_sent) add up to at most 128 share a span, a longer sentence is a span of its own (the packed spans of
 * attention_tc.cu / attention_tc2.cu; no counterpart in the reference, whose ggml_soft_max pads every sentence to the
 * batch maximum, bert.cpp:1018-1036).  Writes the span boundaries as token offsets, span_cu[0..n_spans], capacity
 * n_sent + 1; returns n_spans. */
BERT_API int32_t bertx_pack_spans(int32_t n_sent, const int32_t * n_tokens, int32_t * span_cu);

/* Stage a whole job on device `dev` (index into the context's device list): packed ids
 * + lengths are copied to HBM and cut into micro-batches.  Returns 0 on success. */
BERT_API int32_t bertx_stage(struct bert_ctx * ctx, int32_t dev, int32_t n_sent,
                             const bert_vocab_id * tokens_packed, const int32_t * n_tokens);

/* Run the encoder forward over the staged job `reps` times with inputs already resident
 * in HBM.  The timed region (CUDA events on the launch stream) covers all kernels of all
 * micro-batches; *ms_total receives the elapsed milliseconds for all reps.  If out is
 * non-NULL the [n_sent, n_embd] f32 embeddings of the last rep are copied to it afterwards
 * (outside the timed region), in staging order.  Returns 0 on success. */
BERT_API int32_t bertx_run_staged(struct bert_ctx * ctx, int32_t dev, int32_t reps,
                                  float * ms_total, float * out);

/* Same job, but every kernel class is bracketed by its own events (serialised):
 * ms_by_kind[k] = total ms, launches_by_kind[k] = launch count, for k in
 * 0 embed_ln, 1 gemm, 2 attention, 3 add_ln, 4 pool_norm, 5 other.  Arrays of 8. */
BERT_API int32_t bertx_profile_staged(struct bert_ctx * ctx, int32_t dev,
                                      float * ms_by_kind, int64_t * launches_by_kind);

/* Pooling of the final hidden states: 0 = masked mean over all valid tokens (the reference's only mode,
 * bert.cpp:1087-1089; default), 1 = the [CLS] row (what the BGE model cards specify; opt-in extension, also
 * BERT_POOLING=cls at load).  Both are followed by the L2 normalisation.  Returns the previous mode, -1 on a bad argument. */
BERT_API int32_t bertx_set_pooling(struct bert_ctx * ctx, int32_t mode);

/* Kernel launches issued by this context since load (all devices). */
BERT_API int64_t bertx_kernel_launches(struct bert_ctx * ctx);

/* Kernel self-tests: run the production kernel and a simple SIMT reference kernel on the
 * same random device data and report the largest absolute difference and the largest
 * reference magnitude.  wtype: 1 f16, 2 q4_0, 3 q4_1.  epilogue bit0 bias, bit1 GELU,
 * bit2 residual, bit3 q8 fake-quant of the output.  Return 0 on success (ran), <0 error. */
BERT_API int32_t bertx_selftest_gemm(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue,
                                     int32_t seed, float * max_abs_err, float * max_ref);
/* Time `reps` back-to-back launches of the tcgen05 GEMM on random data (CUDA events). */
BERT_API int32_t bertx_bench_gemm(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue, int32_t reps,
                                  float * ms_total);
BERT_API int32_t bertx_selftest_attention(int32_t n_sent, int32_t max_len, int32_t n_head, int32_t d_head,
                                          int32_t seed, float * max_abs_err, float * max_ref);

/* Debug aid for a per-kernel check against an independent implementation: the production attention kernel on seeded random
 * fp16 q / k / v for sentences of lens[0..n_sent) tokens (flags: the kernel-choice bits of bertx_selftest_attention's seed);
 * the inputs ([T][3 n_head d_head] fp16 bit patterns, q | k | v per token as the QKV GEMM writes them) and the kernel's
 * output ([T][n_head d_head] fp16 bit patterns) are copied to the caller's host arrays.  Replaces nothing in the reference;
 * the arithmetic checked is bert.cpp:1018-1036 (scale, masked soft-max, weighted sum of V) per sentence and head. */
BERT_API int32_t bertx_debug_attention(int32_t n_sent, const int32_t * lens, int32_t n_head, int32_t d_head, int32_t seed,
                                       int32_t flags, uint16_t * qkv_out, uint16_t * out);

/* Time `reps` back-to-back launches of the production attention kernel over n_sent sentences of
 * exactly len tokens (random q/k/v, CUDA events). */
BERT_API int32_t bertx_bench_attention(int32_t n_sent, int32_t len, int32_t n_head, int32_t d_head, int32_t reps,
                                       float * ms_total);

/* Debug aid for kernel bisects: one GEMM of ANY size on device-generated operands; the SIMT validation kernel's output
 * and `runs` outputs of the tcgen05 kernel ([M][N] fp16 bit patterns each) are copied to host arrays the caller owns,
 * ms_tc[runs] receives the kernel times.  out_ref NULL: no validation run.  out_ref2 / out_tc2 (may be NULL): the q8 twin
 * a fused-LayerNorm epilogue (bit4) writes; a_out / w_out (may be NULL): the A operand ([M][K] fp16 bits) and the weights as
 * the kernels decode them ([N][K] f32). */
BERT_API int32_t bertx_debug_gemm(int32_t M, int32_t N, int32_t K, int32_t wtype, int32_t epilogue, int32_t seed,
                                  int32_t runs, uint16_t * out_ref, uint16_t * out_tc, float * ms_tc,
                                  uint16_t * out_ref2, uint16_t * out_tc2, uint16_t * a_out, float * w_out);

/* Model quantiser: fname_inp (f32 / f16 file) -> fname_out with every 2-D "*weight" tensor as
 * q4_0 (itype 2) or q4_1 (itype 3) blocks, byte-identical to the reference's `quantize` tool
 * (models/quantize.cpp:27-268, `./quantize in.bin out.bin type`; block arithmetic ggml.c:942-1018).  The block
 * quantisation runs as a CUDA kernel; returns 0 on success, non-zero (+ message on stderr) otherwise. */
BERT_API int32_t bertx_quantize_file(const char * fname_inp, const char * fname_out, int32_t itype);

/* Checkpoint conversion on the box: a Hugging Face BERT checkpoint (model.safetensors, tensors F32 / F16 /
 * BF16) + config.json + vocab.txt -> fname_out in the reference's model format, byte-for-byte what
 * `python models/convert-to-ggml.py dir ftype` writes (models/convert-to-ggml.py:68-110; ftype 0 f32, 1 f16) and, for
 * ftype 2 (q4_0) / 3 (q4_1), what `quantize ggml-model-f16.bin out.bin ftype` then makes of the f16 file
 * (models/run_conversions.sh, models/quantize.cpp:27-268).  ftype 0 / 1 run on the host alone; the q4 block kernel is
 * CUDA.  Returns 0 on success, non-zero (+ message on stderr) otherwise. */
BERT_API int32_t bertx_convert_safetensors(const char * safetensors, const char * config_json, const char * vocab_txt,
                                           int32_t ftype, const char * fname_out);

/* The same conversion straight into an engine: the file image is built in an anonymous memory file and handed to
 * bert_load_from_file, so nothing touches the disk and the q4 blocks on the device are the ones the reference's tools
 * would have written.  Returns NULL on failure (message on stderr).  Free with bert_free. */
BERT_API struct bert_ctx * bertx_load_from_safetensors(const char * safetensors, const char * config_json,
                                                       const char * vocab_txt, int32_t ftype);

/* No counterpart in the reference; added for callers that cannot allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The block quantiser of
 * bertx_quantize_file on caller-owned HOST buffers: src holds n_elems values (src_ftype 0 = f32, 1 = f16; a whole
 * number of 32-value blocks), dst receives n_elems / 32 blocks of 18 (itype 2, q4_0) or 20 (itype 3, q4_1) bytes, the
 * same bytes bertx_quantize_file writes for those values.  Calls no cudaMalloc / cudaFree: the device staging lives in a
 * fixed __device__ array of 4.625 MiB that every GPU using the entry point keeps until exit.  Limit: n_elems <=
 * 1 048 576 per call (larger inputs: split on block boundaries; return code 9 and a message otherwise).  Runs on the
 * current device, returns after the result is in dst.  All calls share the array (calls of one process are serialised
 * inside; work that overlapped on the GPU would race on it).  Returns 0 on success. */
BERT_API int32_t bertx_quantize_rows_noalloc(const void * src, int32_t src_ftype, int64_t n_elems, int32_t itype, void * dst);

#ifdef __cplusplus
}
#endif
#endif /* BERT_B200_H */
