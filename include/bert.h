/*
 * bert.h -- the drop-in C ABI of the B200-native BERT embedding engine.
 *
 * This header declares exactly the entry points the reference library exports for the
 * hot path (snowyu/embeddings.cpp, /root/reference/bert.h:18-90), with the same names,
 * argument order, types and ownership rules, so that the reference's consumers
 * (examples/main.cpp, examples/server.cpp, examples/test_*.cpp, examples/dylib.cpp,
 * examples/sample_dylib.py, benchmarks/run_mteb.py) bind to libbert.so unchanged.
 * The implementation behind it is hand-written CUDA for sm_100a; there is no CPU
 * fallback: every entry that needs the GPU fails loudly (stderr + untouched outputs /
 * NULL context) when no B200 is present.
 *
 * Conventions kept from the reference:
 *   - load errors  -> NULL + message on stderr     (bert.cpp:430,440,516,569,686-750)
 *   - encode/forward errors -> message on stderr, outputs left untouched (bert.cpp:850,869)
 *   - load progress lines go to stdout             (bert.cpp:425,461-467,765)
 *   - a context is not re-entrant; calls on one context are serialised internally.
 *   - n_threads is accepted for ABI compatibility; it caps host tokenizer threads only.
 *   - n_batch_size is a hint (results do not depend on it beyond fp reassociation).
 */
#ifndef BERT_H
#define BERT_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#if defined(_WIN32)
#  define BERT_API __declspec(dllexport)
#else
#  define BERT_API __attribute__((visibility("default")))
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Laid out by callers (examples/main.cpp:12, server.cpp:37): field order and types are
 * ABI.  Replaces reference bert.h:18-25. */
struct bert_params {
    int32_t n_threads = 6;
    int32_t port      = 8080;   /* TCP port for examples/server.cpp */
    const char * model  = "models/all-MiniLM-L6-v2/ggml-model-q4_0.bin";
    const char * prompt = "test prompt";
};

/* argv parser shared by all example binaries: -t/--threads, -p/--prompt, --port,
 * -m/--model, -h/--help; unknown arguments print usage and exit(0).
 * Replaces reference bert.h:27 / bert.cpp:157-193. */
BERT_API bool bert_params_parse(int argc, char ** argv, bert_params & params);

struct bert_ctx;                 /* opaque */
typedef int32_t bert_vocab_id;   /* reference bert.h:31 */

/* Parse a legacy-ggml model file (f32 / f16 / q4_0 / q4_1), build the tokenizer and
 * upload one weight replica to every selected GPU.  Replaces bert.h:33 / bert.cpp:423-786. */
BERT_API struct bert_ctx * bert_load_from_file(const char * fname);
/* Replaces bert.h:34 / bert.cpp:812-815. */
BERT_API void bert_free(bert_ctx * ctx);

/* Tokenise + encode one NUL-terminated UTF-8 text into embeddings[bert_n_embd()].
 * Replaces bert.h:38-42 / bert.cpp:1365-1372. */
BERT_API void bert_encode(
    struct bert_ctx * ctx,
    int32_t n_threads,
    const char * texts,
    float * embeddings);

/* Tokenise + encode n_inputs texts; embeddings[i] is a caller-owned float[bert_n_embd()]
 * (rows need not be contiguous).  Output order == input order.
 * Replaces bert.h:46-52 / bert.cpp:1374-1444. */
BERT_API void bert_encode_batch(
    struct bert_ctx * ctx,
    int32_t n_threads,
    int32_t n_batch_size,
    int32_t n_inputs,
    const char ** texts,
    float ** embeddings);

/* WordPiece ids of `text`, bit-exact with the reference tokenizer; writes at most
 * n_max_tokens ids (the reference overruns for longer inputs; this one truncates and
 * keeps [SEP]).  Replaces bert.h:56-61 / bert.cpp:297-417. */
BERT_API void bert_tokenize(
    struct bert_ctx * ctx,
    const char * text,
    bert_vocab_id * tokens,
    int32_t * n_tokens,
    int32_t n_max_tokens);

/* Encoder forward of one id sequence.  Replaces bert.h:63-68 / bert.cpp:817-825. */
BERT_API void bert_forward(
    struct bert_ctx * ctx,
    int32_t n_threads,
    bert_vocab_id * tokens,
    int32_t n_tokens,
    float * embeddings);

/* Encoder forward of n_batch_size id sequences (any order: packing replaces padding).
 * Replaces bert.h:71-77 / bert.cpp:827-1147. */
BERT_API void bert_forward_batch(
    struct bert_ctx * ctx,
    int32_t n_threads,
    int32_t n_batch_size,
    bert_vocab_id ** batch_tokens,
    int32_t * n_tokens,
    float ** batch_embeddings);

/* Same results as bert_forward_batch (the reference's per-sentence loop is
 * mathematically identical).  Replaces bert.h:79-85 / bert.cpp:1151-1363. */
BERT_API void bert_forward_fake_batch(
    struct bert_ctx * ctx,
    int32_t n_threads,
    int32_t n_batch_size,
    bert_vocab_id ** batch_tokens,
    int32_t * n_tokens,
    float ** batch_embeddings);

/* Replaces bert.h:87-88 / bert.cpp:111-119. */
BERT_API int32_t bert_n_embd(bert_ctx * ctx);
BERT_API int32_t bert_n_max_tokens(bert_ctx * ctx);

/* Pointer valid for the context's lifetime; unknown ids give the reference's literal
 * "[UNK TOKEN from bert_vocab]".  Replaces bert.h:90 / bert.cpp:121-134. */
BERT_API const char * bert_vocab_id_to_token(bert_ctx * ctx, bert_vocab_id id);

#ifdef __cplusplus
}
#endif

#endif /* BERT_H */
