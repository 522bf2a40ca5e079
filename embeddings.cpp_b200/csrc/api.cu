// The bert.h C ABI (reference bert.h:18-90) on top of the B200 engine.  No exceptions or
// CUDA errors cross this boundary: failures print to stderr and leave outputs untouched.
#include <cstdio>
#include <cstdlib>
#include <condition_variable>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/bert_b200.h"
#include "engine.h"

using namespace b200;

namespace {

void print_usage(char ** argv, const bert_params & params) {
    fprintf(stderr, "usage: %s [options]\n", argv[0]);
    fprintf(stderr, "\n");
    fprintf(stderr, "options:\n");
    fprintf(stderr, "  -h, --help            show this help message and exit\n");
    fprintf(stderr, "  -t N, --threads N     host tokenizer threads (default: %d; the GPU does the rest)\n", params.n_threads);
    fprintf(stderr, "  -p PROMPT, --prompt PROMPT\n");
    fprintf(stderr, "                        text to embed (default: '%s')\n", params.prompt);
    fprintf(stderr, "  --port p              port to bind in server mode (default: %d)\n", params.port);
    fprintf(stderr, "  -m FNAME, --model FNAME\n");
    fprintf(stderr, "                        model path (default: %s)\n", params.model);
    fprintf(stderr, "\n");
}

// tokenise texts[0..n) into one packed id array, threaded over sentences
bool tokenize_all(Engine & e, int n_threads, int n, const char ** texts, std::vector<int32_t> & ids,
                  std::vector<int32_t> & lens, std::vector<size_t> & offs) {
    const int N = e.hp.n_max_tokens;
    lens.assign(n, 0);
    offs.assign(n + 1, 0);
    // like the reference (bert.cpp:1398) an upper bound of N ids per text -- but not zero-filled: pages nobody writes are never mapped
    std::unique_ptr<int32_t[]> tmp(new int32_t[(size_t) n * N]);
    int nt = n_threads > 0 ? n_threads : 1;
    const int hw = (int) std::thread::hardware_concurrency();
    if (hw > 0 && nt > hw) nt = hw;
    if (nt > n / 64 + 1) nt = n / 64 + 1;
    auto work = [&](int lo, int hi) {
        for (int i = lo; i < hi; ++i) lens[i] = e.tok.tokenize(texts[i], tmp.get() + (size_t) i * N, N);
    };
    if (nt <= 1) {
        work(0, n);
    } else {
        std::vector<std::thread> th;
        const int per = (n + nt - 1) / nt;
        for (int t = 0; t < nt; ++t) {
            const int lo = t * per, hi = lo + per < n ? lo + per : n;
            if (lo < hi) th.emplace_back(work, lo, hi);
        }
        for (auto & t : th) t.join();
    }
    for (int i = 0; i < n; ++i) offs[i + 1] = offs[i] + (size_t) lens[i];
    ids.resize(offs[n]);
    for (int i = 0; i < n; ++i) memcpy(ids.data() + offs[i], tmp.get() + (size_t) i * N, (size_t) lens[i] * 4);
    return true;
}


}  // namespace

extern "C" {

bool bert_params_parse(int argc, char ** argv, bert_params & params) {
    for (int i = 1; i < argc; i++) {
        const std::string arg = argv[i];
        const bool has_val = i + 1 < argc;
        if ((arg == "-t" || arg == "--threads") && has_val) {
            params.n_threads = atoi(argv[++i]);
        } else if ((arg == "-p" || arg == "--prompt") && has_val) {
            params.prompt = argv[++i];
        } else if (arg == "--port" && has_val) {
            params.port = atoi(argv[++i]);
        } else if ((arg == "-m" || arg == "--model") && has_val) {
            params.model = argv[++i];
        } else if (arg == "-h" || arg == "--help") {
            print_usage(argv, params);
            exit(0);
        } else {
            fprintf(stderr, "error: unknown argument: %s\n", arg.c_str());
            print_usage(argv, params);
            exit(0);                                   // the reference exits 0 here too (bert.cpp:184-189)
        }
    }
    return true;
}

struct bert_ctx * bert_load_from_file(const char * fname) {
    printf("%s: loading model from '%s' - please wait ...\n", __func__, fname);
    bert_ctx * ctx = nullptr;
    try {
        ModelFile mf;
        if (!mf.load(fname)) return nullptr;
        ctx = new bert_ctx;
        std::vector<std::string> vocab = mf.vocab;
        ctx->eng.tok.build(std::move(vocab));
        if (!ctx->eng.init_devices(mf)) { delete ctx; return nullptr; }
        printf("%s: model size = %8.2f MB / num tensors = %d\n", __func__, mf.total_tensor_bytes / 1024.0 / 1024.0, mf.n_tensors);
        printf("%s: %zu x sm_100a replica(s), micro-batch %d tokens, activation quant emulation %s\n", __func__,
               ctx->eng.reps.size(), ctx->eng.t_cap, ctx->eng.fq_type >= 0 ? "on" : "off");
        if (mf.hp.ftype == 0) printf("%s: f32 weights are held as fp16 on the device (fp16 operands, f32 accumulation)\n", __func__);
        fflush(stdout);
        return ctx;
    } catch (const std::exception & ex) {
        fprintf(stderr, "%s: %s\n", __func__, ex.what());
    } catch (...) {
        fprintf(stderr, "%s: unknown failure\n", __func__);
    }
    delete ctx;
    return nullptr;
}

// Host-only context (tokenizer + header, no GPU): lets the host logic be tested on a box
// without a device.  Every forward/encode call on it fails loudly.
struct bert_ctx * bertx_load_host_only(const char * fname) {
    try {
        ModelFile mf;
        if (!mf.load(fname)) return nullptr;
        bert_ctx * ctx = new bert_ctx;
        ctx->eng.hp = mf.hp;
        ctx->eng.init_limits();
        std::vector<std::string> vocab = mf.vocab;
        ctx->eng.tok.build(std::move(vocab));
        fflush(stdout);
        return ctx;
    } catch (...) {
        return nullptr;
    }
}

void bert_free(bert_ctx * ctx) { delete ctx; }

int32_t bert_n_embd(bert_ctx * ctx) { return ctx->eng.hp.n_embd; }
int32_t bert_n_max_tokens(bert_ctx * ctx) { return ctx->eng.hp.n_max_tokens; }
const char * bert_vocab_id_to_token(bert_ctx * ctx, bert_vocab_id id) { return ctx->eng.tok.id_to_token(id); }

void bert_tokenize(struct bert_ctx * ctx, const char * text, bert_vocab_id * tokens, int32_t * n_tokens,
                   int32_t n_max_tokens) {
    *n_tokens = ctx->eng.tok.tokenize(text, tokens, n_max_tokens);
}

void bert_forward_batch(bert_ctx * ctx, int32_t n_threads, int32_t n_batch_size, bert_vocab_id ** batch_tokens,
                        int32_t * n_tokens, float ** batch_embeddings) {
    (void) n_threads;
    if (!batch_embeddings || n_batch_size <= 0) return;       // the reference's memory-probe mode has no meaning here
    try {
        std::lock_guard<std::mutex> lock(ctx->eng.mu);
        if (!ctx->eng.run_tokens(batch_tokens, n_tokens, n_batch_size, batch_embeddings)) {
            fprintf(stderr, "%s: failed, outputs of unfinished inputs left untouched\n", __func__);
        }
    } catch (const std::exception & ex) {
        fprintf(stderr, "%s: %s\n", __func__, ex.what());
    } catch (...) {
        fprintf(stderr, "%s: unknown failure\n", __func__);
    }
}

void bert_forward_fake_batch(bert_ctx * ctx, int32_t n_threads, int32_t n_batch_size, bert_vocab_id ** batch_tokens,
                             int32_t * n_tokens, float ** batch_embeddings) {
    bert_forward_batch(ctx, n_threads, n_batch_size, batch_tokens, n_tokens, batch_embeddings);
}

void bert_forward(struct bert_ctx * ctx, int32_t n_threads, bert_vocab_id * tokens, int32_t n_tokens, float * embeddings) {
    bert_forward_batch(ctx, n_threads, 1, &tokens, &n_tokens, embeddings ? &embeddings : nullptr);
}

void bert_encode_batch(struct bert_ctx * ctx, int32_t n_threads, int32_t n_batch_size, int32_t n_inputs,
                       const char ** texts, float ** embeddings) {
    (void) n_batch_size;                                      // a hint only: micro-batches are cut by token budget
    if (n_inputs <= 0) return;
    try {
        std::lock_guard<std::mutex> lock(ctx->eng.mu);
        if (!ctx->eng.encode_texts(n_threads, n_inputs, texts, embeddings)) {
            fprintf(stderr, "%s: failed, outputs of unfinished inputs left untouched\n", __func__);
        }
    } catch (const std::exception & ex) {
        fprintf(stderr, "%s: %s\n", __func__, ex.what());
    } catch (...) {
        fprintf(stderr, "%s: unknown failure\n", __func__);
    }
}

void bert_encode(struct bert_ctx * ctx, int32_t n_threads, const char * texts, float * embeddings) {
    bert_encode_batch(ctx, n_threads, 1, 1, &texts, &embeddings);
}

// ------------------------------------------------------------------ extension entry points
int32_t bertx_n_devices(struct bert_ctx * ctx) { return (int32_t) ctx->eng.reps.size(); }

int32_t bertx_hparam(struct bert_ctx * ctx, int32_t index) {
    const HParams & h = ctx->eng.hp;
    const int32_t v[7] = {h.n_vocab, h.n_max_tokens, h.n_embd, h.n_intermediate, h.n_head, h.n_layer, h.ftype};
    return index >= 0 && index < 7 ? v[index] : -1;
}

int64_t bertx_kernel_launches(struct bert_ctx * ctx) { return ctx->eng.launches.load(); }

int32_t bertx_plan(struct bert_ctx * ctx, int32_t n_devices, int32_t n_sent, const int32_t * n_tokens,
                   int32_t * mb_of_sentence, int32_t * mb_tokens, int32_t cap) {
    try {
        if (!ctx || n_sent <= 0 || n_devices <= 0) return -1;
        std::vector<int> mbt;
        const int n = ctx->eng.plan_encode(n_tokens, n_sent, n_devices, mb_of_sentence, mbt);
        if (n > cap) return -2;
        for (int i = 0; i < n; ++i) mb_tokens[i] = mbt[i];
        return n;
    } catch (...) {
        return -3;
    }
}

int32_t bertx_pack_spans(int32_t n_sent, const int32_t * n_tokens, int32_t * span_cu) {
    if (n_sent < 0 || (n_sent > 0 && (!n_tokens || !span_cu))) return -1;
    std::vector<int32_t> cu((size_t) n_sent + 1, 0);
    for (int i = 0; i < n_sent; ++i) cu[i + 1] = cu[i] + n_tokens[i];
    return b200::pack_spans(cu.data(), n_sent, span_cu);
}

int32_t bertx_set_pooling(struct bert_ctx * ctx, int32_t mode) {
    if (!ctx || (mode != 0 && mode != 1)) return -1;
    std::lock_guard<std::mutex> lk(ctx->eng.mu);
    const int32_t prev = ctx->eng.pooling;
    ctx->eng.pooling = mode;
    return prev;
}

int64_t bertx_tokenize_batch(struct bert_ctx * ctx, int32_t n_threads, int32_t n_inputs, const char ** texts,
                             bert_vocab_id * tokens_out, int64_t cap, int32_t * n_tokens_out) {
    try {
        std::vector<int32_t> ids, lens;
        std::vector<size_t> offs;
        tokenize_all(ctx->eng, n_threads, n_inputs, texts, ids, lens, offs);
        if ((int64_t) ids.size() > cap) return -1;
        memcpy(tokens_out, ids.data(), ids.size() * 4);
        memcpy(n_tokens_out, lens.data(), (size_t) n_inputs * 4);
        return (int64_t) ids.size();
    } catch (...) {
        return -1;
    }
}

}  // extern "C"
