// The encoder forward as a fixed sequence of sm_100a kernels on one stream per GPU (the host
// runtime that feeds it -- tokenizer pool, micro-batch queue, per-GPU workers -- is pipeline.cu).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <numeric>
#include <thread>

#include "engine.h"

namespace b200 {

namespace {

struct Launcher {
    Engine & e; Replica & r; bool ok = true;
    void operator()(int kind, cudaError_t err, const char * what) {
        e.launches.fetch_add(1, std::memory_order_relaxed);
        (void) kind;
        if (err != cudaSuccess) { ok = false; cuda_ok(err, what); }
    }
};

}  // namespace

// One micro-batch.  Kernel sequence per layer (reference graph nodes in brackets):
//   GEMM qkv  [bert.cpp:993-1016]  ->  attention [957-961, 1018-1036]  ->  GEMM o + bias +
//   residual [1040-1045]  ->  LN [1047-1056]  ->  GEMM ffn-in + bias + GELU [1059-1063]  ->
//   GEMM ffn-out + bias + residual [1066-1071]  ->  LN [1073-1082]
bool Engine::forward(Replica & r, const int32_t * d_ids, const int32_t * d_cu, int n_sent, int T, int max_len,
                     float * d_out, int n_spans) {
    const int H = hp.n_embd;
    const bool fq = fq_type >= 0;
    cudaStream_t st = r.stream;
    Launcher L{*this, r};

    auto begin = [&](int kind) {
        if (!r.profiling) return;
        Replica::Span s; s.kind = kind;
        cudaEventCreate(&s.a); cudaEventCreate(&s.b);
        cudaEventRecord(s.a, st);
        r.spans.push_back(s);
    };
    auto end = [&]() { if (r.profiling) cudaEventRecord(r.spans.back().b, st); };

    // latency path: with few tokens the GEMMs are bound by how fast ONE CTA pair can stream and dequantise its
    // weight rows (256-wide N-tiles leave 3-12 pairs busy); 64-wide tiles spread the weights over 4x as many
    // pairs.  The LayerNorm fusion needs the wide tiles, so small batches take the stand-alone LN kernel.
    const bool small = T <= small_t && (hp.ftype == 2 || hp.ftype == 3);      // the N-tile override exists for q4 weights only
    auto gemm = [&](const CUtensorMap * tm, const CUtensorMap * tmo, const __half * a, const DevWeight & w,
                    const __half * res, const CUtensorMap * tmr, __half * out, int epi) {
        GemmArgs g{tm, tmo, tmr, a, &w, res, out, T, epi, fq_type};
        g.bn_override = small ? 64 : 0;
        begin(KK_GEMM);
        L(KK_GEMM, gemm_simt ? launch_gemm_simt(g, st) : launch_gemm_tc(g, r.num_sms, st), "gemm");
        end();
    };

    // o / ffn-out GEMM with the LayerNorm that follows it fused into the epilogue (no y round trip)
    auto gemm_ln = [&](const CUtensorMap * tm, const __half * a, const DevWeight & w, const __half * res,
                       const CUtensorMap * tmr, const float * g_, const float * b_, __half * out,
                       const CUtensorMap * tmo, const CUtensorMap * tmo2) {
        GemmArgs g{tm, tmo, tmr, a, &w, res, out, T, EPI_BIAS | EPI_RESIDUAL | EPI_LN | (fq ? EPI_FAKEQUANT : 0), fq_type};
        g.tmap_out2 = fq ? tmo2 : nullptr; g.ln_gamma = g_; g.ln_beta = b_;
        begin(KK_GEMM);
        L(KK_GEMM, launch_gemm_tc(g, r.num_sms, st), "gemm+ln");
        end();
    };
    begin(KK_EMBED);
    // packed spans: several short sentences share one attention work item; the embedding kernel
    // leaves every token's sentence bounds behind for the masks.  BERT_ATTN_PACK=0 turns it off (A/B).
    const int dh_ = H / hp.n_head;
    const bool packed = n_spans > 0 && n_spans < n_sent && attn_pack && !attn_ref && !(dh_ == 64 && attn_v3);
    L(KK_EMBED, launch_embed_ln(r.word, r.postype, r.ln_e_g, r.ln_e_b, d_ids, d_cu, n_sent, T, H,
                                r.x, fq ? r.xq : nullptr, fq_type, packed ? r.tok_bounds : nullptr, st), "embed_ln");
    end();
    for (int il = 0; il < hp.n_layer && L.ok; ++il) {
        const LayerDev & ly = r.layers[il];
        gemm(&r.tm_xin, &r.to_qkv, r.xq, ly.qkv, nullptr, nullptr, r.qkv, EPI_BIAS);
        begin(KK_ATTN);
        {
            const int dh = H / hp.n_head;
            cudaError_t ae;
            // head size 64 -> attention_tc.cu, head size 32 -> attention_tc2.cu (two heads per work item); BERT_ATTN_IMPL = tc2
            // (eight softmax warps, either head size) | tc3 (the round-2 study kernel for head size 64, 12 % slower:
            // profiles/r2_attention.md) | ref (SIMT validation kernel)
            if (attn_ref) ae = launch_attention_ref(r.qkv, d_cu, n_sent, max_len, hp.n_head, dh, r.ctx, fq_type, st);
            else if (dh == 64 && attn_v3) ae = launch_attention_tc3(&r.ta_q, &r.ta_kv, d_cu, n_sent, max_len, hp.n_head, dh, r.ctx, fq_type, r.num_sms, st);
            else if (dh == 32 || attn_v2) {
                if (packed) ae = launch_attention_tc2(&r.ta_q, &r.ta_kv, d_cu + n_sent + 1, n_spans, max_len, hp.n_head, dh, r.ctx, fq_type, r.num_sms, st, r.tok_bounds);
                else ae = launch_attention_tc2(&r.ta_q, &r.ta_kv, d_cu, n_sent, max_len, hp.n_head, dh, r.ctx, fq_type, r.num_sms, st);
            }
            else if (packed) ae = launch_attention_tc(&r.ta_q, &r.ta_kv, d_cu + n_sent + 1, n_spans, max_len, hp.n_head, dh, r.ctx, fq_type, r.num_sms, st, r.tok_bounds);
            else ae = launch_attention_tc(&r.ta_q, &r.ta_kv, d_cu, n_sent, max_len, hp.n_head, dh, r.ctx, fq_type, r.num_sms, st);
            L(KK_ATTN, ae, "attention");
        }
        end();
        const bool fuse = fuse_ln && !small && !gemm_simt && gemm_ln_fusable(ly.o) && gemm_ln_fusable(ly.ffo);
        if (fuse && fuse_ln_att) {
            gemm_ln(&r.tm_ctx, r.ctx, ly.o, r.x, &r.tr_x, ly.ln_att_g, ly.ln_att_b, r.x1, &r.to_x1, &r.to_x1q);
        } else {
            gemm(&r.tm_ctx, &r.to_y, r.ctx, ly.o, r.x, &r.tr_x, r.y, EPI_BIAS | EPI_RESIDUAL);
            begin(KK_LN);
            L(KK_LN, launch_ln(r.y, ly.ln_att_g, ly.ln_att_b, T, H, r.x1, fq ? r.x1q : nullptr, fq_type, st), "ln_att");
            end();
        }
        gemm(&r.tm_x1in, &r.to_mid, r.x1q, ly.ffi, nullptr, nullptr, r.mid, EPI_BIAS | EPI_GELU | (gelu_h2 ? EPI_GELU_H2 : 0) | (fq ? EPI_FAKEQUANT : 0));
        if (fuse) {
            gemm_ln(&r.tm_mid, r.mid, ly.ffo, r.x1, &r.tr_x1, ly.ln_out_g, ly.ln_out_b, r.x, &r.to_x, &r.to_xq);
        } else {
            gemm(&r.tm_mid, &r.to_y, r.mid, ly.ffo, r.x1, &r.tr_x1, r.y, EPI_BIAS | EPI_RESIDUAL);
            begin(KK_LN);
            L(KK_LN, launch_ln(r.y, ly.ln_out_g, ly.ln_out_b, T, H, r.x, fq ? r.xq : nullptr, fq_type, st), "ln_out");
            end();
        }
    }
    begin(KK_POOL);
    L(KK_POOL, launch_pool_norm(r.x, d_cu, n_sent, H, d_out, pooling, st), "pool_norm");
    end();
    return L.ok;
}

void Engine::plan(const int32_t * n_tokens, int n, std::vector<int> & order, std::vector<MicroBatch> & mbs) const {
    order.resize(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return n_tokens[a] > n_tokens[b]; });
    mbs.clear();
    MicroBatch cur;
    size_t off = 0;
    for (int i = 0; i < n; ++i) {
        const int len = n_tokens[order[i]];
        if (cur.count > 0 && (cur.tokens + len > t_cap || cur.count + 1 > b_cap)) {
            mbs.push_back(cur);
            cur = MicroBatch();
            cur.first = i; cur.tok_offset = off;
        }
        if (cur.count == 0) { cur.first = i; cur.tok_offset = off; }
        cur.count++; cur.tokens += len; cur.max_len = std::max(cur.max_len, len);
        off += len;
    }
    if (cur.count > 0) mbs.push_back(cur);
}

void Engine::free_job(Replica & r) {
    if (r.job_ids) cudaFree(r.job_ids);
    if (r.job_cu) cudaFree(r.job_cu);
    if (r.job_out) cudaFree(r.job_out);
    r.job_ids = nullptr; r.job_cu = nullptr; r.job_out = nullptr;
    r.job_mbs.clear(); r.job_order.clear(); r.job_cu_off.clear(); r.job_sent = 0;
}

}  // namespace b200
