// Model upload: file tensors -> one GEMM-ready weight replica per GPU, plus workspace.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "engine.h"

namespace b200 {

bool cuda_ok(cudaError_t e, const char * what) {
    if (e == cudaSuccess) return true;
    fprintf(stderr, "libbert(b200): CUDA error at %s: %s\n", what, cudaGetErrorString(e));
    return false;
}

namespace {

template <typename T> T * dev_alloc(Replica & r, size_t n_elems) {
    void * p = nullptr;
    const size_t bytes = (n_elems * sizeof(T) + 255) & ~(size_t) 255;
    if (!cuda_ok(cudaMalloc(&p, bytes ? bytes : 256), "cudaMalloc")) return nullptr;
    r.allocs.push_back(p);
    return (T *) p;
}

template <typename T> T * pinned_alloc(Replica & r, size_t n_elems) {
    void * p = nullptr;
    if (!cuda_ok(cudaMallocHost(&p, n_elems * sizeof(T) + 64), "cudaMallocHost")) return nullptr;
    r.pinned.push_back(p);
    return (T *) p;
}

const TensorEntry * find(const ModelFile & mf, const std::string & name, int n_dims, int64_t ne0, int64_t ne1) {
    auto it = mf.tensors.find(name);
    if (it == mf.tensors.end()) {
        fprintf(stderr, "bert_load_from_file: tensor '%s' missing in model file\n", name.c_str());
        return nullptr;
    }
    const TensorEntry & t = it->second;
    if (t.n_dims != n_dims || t.ne[0] != ne0 || (n_dims == 2 && t.ne[1] != ne1)) {
        fprintf(stderr, "bert_load_from_file: tensor '%s' has wrong shape in model file: got [%lld, %lld], expected [%lld, %lld]\n",
                name.c_str(), (long long) t.ne[0], (long long) t.ne[1], (long long) ne0, (long long) ne1);
        return nullptr;
    }
    if (n_dims == 1 && t.ftype != 0) {
        fprintf(stderr, "bert_load_from_file: 1-D tensor '%s' must be f32\n", name.c_str());
        return nullptr;
    }
    return &t;
}

// raw bytes of a file tensor -> temporary device buffer
uint8_t * stage_raw(Replica & r, const TensorEntry & t) {
    void * p = nullptr;
    if (!cuda_ok(cudaMalloc(&p, t.nbytes + 256), "cudaMalloc(stage)")) return nullptr;
    if (!cuda_ok(cudaMemcpyAsync(p, t.data, t.nbytes, cudaMemcpyHostToDevice, r.stream), "H2D weights")) {
        cudaFree(p);
        return nullptr;
    }
    return (uint8_t *) p;
}

const float * upload_vec(Replica & r, const TensorEntry & t) {
    float * d = dev_alloc<float>(r, (size_t) t.ne[0]);
    if (!d) return nullptr;
    if (!cuda_ok(cudaMemcpyAsync(d, t.data, t.nbytes, cudaMemcpyHostToDevice, r.stream), "H2D vec")) return nullptr;
    return d;
}

// Build one GEMM weight [sum(N_i)][K] from 1..3 file tensors stacked along N (Q|K|V share a GEMM).
// ln_target: the GEMM is followed by residual + LayerNorm, which the epilogue can only absorb for 256- or 128-wide column blocks
// held by one cluster (gemm_ln_fusable) -- so N = 384 (all-MiniLM, bge-small-en) takes three 128-wide blocks instead of two
// 192-wide ones (unfused, the stand-alone LayerNorm kernel was 15 % of those models' step).
bool build_weight(Replica & r, DevWeight & w, const TensorEntry * const * parts, const TensorEntry * const * biases,
                  int n_parts, int K, bool ln_target = false) {
    int N = 0;
    for (int i = 0; i < n_parts; ++i) N += (int) parts[i]->ne[1];
    const int ft = parts[0]->ftype;
    for (int i = 0; i < n_parts; ++i) {
        if (parts[i]->ftype != ft || parts[i]->ne[0] != K) {
            fprintf(stderr, "bert_load_from_file: stacked weights disagree in type/shape\n");
            return false;
        }
    }
    w.N = N; w.K = K;
    w.type = ft <= 1 ? W_F16 : ft;
    w.bn = ln_target ? gemm_pick_bn_ln(N) : gemm_pick_bn(N);
    if (w.bn == 0 || K % 64 != 0) {
        fprintf(stderr, "bert_load_from_file: unsupported GEMM shape N=%d K=%d (need N %% 64 == 0, K %% 64 == 0)\n", N, K);
        return false;
    }
    float * bias = dev_alloc<float>(r, (size_t) N);
    if (!bias) return false;
    w.bias = bias;
    int n_off = 0;
    if (w.type == W_F16) {
        __half * w16 = dev_alloc<__half>(r, (size_t) N * K);
        if (!w16) return false;
        w.w16 = w16;
        for (int i = 0; i < n_parts; ++i) {
            const TensorEntry & t = *parts[i];
            __half * dst = w16 + (size_t) n_off * K;
            if (ft == 1) {
                if (!cuda_ok(cudaMemcpyAsync(dst, t.data, t.nbytes, cudaMemcpyHostToDevice, r.stream), "H2D f16")) return false;
            } else {
                uint8_t * raw = stage_raw(r, t);
                if (!raw) return false;
                bool ok = cuda_ok(launch_f32_to_f16((const float *) raw, dst, (size_t) t.ne[0] * t.ne[1], r.stream), "f32->f16");
                cudaStreamSynchronize(r.stream);
                cudaFree(raw);
                if (!ok) return false;
            }
            if (!cuda_ok(cudaMemcpyAsync(bias + n_off, biases[i]->data, biases[i]->nbytes, cudaMemcpyHostToDevice, r.stream), "H2D bias")) return false;
            n_off += (int) t.ne[1];
        }
        if (!make_tmap_f16_2d(&w.tmap, w16, (uint64_t) N, (uint64_t) K, (uint32_t) w.bn / 2)) {
            fprintf(stderr, "bert_load_from_file: cuTensorMapEncodeTiled failed for a weight map\n");
            return false;
        }
    } else {
        const size_t nblk = (size_t) N * (K / 32);
        uint4 * qnib = dev_alloc<uint4>(r, nblk);
        __half * qd = dev_alloc<__half>(r, nblk);
        __half * qm = w.type == W_Q4_1 ? dev_alloc<__half>(r, nblk) : nullptr;
        if (!qnib || !qd || (w.type == W_Q4_1 && !qm)) return false;
        w.qnib = qnib; w.qd = qd; w.qm = qm;
        memset(&w.tmap, 0, sizeof(w.tmap));
        for (int i = 0; i < n_parts; ++i) {
            const TensorEntry & t = *parts[i];
            uint8_t * raw = stage_raw(r, t);
            if (!raw) return false;
            bool ok = cuda_ok(launch_repack_q4(raw, w.type, (int) t.ne[1], K, n_off, N, qnib, qd, qm, r.stream), "repack_q4");
            cudaStreamSynchronize(r.stream);
            cudaFree(raw);
            if (!ok) return false;
            if (!cuda_ok(cudaMemcpyAsync(bias + n_off, biases[i]->data, biases[i]->nbytes, cudaMemcpyHostToDevice, r.stream), "H2D bias")) return false;
            n_off += (int) t.ne[1];
        }
    }
    return true;
}

}  // namespace

Engine::~Engine() {
    stop_workers();
    for (Replica & r : reps) {
        if (r.device < 0) continue;
        cudaSetDevice(r.device);
        if (r.stream) cudaStreamSynchronize(r.stream);
        free_job(r);
        for (Replica::Span & s : r.spans) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
        for (int i = 0; i < Replica::kSlots; ++i) {
            if (r.slot_done[i]) cudaEventDestroy(r.slot_done[i]);
            if (r.h2d_done[i]) cudaEventDestroy(r.h2d_done[i]);
            if (r.fwd_done[i]) cudaEventDestroy(r.fwd_done[i]);
        }
        for (void * p : r.allocs) cudaFree(p);
        for (void * p : r.pinned) cudaFreeHost(p);
        if (r.copy_in) { cudaStreamSynchronize(r.copy_in); cudaStreamDestroy(r.copy_in); }
        if (r.copy_out) { cudaStreamSynchronize(r.copy_out); cudaStreamDestroy(r.copy_out); }
        if (r.stream) cudaStreamDestroy(r.stream);
    }
}

// micro-batch limits (also valid for a host-only context: the planner is testable without a GPU)
void Engine::init_limits() {
    t_cap = 262144;                                           // (131072 until round 2; the larger micro-batch is 1 % faster at the headline shape:
                                                              // fewer partial waves in the 22-cluster fused-LayerNorm GEMMs, half as many launches)
    if (const char * e = getenv("BERT_MICROBATCH_TOKENS")) { int v = atoi(e); if (v >= hp.n_max_tokens) t_cap = v; }
    if (t_cap < hp.n_max_tokens) t_cap = hp.n_max_tokens;
    b_cap = t_cap / 2 < 16384 ? t_cap / 2 : 16384;                // every sentence has >= 2 ids
}

bool Engine::init_devices(const ModelFile & mf) {
    hp = mf.hp;
    const int H = hp.n_embd;
    if (H % 64 != 0 || H > 1024 || hp.n_intermediate % 64 != 0 || H % hp.n_head != 0 ||
        !(H / hp.n_head == 64 || H / hp.n_head == 32)) {
        fprintf(stderr, "bert_load_from_file: unsupported shape for the sm_100a kernels (n_embd=%d, n_head=%d, "
                        "n_intermediate=%d): need n_embd %% 64 == 0, n_embd <= 1024, head size 32 or 64\n",
                H, hp.n_head, hp.n_intermediate);
        return false;
    }
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        fprintf(stderr, "bert_load_from_file: no CUDA device visible -- this library has no CPU path\n");
        return false;
    }
    std::vector<int> devs;
    if (const char * e = getenv("BERT_CUDA_DEVICES")) {
        const char * p = e;
        while (*p) {
            char * end = nullptr;
            long v = strtol(p, &end, 10);
            if (end == p) break;
            if (v >= 0 && v < n_dev) devs.push_back((int) v);
            p = *end ? end + 1 : end;
        }
    }
    if (devs.empty()) for (int i = 0; i < n_dev; ++i) devs.push_back(i);

    if (const char * e = getenv("BERT_GEMM_IMPL")) gemm_simt = strcmp(e, "simt") == 0;
    if (const char * e = getenv("BERT_ATTN_PACK")) attn_pack = atoi(e) != 0;
    if (const char * e = getenv("BERT_FUSE_LN")) { fuse_ln = atoi(e) != 0; fuse_ln_att = atoi(e) >= 2; }
    if (const char * e = getenv("BERT_ATTN_IMPL")) { attn_ref = strcmp(e, "ref") == 0; attn_v2 = strcmp(e, "tc2") == 0; attn_v3 = strcmp(e, "tc3") == 0; }
    fq_type = (hp.ftype == W_Q4_0 || hp.ftype == W_Q4_1) ? hp.ftype : -1;
    if (const char * e = getenv("BERT_ACT_QUANT")) { if (atoi(e) == 0) fq_type = -1; }
    init_limits();
    if (const char * e = getenv("BERT_TRACE")) tracing = atoi(e) != 0;
    if (const char * e = getenv("BERT_POOLING")) pooling = strcmp(e, "cls") == 0 ? 1 : 0;
    if (const char * e = getenv("BERT_SMALL_T")) small_t = atoi(e);
    if (const char * e = getenv("BERT_GELU_H2")) gelu_h2 = atoi(e) != 0;

    reps.resize(devs.size());
    for (size_t i = 0; i < devs.size(); ++i) {
        Replica & r = reps[i];
        r.device = devs[i];
        if (!cuda_ok(cudaSetDevice(r.device), "cudaSetDevice")) return false;
        cudaDeviceProp prop;
        if (!cuda_ok(cudaGetDeviceProperties(&prop, r.device), "cudaGetDeviceProperties")) return false;
        if (prop.major != 10) {
            fprintf(stderr, "bert_load_from_file: device %d is sm_%d%d; this library is built for sm_100a only\n",
                    r.device, prop.major, prop.minor);
            return false;
        }
        r.num_sms = prop.multiProcessorCount;
        if (!cuda_ok(cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
        if (!cuda_ok(cudaStreamCreateWithFlags(&r.copy_in, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
        if (!cuda_ok(cudaStreamCreateWithFlags(&r.copy_out, cudaStreamNonBlocking), "cudaStreamCreate")) return false;
        if (!upload(r, mf)) return false;
        if (!alloc_workspace(r)) return false;
        if (!cuda_ok(cudaStreamSynchronize(r.stream), "upload sync")) return false;
    }
    start_workers();
    return true;
}

bool Engine::upload(Replica & r, const ModelFile & mf) {
    const int H = hp.n_embd, I = hp.n_intermediate;
    const TensorEntry * we = find(mf, "embeddings.word_embeddings.weight", 2, H, hp.n_vocab);
    const TensorEntry * pe = find(mf, "embeddings.position_embeddings.weight", 2, H, hp.n_max_tokens);
    const TensorEntry * te = find(mf, "embeddings.token_type_embeddings.weight", 2, H, 2);
    const TensorEntry * lg = find(mf, "embeddings.LayerNorm.weight", 1, H, 1);
    const TensorEntry * lb = find(mf, "embeddings.LayerNorm.bias", 1, H, 1);
    if (!we || !pe || !te || !lg || !lb) return false;

    // word table stays in its file type (4-bit tables are gathered and decoded per token)
    r.word.type = we->ftype; r.word.rows = hp.n_vocab; r.word.H = H;
    if (we->ftype <= 1) {
        uint8_t * d = dev_alloc<uint8_t>(r, we->nbytes);
        if (!d || !cuda_ok(cudaMemcpyAsync(d, we->data, we->nbytes, cudaMemcpyHostToDevice, r.stream), "H2D word")) return false;
        r.word.data = d;
    } else {
        const size_t nblk = (size_t) hp.n_vocab * (H / 32);
        uint4 * nib = dev_alloc<uint4>(r, nblk);
        __half * qd = dev_alloc<__half>(r, nblk);
        __half * qm = we->ftype == W_Q4_1 ? dev_alloc<__half>(r, nblk) : nullptr;
        uint8_t * raw = stage_raw(r, *we);
        if (!nib || !qd || !raw) return false;
        bool ok = cuda_ok(launch_split_q4_table(raw, we->ftype, hp.n_vocab, H, nib, qd, qm, r.stream), "split_q4_table");
        cudaStreamSynchronize(r.stream);
        cudaFree(raw);
        if (!ok) return false;
        r.word.data = nib; r.word.qd = qd; r.word.qm = qm;
    }
    // position + token_type[0] rows, decoded on the host exactly as ggml's get_rows would
    // (ggml.c:11614-11696) and pre-summed in f32: 512 x H floats, L2-resident.
    {
        std::vector<float> pt((size_t) hp.n_max_tokens * H);
        auto decode_row = [&](const TensorEntry & t, int row, float * dst) {
            const uint8_t * p = t.data + tensor_row_bytes(t.ftype, H) * (size_t) row;
            if (t.ftype == 0) { memcpy(dst, p, (size_t) H * 4); return; }
            if (t.ftype == 1) { for (int i = 0; i < H; ++i) dst[i] = __half2float(((const __half *) p)[i]); return; }
            const int bs = t.ftype == 2 ? 18 : 20;
            for (int b = 0; b < H / 32; ++b) {
                const uint8_t * q = p + (size_t) b * bs;
                __half dh, mh; memcpy(&dh, q, 2);
                float d = __half2float(dh), m = 0.f;
                int off = 2;
                if (t.ftype == 3) { memcpy(&mh, q + 2, 2); m = __half2float(mh); off = 4; }
                for (int j = 0; j < 16; ++j) {
                    const int lo = q[off + j] & 15, hi = q[off + j] >> 4;
                    dst[b * 32 + j] = t.ftype == 2 ? (float) (lo - 8) * d : (float) lo * d + m;
                    dst[b * 32 + j + 16] = t.ftype == 2 ? (float) (hi - 8) * d : (float) hi * d + m;
                }
            }
        };
        std::vector<float> type0(H), row(H);
        decode_row(*te, 0, type0.data());
        for (int p = 0; p < hp.n_max_tokens; ++p) {
            decode_row(*pe, p, row.data());
            for (int i = 0; i < H; ++i) pt[(size_t) p * H + i] = type0[i] + row[i];   // reference adds type first (bert.cpp:968-973)
        }
        float * d = dev_alloc<float>(r, pt.size());
        if (!d || !cuda_ok(cudaMemcpy(d, pt.data(), pt.size() * 4, cudaMemcpyHostToDevice), "H2D postype")) return false;
        r.postype = d;
    }
    r.ln_e_g = upload_vec(r, *lg);
    r.ln_e_b = upload_vec(r, *lb);
    if (!r.ln_e_g || !r.ln_e_b) return false;

    r.layers.resize(hp.n_layer);
    for (int il = 0; il < hp.n_layer; ++il) {
        const std::string p = "encoder.layer." + std::to_string(il) + ".";
        const TensorEntry * qw = find(mf, p + "attention.self.query.weight", 2, H, H);
        const TensorEntry * qb = find(mf, p + "attention.self.query.bias", 1, H, 1);
        const TensorEntry * kw = find(mf, p + "attention.self.key.weight", 2, H, H);
        const TensorEntry * kb = find(mf, p + "attention.self.key.bias", 1, H, 1);
        const TensorEntry * vw = find(mf, p + "attention.self.value.weight", 2, H, H);
        const TensorEntry * vb = find(mf, p + "attention.self.value.bias", 1, H, 1);
        const TensorEntry * ow = find(mf, p + "attention.output.dense.weight", 2, H, H);
        const TensorEntry * ob = find(mf, p + "attention.output.dense.bias", 1, H, 1);
        const TensorEntry * ag = find(mf, p + "attention.output.LayerNorm.weight", 1, H, 1);
        const TensorEntry * ab = find(mf, p + "attention.output.LayerNorm.bias", 1, H, 1);
        const TensorEntry * iw = find(mf, p + "intermediate.dense.weight", 2, H, I);
        const TensorEntry * ib = find(mf, p + "intermediate.dense.bias", 1, I, 1);
        const TensorEntry * fw = find(mf, p + "output.dense.weight", 2, I, H);
        const TensorEntry * fb = find(mf, p + "output.dense.bias", 1, H, 1);
        const TensorEntry * og = find(mf, p + "output.LayerNorm.weight", 1, H, 1);
        const TensorEntry * obb = find(mf, p + "output.LayerNorm.bias", 1, H, 1);
        if (!qw || !qb || !kw || !kb || !vw || !vb || !ow || !ob || !ag || !ab || !iw || !ib || !fw || !fb || !og || !obb) return false;
        LayerDev & L = r.layers[il];
        const TensorEntry * qkv_w[3] = {qw, kw, vw};
        const TensorEntry * qkv_b[3] = {qb, kb, vb};
        if (!build_weight(r, L.qkv, qkv_w, qkv_b, 3, H)) return false;
        if (!build_weight(r, L.o, &ow, &ob, 1, H, true)) return false;
        if (!build_weight(r, L.ffi, &iw, &ib, 1, H)) return false;
        if (!build_weight(r, L.ffo, &fw, &fb, 1, I, true)) return false;
        L.ln_att_g = upload_vec(r, *ag); L.ln_att_b = upload_vec(r, *ab);
        L.ln_out_g = upload_vec(r, *og); L.ln_out_b = upload_vec(r, *obb);
        if (!L.ln_att_g || !L.ln_att_b || !L.ln_out_g || !L.ln_out_b) return false;
    }
    return cuda_ok(cudaStreamSynchronize(r.stream), "upload");
}

bool Engine::alloc_workspace(Replica & r) {
    const int H = hp.n_embd, I = hp.n_intermediate;
    r.t_cap = t_cap; r.b_cap = b_cap;
    const size_t rows = ((size_t) t_cap + 127) / 128 * 128;        // whole 128-row GEMM tiles
    const bool fq = fq_type >= 0;
    r.x = dev_alloc<__half>(r, rows * H);
    r.xq = fq ? dev_alloc<__half>(r, rows * H) : r.x;
    r.qkv = dev_alloc<__half>(r, rows * 3 * H);
    r.ctx = dev_alloc<__half>(r, rows * H);
    r.y = dev_alloc<__half>(r, rows * H);
    r.x1 = dev_alloc<__half>(r, rows * H);
    r.x1q = fq ? dev_alloc<__half>(r, rows * H) : r.x1;
    r.mid = dev_alloc<__half>(r, rows * I);
    if (!r.x || !r.xq || !r.qkv || !r.ctx || !r.y || !r.x1 || !r.x1q || !r.mid) return false;
    // deterministic contents for the never-written tail rows the GEMM tiles may read
    cudaMemsetAsync(r.x, 0, rows * H * 2, r.stream);
    if (fq) { cudaMemsetAsync(r.xq, 0, rows * H * 2, r.stream); cudaMemsetAsync(r.x1q, 0, rows * H * 2, r.stream); }
    cudaMemsetAsync(r.qkv, 0, rows * 3 * H * 2, r.stream);       // attention may read (masked) rows past T: keep them finite
    cudaMemsetAsync(r.ctx, 0, rows * H * 2, r.stream);
    cudaMemsetAsync(r.x1, 0, rows * H * 2, r.stream);
    cudaMemsetAsync(r.mid, 0, rows * I * 2, r.stream);
    if (!make_tmap_f16_2d(&r.tm_xin, r.xq, rows, H, 128) || !make_tmap_f16_2d(&r.tm_ctx, r.ctx, rows, H, 128) ||
        !make_tmap_f16_2d(&r.tm_x1in, r.x1q, rows, H, 128) || !make_tmap_f16_2d(&r.tm_mid, r.mid, rows, I, 128) ||
        !make_tmap_f16_2d(&r.to_qkv, r.qkv, rows, 3 * H, 32) || !make_tmap_f16_2d(&r.to_y, r.y, rows, H, 32) ||
        !make_tmap_f16_2d(&r.to_mid, r.mid, rows, I, 32) ||
        !make_attention_tmaps(&r.ta_q, &r.ta_kv, r.qkv, rows, H) ||
        !make_tmap_f16_2d(&r.to_x, r.x, rows, H, 32) || !make_tmap_f16_2d(&r.to_xq, r.xq, rows, H, 32) ||
        !make_tmap_f16_2d(&r.to_x1, r.x1, rows, H, 32) || !make_tmap_f16_2d(&r.to_x1q, r.x1q, rows, H, 32) ||
        !make_tmap_f16_2d(&r.tr_x, r.x, rows, H, 32) || !make_tmap_f16_2d(&r.tr_x1, r.x1, rows, H, 32)) {
        fprintf(stderr, "bert_load_from_file: cuTensorMapEncodeTiled failed for an activation map\n");
        return false;
    }
    r.tok_bounds = dev_alloc<int2>(r, t_cap);
    if (!r.tok_bounds) return false;
    for (int s = 0; s < Replica::kSlots; ++s) {
        r.d_ids[s] = dev_alloc<int32_t>(r, t_cap);
        r.d_cu[s] = dev_alloc<int32_t>(r, 2 * (b_cap + 1));
        r.d_out[s] = dev_alloc<float>(r, (size_t) b_cap * H);
        r.h_ids[s] = pinned_alloc<int32_t>(r, t_cap);
        r.h_cu[s] = pinned_alloc<int32_t>(r, 2 * (b_cap + 1));
        r.h_out[s] = pinned_alloc<float>(r, (size_t) b_cap * H);
        if (!r.d_ids[s] || !r.d_cu[s] || !r.d_out[s] || !r.h_ids[s] || !r.h_cu[s] || !r.h_out[s]) return false;
        if (!cuda_ok(cudaEventCreateWithFlags(&r.slot_done[s], cudaEventDisableTiming | cudaEventBlockingSync), "cudaEventCreate")) return false;
        if (!cuda_ok(cudaEventCreateWithFlags(&r.h2d_done[s], cudaEventDisableTiming), "cudaEventCreate")) return false;
        if (!cuda_ok(cudaEventCreateWithFlags(&r.fwd_done[s], cudaEventDisableTiming), "cudaEventCreate")) return false;
    }
    return true;
}

}  // namespace b200
