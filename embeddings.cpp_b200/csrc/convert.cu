// Hugging Face BERT checkpoint (model.safetensors + config.json + vocab.txt) -> the reference's legacy-ggml model file, or
// straight into an engine.  Replaces models/convert-to-ggml.py:27-110 (header :68-75, vocabulary :77-82, tensors :84-110)
// and, for q4_0 / q4_1, the `quantize` step models/run_conversions.sh chains behind the f16 conversion
// (models/quantize.cpp:27-268; block arithmetic ggml.c:942-1018 -- the CUDA kernel of quantize.cu).  SURVEY.md 8f-2.
//
// The output is byte-for-byte what the reference's two tools write for the same checkpoint:
//   * header: magic, vocab_size, max_position_embeddings, hidden_size, intermediate_size, num_attention_heads,
//     num_hidden_layers (config.json), ftype;
//   * vocabulary: the first vocab_size lines of vocab.txt, each WITHOUT ITS LAST CHARACTER (the script's `vocab[i][:-1]`
//     strips the newline that way, so a final line without a newline loses a letter -- kept, it is what the reference
//     ships), as int32 length + bytes;
//   * tensors in the order of transformers' BertModel.state_dict(), `pooler.*` and `embeddings.position_ids` skipped;
//     per tensor n_dims, name length, type, the dimensions innermost first, the name, the data.  ftype 1 turns 2-D
//     "*.weight" tensors into fp16 (round-to-nearest-even, numpy's astype), everything else stays f32;
//   * ftype 2 / 3: the f16 image's 2-D "*weight" tensors (the three embedding tables included) become q4_0 / q4_1 blocks.
// Checkpoint tensors may be F32, F16 or BF16 (widened exactly to f32 first; the script only works for F32) and may carry
// a "bert." prefix.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/bert_b200.h"

namespace {

const char * FN = "bertx_convert_safetensors";

// ------------------------------------------------------------------ a file, mapped read-only
struct Mapped {
    const uint8_t * p = nullptr;
    size_t n = 0;
    ~Mapped() { if (p) munmap((void *) p, n); }
    bool open(const char * path) {
        int fd = ::open(path, O_RDONLY);
        if (fd < 0) { fprintf(stderr, "%s: failed to open '%s'\n", FN, path); return false; }
        struct stat st;
        if (fstat(fd, &st) != 0 || st.st_size <= 0) { fprintf(stderr, "%s: '%s' is empty\n", FN, path); close(fd); return false; }
        n = (size_t) st.st_size;
        void * m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
        close(fd);
        if (m == MAP_FAILED) { fprintf(stderr, "%s: cannot map '%s'\n", FN, path); return false; }
        p = (const uint8_t *) m;
        return true;
    }
};

// ------------------------------------------------------------------ the little JSON both headers need
struct Json {
    enum Kind { NUL, BOOL, NUM, STR, ARR, OBJ } kind = NUL;
    double num = 0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;
    const Json * get(const char * key) const {
        for (const auto & kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
};

struct JsonParser {
    const char * p; const char * end;
    bool ok = true;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p; }
    bool lit(const char * s) { const size_t n = strlen(s); if ((size_t) (end - p) >= n && memcmp(p, s, n) == 0) { p += n; return true; } return false; }
    std::string string() {
        std::string out;
        if (p >= end || *p != '"') { ok = false; return out; }
        ++p;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                ++p;
                switch (*p) {
                    case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                    case 'u': {                                     // \uXXXX -> UTF-8 (tensor names / metadata are ASCII in practice)
                        unsigned cp = 0;
                        for (int i = 0; i < 4 && p + 1 < end; ++i) { ++p; cp = cp * 16 + (unsigned) (*p <= '9' ? *p - '0' : (*p | 32) - 'a' + 10); }
                        if (cp < 0x80) out += (char) cp;
                        else if (cp < 0x800) { out += (char) (0xC0 | (cp >> 6)); out += (char) (0x80 | (cp & 63)); }
                        else { out += (char) (0xE0 | (cp >> 12)); out += (char) (0x80 | ((cp >> 6) & 63)); out += (char) (0x80 | (cp & 63)); }
                        break;
                    }
                    default: out += *p;
                }
                ++p;
            } else out += *p++;
        }
        if (p >= end) { ok = false; return out; }
        ++p;
        return out;
    }
    Json value(int depth = 0) {
        Json j;
        ws();
        if (p >= end || depth > 32) { ok = false; return j; }
        if (*p == '{') {
            j.kind = Json::OBJ; ++p; ws();
            if (p < end && *p == '}') { ++p; return j; }
            while (ok) {
                ws();
                std::string k = string();
                ws();
                if (!ok || p >= end || *p != ':') { ok = false; break; }
                ++p;
                j.obj.emplace_back(std::move(k), value(depth + 1));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; break; }
                ok = false;
            }
        } else if (*p == '[') {
            j.kind = Json::ARR; ++p; ws();
            if (p < end && *p == ']') { ++p; return j; }
            while (ok) {
                j.arr.push_back(value(depth + 1));
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; break; }
                ok = false;
            }
        } else if (*p == '"') { j.kind = Json::STR; j.str = string(); }
        else if (lit("true")) { j.kind = Json::BOOL; j.num = 1; }
        else if (lit("false")) { j.kind = Json::BOOL; }
        else if (lit("null")) { j.kind = Json::NUL; }
        else {
            char * e = nullptr;
            std::string tmp(p, (size_t) std::min<ptrdiff_t>(end - p, 40));
            j.num = strtod(tmp.c_str(), &e);
            if (e == tmp.c_str()) { ok = false; return j; }
            p += e - tmp.c_str();
            j.kind = Json::NUM;
        }
        return j;
    }
};

bool parse_json(const uint8_t * p, size_t n, Json & out, const char * what) {
    JsonParser jp{(const char *) p, (const char *) p + n};
    out = jp.value();
    if (!jp.ok || out.kind != Json::OBJ) { fprintf(stderr, "%s: cannot parse %s\n", FN, what); return false; }
    return true;
}

// ------------------------------------------------------------------ the checkpoint
struct SrcTensor {
    int dtype = 0;                 // 0 F32, 1 F16, 2 BF16
    std::vector<int64_t> shape;
    const uint8_t * data = nullptr;
    size_t nbytes = 0;
};

bool read_safetensors(const Mapped & f, std::map<std::string, SrcTensor> & out) {
    if (f.n < 8) { fprintf(stderr, "%s: safetensors file too short\n", FN); return false; }
    uint64_t hlen = 0;
    memcpy(&hlen, f.p, 8);
    if (hlen > f.n - 8 || hlen > (100u << 20)) { fprintf(stderr, "%s: bad safetensors header length\n", FN); return false; }
    Json hdr;
    if (!parse_json(f.p + 8, (size_t) hlen, hdr, "the safetensors header")) return false;
    const uint8_t * base = f.p + 8 + hlen;
    const size_t avail = f.n - 8 - (size_t) hlen;
    for (const auto & kv : hdr.obj) {
        if (kv.first == "__metadata__") continue;
        const Json * dt = kv.second.get("dtype"), * sh = kv.second.get("shape"), * off = kv.second.get("data_offsets");
        if (!dt || !sh || !off || off->arr.size() != 2) { fprintf(stderr, "%s: malformed entry '%s'\n", FN, kv.first.c_str()); return false; }
        SrcTensor t;
        if (dt->str == "F32") t.dtype = 0; else if (dt->str == "F16") t.dtype = 1; else if (dt->str == "BF16") t.dtype = 2;
        else continue;                                             // integer buffers (position_ids) are not weights
        size_t nel = 1;
        for (const Json & d : sh->arr) { t.shape.push_back((int64_t) d.num); nel *= (size_t) d.num; }
        const size_t b0 = (size_t) off->arr[0].num, b1 = (size_t) off->arr[1].num;
        if (b1 < b0 || b1 > avail || b1 - b0 != nel * (t.dtype == 0 ? 4 : 2)) {
            fprintf(stderr, "%s: data range of '%s' does not match its shape\n", FN, kv.first.c_str());
            return false;
        }
        t.data = base + b0; t.nbytes = b1 - b0;
        std::string name = kv.first;
        if (name.compare(0, 5, "bert.") == 0) name = name.substr(5);
        out[name] = std::move(t);
    }
    return true;
}

// transformers' BertModel.state_dict() order (what convert-to-ggml.py:84 iterates), pooler and position_ids left out
std::vector<std::string> tensor_order(int n_layer) {
    std::vector<std::string> v = {"embeddings.word_embeddings.weight", "embeddings.position_embeddings.weight",
                                  "embeddings.token_type_embeddings.weight", "embeddings.LayerNorm.weight", "embeddings.LayerNorm.bias"};
    static const char * per_layer[] = {"attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense",
                                       "attention.output.LayerNorm", "intermediate.dense", "output.dense", "output.LayerNorm"};
    for (int i = 0; i < n_layer; ++i)
        for (const char * m : per_layer)
            for (const char * wb : {".weight", ".bias"}) v.push_back("encoder.layer." + std::to_string(i) + "." + m + wb);
    return v;
}

void put(std::vector<uint8_t> & img, const void * p, size_t n) { img.insert(img.end(), (const uint8_t *) p, (const uint8_t *) p + n); }
void put_i32(std::vector<uint8_t> & img, int32_t v) { put(img, &v, 4); }

bool ends_with(const std::string & s, const char * suf) {
    const size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// values of a checkpoint tensor as f32 (exact for all three source types)
void widen(const SrcTensor & t, std::vector<float> & out) {
    const size_t nel = t.nbytes / (t.dtype == 0 ? 4 : 2);
    out.resize(nel);
    if (t.dtype == 0) { memcpy(out.data(), t.data, t.nbytes); return; }
    const uint16_t * h = (const uint16_t *) t.data;
    for (size_t i = 0; i < nel; ++i) {
        if (t.dtype == 1) out[i] = __half2float(__ushort_as_half(h[i]));
        else { const uint32_t u = (uint32_t) h[i] << 16; memcpy(&out[i], &u, 4); }
    }
}

// The whole model file image in memory.  ftype 0 / 1: what convert-to-ggml.py writes; 2 / 3: what `quantize` makes of the
// f16 file.
int32_t build_image(const char * safetensors, const char * config_json, const char * vocab_txt, int32_t ftype,
                    std::vector<uint8_t> & img) {
    if (ftype < 0 || ftype > 3) { fprintf(stderr, "%s: invalid ftype %d (0 f32, 1 f16, 2 q4_0, 3 q4_1)\n", FN, ftype); return 1; }
    Mapped fc, fv, fs;
    if (!fc.open(config_json) || !fv.open(vocab_txt) || !fs.open(safetensors)) return 3;
    Json cfg;
    if (!parse_json(fc.p, fc.n, cfg, "config.json")) return 4;
    static const char * keys[6] = {"vocab_size", "max_position_embeddings", "hidden_size", "intermediate_size",
                                   "num_attention_heads", "num_hidden_layers"};
    int32_t hp[6];
    for (int i = 0; i < 6; ++i) {
        const Json * v = cfg.get(keys[i]);
        if (!v || v->kind != Json::NUM || v->num <= 0) { fprintf(stderr, "%s: config.json has no usable \"%s\"\n", FN, keys[i]); return 4; }
        hp[i] = (int32_t) v->num;
    }
    std::map<std::string, SrcTensor> src;
    if (!read_safetensors(fs, src)) return 4;

    img.clear();
    put_i32(img, 0x67676d6c);
    for (int i = 0; i < 6; ++i) put_i32(img, hp[i]);
    put_i32(img, ftype);
    // vocabulary: readlines() + [:-1]
    {
        const char * p = (const char *) fv.p, * end = p + fv.n;
        int n = 0;
        while (n < hp[0] && p < end) {
            const char * nl = (const char *) memchr(p, '\n', (size_t) (end - p));
            const char * line_end = nl ? nl + 1 : end;              // the line as readlines() returns it
            const size_t len = (size_t) (line_end - p) - 1;         // ... minus its last character
            put_i32(img, (int32_t) len);
            put(img, p, len);
            p = line_end;
            ++n;
        }
        if (n < hp[0]) { fprintf(stderr, "%s: vocab.txt has %d lines, config.json wants %d\n", FN, n, hp[0]); return 4; }
    }
    const bool want_f16 = ftype >= 1;
    const int32_t qtype = ftype >= 2 ? ftype : 0;
    std::vector<float> f32;
    std::vector<uint16_t> f16;
    std::vector<uint8_t> blocks;
    for (const std::string & name : tensor_order(hp[5])) {
        auto it = src.find(name);
        if (it == src.end()) { fprintf(stderr, "%s: checkpoint has no tensor '%s'\n", FN, name.c_str()); return 4; }
        const SrcTensor & t = it->second;
        std::vector<int64_t> shape;                                 // .squeeze()
        for (int64_t d : t.shape) if (d != 1) shape.push_back(d);
        const int n_dims = (int) shape.size();
        if (n_dims < 1 || n_dims > 2) { fprintf(stderr, "%s: tensor '%s' has %d dimensions\n", FN, name.c_str(), n_dims); return 4; }
        widen(t, f32);
        const bool to_f16 = want_f16 && n_dims == 2 && ends_with(name, ".weight");
        const bool to_q4 = qtype && n_dims == 2 && ends_with(name, "weight");
        put_i32(img, n_dims);
        put_i32(img, (int32_t) name.size());
        put_i32(img, to_q4 ? qtype : to_f16 ? 1 : 0);
        for (int i = 0; i < n_dims; ++i) put_i32(img, (int32_t) shape[n_dims - 1 - i]);
        put(img, name.data(), name.size());
        if (!to_f16) { put(img, f32.data(), f32.size() * 4); continue; }
        f16.resize(f32.size());
        for (size_t i = 0; i < f32.size(); ++i) f16[i] = __half_as_ushort(__float2half_rn(f32[i]));
        if (!to_q4) { put(img, f16.data(), f16.size() * 2); continue; }
        if (shape[n_dims - 1] % 32 != 0) { fprintf(stderr, "%s: row length of '%s' is not a multiple of 32\n", FN, name.c_str()); return 6; }
        // the block kernel of quantize.cu, in pieces of at most 2^20 values through the allocation-free entry point
        const size_t bpb = qtype == 2 ? 18 : 20;
        blocks.resize(f16.size() / 32 * bpb);
        for (size_t off = 0; off < f16.size(); off += (size_t) 1 << 20) {
            const size_t n = std::min(f16.size() - off, (size_t) 1 << 20);
            const int32_t rc = bertx_quantize_rows_noalloc(f16.data() + off, 1, (int64_t) n, qtype, blocks.data() + off / 32 * bpb);
            if (rc != 0) return rc;
        }
        put(img, blocks.data(), blocks.size());
    }
    return 0;
}

}  // namespace

extern "C" int32_t bertx_convert_safetensors(const char * safetensors, const char * config_json, const char * vocab_txt,
                                             int32_t ftype, const char * fname_out) {
    if (!safetensors || !config_json || !vocab_txt || !fname_out) return 1;
    try {
        std::vector<uint8_t> img;
        const int32_t rc = build_image(safetensors, config_json, vocab_txt, ftype, img);
        if (rc != 0) return rc;
        FILE * f = fopen(fname_out, "wb");
        if (!f) { fprintf(stderr, "%s: failed to open '%s' for writing\n", FN, fname_out); return 3; }
        const bool ok = fwrite(img.data(), 1, img.size(), f) == img.size();
        if (fclose(f) != 0 || !ok) { fprintf(stderr, "%s: write to '%s' failed\n", FN, fname_out); return 5; }
        return 0;
    } catch (const std::exception & ex) {
        fprintf(stderr, "%s: %s\n", FN, ex.what());
    } catch (...) {
        fprintf(stderr, "%s: unknown failure\n", FN);
    }
    return 10;
}

extern "C" struct bert_ctx * bertx_load_from_safetensors(const char * safetensors, const char * config_json,
                                                         const char * vocab_txt, int32_t ftype) {
    if (!safetensors || !config_json || !vocab_txt) return nullptr;
    int fd = -1;
    try {
        std::vector<uint8_t> img;
        if (build_image(safetensors, config_json, vocab_txt, ftype, img) != 0) return nullptr;
        // the image goes through the one loader there is (bert_load_from_file) by way of an anonymous memory file
        fd = memfd_create("bert_b200_model", 0);
        if (fd < 0) { fprintf(stderr, "bertx_load_from_safetensors: memfd_create failed\n"); return nullptr; }
        size_t done = 0;
        while (done < img.size()) {
            const ssize_t w = write(fd, img.data() + done, img.size() - done);
            if (w <= 0) { fprintf(stderr, "bertx_load_from_safetensors: cannot fill the memory file\n"); close(fd); return nullptr; }
            done += (size_t) w;
        }
        img.clear(); img.shrink_to_fit();
        const std::string path = "/proc/self/fd/" + std::to_string(fd);
        struct bert_ctx * ctx = bert_load_from_file(path.c_str());
        close(fd);
        return ctx;
    } catch (const std::exception & ex) {
        fprintf(stderr, "bertx_load_from_safetensors: %s\n", ex.what());
    } catch (...) {
        fprintf(stderr, "bertx_load_from_safetensors: unknown failure\n");
    }
    if (fd >= 0) close(fd);
    return nullptr;
}
