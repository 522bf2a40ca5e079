#include "model_file.h"

#include <cstdio>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace b200 {

size_t tensor_row_bytes(int ftype, int64_t ne0) {
    switch (ftype) {
        case 0: return (size_t) ne0 * 4;
        case 1: return (size_t) ne0 * 2;
        case 2: return (size_t) (ne0 / 32) * 18;     // block_q4_0, ggml.c:893-899
        case 3: return (size_t) (ne0 / 32) * 20;     // block_q4_1, ggml.c:901-907
    }
    return 0;
}

ModelFile::~ModelFile() {
    if (map_) munmap(map_, map_len_);
}

namespace {
struct Cursor {
    const uint8_t * p; const uint8_t * end;
    bool read(void * dst, size_t n) {
        if ((size_t) (end - p) < n) return false;
        memcpy(dst, p, n); p += n; return true;
    }
};
}  // namespace

bool ModelFile::load(const char * fname) {
    const char * fn = "bert_load_from_file";
    int fd = open(fname, O_RDONLY);
    if (fd < 0) { fprintf(stderr, "%s: failed to open '%s'\n", fn, fname); return false; }
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 32) {
        fprintf(stderr, "%s: invalid model file '%s' (bad magic)\n", fn, fname);
        close(fd); return false;
    }
    map_len_ = (size_t) st.st_size;
    map_ = mmap(nullptr, map_len_, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map_ == MAP_FAILED) { map_ = nullptr; fprintf(stderr, "%s: failed to open '%s'\n", fn, fname); return false; }
    Cursor c{(const uint8_t *) map_, (const uint8_t *) map_ + map_len_};

    uint32_t magic = 0;
    c.read(&magic, 4);
    if (magic != 0x67676d6c) { fprintf(stderr, "%s: invalid model file '%s' (bad magic)\n", fn, fname); return false; }
    int32_t h[7];
    if (!c.read(h, sizeof(h))) { fprintf(stderr, "%s: invalid model file '%s' (truncated header)\n", fn, fname); return false; }
    hp = {h[0], h[1], h[2], h[3], h[4], h[5], h[6]};
    printf("%s: n_vocab = %d\n", fn, hp.n_vocab);
    printf("%s: n_max_tokens   = %d\n", fn, hp.n_max_tokens);
    printf("%s: n_embd  = %d\n", fn, hp.n_embd);
    printf("%s: n_intermediate  = %d\n", fn, hp.n_intermediate);
    printf("%s: n_head  = %d\n", fn, hp.n_head);
    printf("%s: n_layer = %d\n", fn, hp.n_layer);
    printf("%s: f16     = %d\n", fn, hp.ftype);
    if (hp.ftype < 0 || hp.ftype > 3) {
        fprintf(stderr, "%s: invalid model file '%s' (bad f16 value %d)\n", fn, fname, hp.ftype);
        return false;
    }
    if (hp.n_vocab <= 0 || hp.n_vocab > (1 << 24) || hp.n_layer < 0 || hp.n_embd <= 0 || hp.n_head <= 0 ||
        hp.n_max_tokens <= 0 || hp.n_intermediate <= 0) {
        fprintf(stderr, "%s: invalid model file '%s' (bad hparams)\n", fn, fname);
        return false;
    }
    vocab.resize(hp.n_vocab);
    for (int i = 0; i < hp.n_vocab; ++i) {
        uint32_t len = 0;
        if (!c.read(&len, 4) || (size_t) (c.end - c.p) < len) {
            fprintf(stderr, "%s: invalid model file '%s' (truncated vocab)\n", fn, fname);
            return false;
        }
        vocab[i].assign((const char *) c.p, len);
        c.p += len;
    }
    while (c.p < c.end) {
        int32_t n_dims, length, ftype;
        if (!c.read(&n_dims, 4) || !c.read(&length, 4) || !c.read(&ftype, 4)) break;   // EOF (bert.cpp:666-669)
        if (n_dims < 1 || n_dims > 2 || length < 0 || length > 4096) {
            fprintf(stderr, "%s: invalid model file '%s' (bad tensor header)\n", fn, fname);
            return false;
        }
        TensorEntry t;
        t.n_dims = n_dims; t.ftype = ftype;
        int64_t nel = 1;
        for (int i = 0; i < n_dims; ++i) {
            int32_t ne = 0;
            if (!c.read(&ne, 4)) { fprintf(stderr, "%s: truncated tensor header\n", fn); return false; }
            t.ne[i] = ne; nel *= ne;
        }
        if ((size_t) (c.end - c.p) < (size_t) length) { fprintf(stderr, "%s: truncated tensor name\n", fn); return false; }
        std::string name((const char *) c.p, (size_t) length);
        c.p += length;
        if (ftype < 0 || ftype > 3) {
            fprintf(stderr, "%s: unknown ftype %d in model file\n", fn, ftype);
            return false;
        }
        if ((ftype == 2 || ftype == 3) && t.ne[0] % 64 != 0) {          // bert.cpp:730,734
            fprintf(stderr, "%s: tensor '%s' has a row length not divisible by 64\n", fn, name.c_str());
            return false;
        }
        t.nbytes = tensor_row_bytes(ftype, t.ne[0]) * (size_t) t.ne[1];
        if ((size_t) (c.end - c.p) < t.nbytes) {
            fprintf(stderr, "%s: tensor '%s' has wrong size in model file\n", fn, name.c_str());
            return false;
        }
        t.data = c.p;
        c.p += t.nbytes;
        total_tensor_bytes += t.nbytes;
        ++n_tensors;
        tensors[name] = t;
    }
    return true;
}

}  // namespace b200
