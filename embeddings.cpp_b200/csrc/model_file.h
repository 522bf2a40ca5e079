// Parser for the reference's legacy-ggml model file ("ggml" magic 0x67676d6c), the only
// weight format the reference reads (bert.cpp:434-468 header, 471-495 vocab, 656-752
// tensors; written by models/convert-to-ggml.py:68-110 and models/quantize.cpp:27-268).
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace b200 {

struct HParams {
    int32_t n_vocab = 0, n_max_tokens = 0, n_embd = 0, n_intermediate = 0, n_head = 0, n_layer = 0, ftype = 0;
};

struct TensorEntry {
    int32_t n_dims = 0;
    int32_t ftype = 0;             // 0 f32, 1 f16, 2 q4_0, 3 q4_1
    int64_t ne[2] = {1, 1};        // ne[0] = fastest (row length)
    const uint8_t * data = nullptr;
    size_t nbytes = 0;
};

struct ModelFile {
    HParams hp;
    std::vector<std::string> vocab;
    std::map<std::string, TensorEntry> tensors;
    size_t total_tensor_bytes = 0;
    int n_tensors = 0;

    ~ModelFile();
    // Returns false (after printing the reference's style of message to stderr) on failure.
    bool load(const char * fname);

private:
    void * map_ = nullptr;
    size_t map_len_ = 0;
};

size_t tensor_row_bytes(int ftype, int64_t ne0);

}  // namespace b200
