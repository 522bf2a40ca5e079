// Host WordPiece tokenizer, bit-exact with the reference's (bert.cpp:199-417) on every
// input the reference handles without overrunning its buffer.  Same byte-level algorithm
// (accent strip -> ASCII lower-case -> isolate ASCII punctuation and 3-byte CJK ->
// split on isspace -> greedy longest-prefix match with the two vocab maps), implemented
// over hash maps with a bounded match length instead of std::map + O(n^2) substr.
#pragma once
#include <cstdint>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

namespace b200 {

class Tokenizer {
public:
    // tokens in id order, exactly as stored in the model file (bert.cpp:471-495)
    void build(std::vector<std::string> && vocab);

    // Writes at most n_max ids ([CLS] ... [SEP]); returns the count.  Inputs that would
    // need more ids are truncated (the reference overruns: bert.cpp:408-415).
    int32_t tokenize(const char * text, int32_t * out, int32_t n_max) const;

    // reference bert.cpp:121-134
    const char * id_to_token(int32_t id) const;

    size_t size() const { return vocab_.size(); }

private:
    std::vector<std::string> vocab_;
    std::unordered_map<std::string_view, int32_t> token_to_id_;     // first occurrence wins
    std::unordered_map<std::string_view, int32_t> subword_to_id_;   // "##x" -> "x", last wins
    std::vector<int8_t> id_in_token_map_;                           // _id_to_token membership
    size_t max_tok_len_ = 0, max_sub_len_ = 0;
};

}  // namespace b200
