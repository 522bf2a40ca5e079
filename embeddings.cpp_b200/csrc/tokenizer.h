// Host WordPiece tokenizer, bit-exact with the reference's (bert.cpp:199-417) on every
// input the reference handles without overrunning its buffer.  Same byte-level algorithm
// (accent strip -> ASCII lower-case -> isolate ASCII punctuation and 3-byte CJK ->
// split on isspace -> greedy longest-prefix match with the two vocab maps), implemented
// as two scans (fold + lower-case; classify + split + match) over flat hash tables with a
// bounded match length instead of std::map + O(n^2) substr and three rewritten strings.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <string_view>
#include <vector>

namespace b200 {

// Open-addressing string -> id table for the vocabulary maps: one 32-byte slot holds the hash, the id and the key itself
// (keys longer than 20 bytes live in a side arena), so a hit is one cache line instead of std::unordered_map's bucket ->
// node -> key bytes chain.  Insert-only; `put` overwrites an existing key's id (the caller implements first-wins by `find`).
class VocabTable {
public:
    void reserve(size_t n_keys);
    void put(std::string_view key, int32_t id);
    // id or -1
    int32_t find(std::string_view key) const {
        if (slots_.empty()) return -1;
        const uint64_t h = hash(key);
        const uint32_t tag = (uint32_t) (h >> 32) | 1u;              // 0 marks an empty slot
        for (size_t i = (size_t) h & mask_;; i = (i + 1) & mask_) {
            const Slot & s = slots_[i];
            if (s.tag == 0) return -1;
            if (s.tag == tag && s.len == key.size() && memcmp(key_of(s), key.data(), key.size()) == 0) return s.id;
        }
    }

private:
    struct Slot {
        uint32_t tag = 0;
        int32_t id = 0;
        uint32_t len = 0;
        char key[20] = {};                                           // the key, or (len > 20) its offset in arena_
    };
    static_assert(sizeof(Slot) == 32, "two slots per cache line");
    const char * key_of(const Slot & s) const {
        if (s.len <= sizeof(s.key)) return s.key;
        uint32_t off; memcpy(&off, s.key, 4);
        return arena_.data() + off;
    }
    static uint64_t hash(std::string_view k) {
        // 8 bytes at a time through a multiply-xorshift mix (the keys are short words; any avalanche-ing mix will do)
        uint64_t h = 0x9E3779B97F4A7C15ull ^ (k.size() * 0xD6E8FEB86659FD93ull);
        const char * p = k.data();
        size_t n = k.size();
        while (n >= 8) { uint64_t w; memcpy(&w, p, 8); h = (h ^ w) * 0xFF51AFD7ED558CCDull; h ^= h >> 32; p += 8; n -= 8; }
        if (n) { uint64_t w = 0; memcpy(&w, p, n); h = (h ^ w) * 0xC4CEB9FE1A85EC53ull; h ^= h >> 29; }
        h *= 0x9FB21C651E98DF25ull;
        return h ^ (h >> 32);
    }
    std::vector<Slot> slots_;
    std::string arena_;
    size_t mask_ = 0;
};

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
    VocabTable token_to_id_;                                        // first occurrence wins
    VocabTable subword_to_id_;                                      // "##x" -> "x", last wins
    std::vector<int8_t> id_in_token_map_;                           // _id_to_token membership
    size_t max_tok_len_ = 0, max_sub_len_ = 0;
};

}  // namespace b200
