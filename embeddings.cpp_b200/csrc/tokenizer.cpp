#include "tokenizer.h"

#include <cstring>

namespace b200 {

namespace {

constexpr int32_t kCls = 101, kSep = 102, kUnk = 100;

// bert.cpp:199-204 -- length from the lead byte's high nibble (continuation bytes count as 1)
inline size_t utf8_len(unsigned char c) {
    static const uint8_t lut[16] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 3, 4};
    return lut[c >> 4];
}

// bert.cpp:209-219 -- the 52 Latin-1 letters (all are 0xC3 0x80..0xBF in UTF-8); 0 = not mapped
inline char accent_fold(unsigned char b0, unsigned char b1) {
    if (b0 != 0xC3) return 0;
    static const char * map =
        /* 0x80 */ "AAAAAA\0CEEEEIIII"
        /* 0x90 */ "\0NOOOOO\0\0UUUUY\0\0"
        /* 0xA0 */ "aaaaaa\0ceeeeiiii"
        /* 0xB0 */ "\0nooooo\0\0uuuuy\0\0";
    if (b1 < 0x80 || b1 > 0xBF) return 0;
    return map[b1 - 0x80];
}

// C-locale ispunct / isspace on a (possibly negative) char: bytes >= 0x80 are neither
inline bool is_punct(unsigned char c) {
    return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
}
inline bool is_space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }

// bert.cpp:253-295 on the (up to) 3 bytes at p; `avail` bytes are readable
inline bool is_chinese3(const unsigned char * p, size_t avail) {
    // caller guarantees utf8_len(p[0]) == 3, i.e. p[0] in 0xE0..0xEF
    uint32_t cp = p[0] & 0x0F;
    for (int j = 1; j < 3; ++j) {
        if ((size_t) j >= avail) return false;
        if ((p[j] >> 6) != 0x02) return false;
        cp = (cp << 6) | (p[j] & 0x3F);
    }
    return (cp >= 0x4E00 && cp <= 0x9FFF) || (cp >= 0x3400 && cp <= 0x4DBF) || (cp >= 0xF900 && cp <= 0xFAFF) ||
           (cp >= 0x3000 && cp <= 0x303F) || (cp >= 0xFF00 && cp <= 0xFFEF);
    // the supplementary-plane ranges of bert.cpp:284-289 cannot be reached by a 3-byte sequence
}

}  // namespace

void VocabTable::reserve(size_t n_keys) {
    size_t cap = 16;
    while (cap < 2 * n_keys + 2) cap <<= 1;                          // load factor <= 0.5
    slots_.assign(cap, Slot());
    arena_.clear();
    mask_ = cap - 1;
}

void VocabTable::put(std::string_view key, int32_t id) {
    const uint64_t h = hash(key);
    const uint32_t tag = (uint32_t) (h >> 32) | 1u;
    for (size_t i = (size_t) h & mask_;; i = (i + 1) & mask_) {
        Slot & s = slots_[i];
        if (s.tag == 0) {
            s.tag = tag; s.id = id; s.len = (uint32_t) key.size();
            if (key.size() <= sizeof(s.key)) {
                memcpy(s.key, key.data(), key.size());
            } else {
                const uint32_t off = (uint32_t) arena_.size();
                arena_.append(key.data(), key.size());
                memcpy(s.key, &off, 4);
            }
            return;
        }
        if (s.tag == tag && s.len == key.size() && memcmp(key_of(s), key.data(), key.size()) == 0) { s.id = id; return; }
    }
}

void Tokenizer::build(std::vector<std::string> && vocab) {
    vocab_ = std::move(vocab);
    token_to_id_.reserve(vocab_.size());
    subword_to_id_.reserve(vocab_.size());
    id_in_token_map_.assign(vocab_.size(), 0);
    max_tok_len_ = max_sub_len_ = 0;
    for (size_t i = 0; i < vocab_.size(); ++i) {
        const std::string & w = vocab_[i];
        if (w.size() >= 2 && w[0] == '#' && w[1] == '#') {
            std::string_view sub(w.data() + 2, w.size() - 2);
            subword_to_id_.put(sub, (int32_t) i);                    // operator[] = : last wins
            if (sub.size() > max_sub_len_) max_sub_len_ = sub.size();
        }
        std::string_view key(w);
        if (token_to_id_.find(key) < 0) {                            // count()==0 : first wins
            token_to_id_.put(key, (int32_t) i);
            id_in_token_map_[i] = 1;
            if (key.size() > max_tok_len_) max_tok_len_ = key.size();
        }
    }
}

const char * Tokenizer::id_to_token(int32_t id) const {
    if (id >= 0 && (size_t) id < vocab_.size()) {
        const std::string & w = vocab_[id];
        if (id_in_token_map_[id]) return w.c_str();
        if (w.size() >= 2 && w[0] == '#' && w[1] == '#') return w.c_str();
    }
    return "[UNK TOKEN from bert_vocab]";
}

int32_t Tokenizer::tokenize(const char * text, int32_t * out, int32_t n_max) const {
    if (n_max <= 0) return 0;
    const size_t len = strlen(text);
    const unsigned char * src = reinterpret_cast<const unsigned char *>(text);

    // The two work strings keep their capacity from call to call (no allocation per sentence) and are written through raw
    // pointers into space sized for the worst case (no per-byte capacity checks).
    static thread_local std::string s;
    bool ascii = true;
    for (size_t i = 0; i < len; ++i) ascii &= src[i] < 0x80;
    s.resize(len + 8);
    char * sd = &s[0];
    size_t sn = 0;
    if (ascii) {
        // 1 + 2 for pure ASCII: nothing to fold, every byte is a character start
        for (size_t i = 0; i < len; ++i) { const char c = text[i]; sd[i] = (c >= 'A' && c <= 'Z') ? (char) (c - 'A' + 'a') : c; }
        sn = len;
    } else {
        // 1 + 2 in one walk.  stripAccents (bert.cpp:221-235) copies whole characters by utf8_len and folds the 52 mapped
        // letters to one ASCII byte; the lower-casing walk (bert.cpp:244-249) then visits exactly the same character starts
        // (a copied character keeps its lead byte, a folded one is a 1-byte character), so both can be applied per source
        // character: an upper-case ASCII byte can only be lower-cased where it IS a character (n == 1) or a folded letter.
        for (size_t i = 0; i < len;) {
            const unsigned char c = src[i];
            if (c < 0x80) {                                           // 1-byte character
                sd[sn++] = (c >= 'A' && c <= 'Z') ? (char) (c - 'A' + 'a') : (char) c;
                ++i;
                continue;
            }
            const size_t n = utf8_len(c);
            if (n == 2 && i + 1 < len) {
                const char a = accent_fold(c, src[i + 1]);
                if (a) { sd[sn++] = (a >= 'A' && a <= 'Z') ? (char) (a - 'A' + 'a') : a; i += 2; continue; }
            }
            if (i + 4 <= len) {                                       // whole character readable: fixed-size copy
                memcpy(sd + sn, text + i, 4);                         // (s has 8 spare bytes; only n of the 4 count)
                sn += n;
            } else {
                const size_t take = i + n <= len ? n : len - i;       // substr clips at the end
                for (size_t k = 0; k < take; ++k) sd[sn + k] = text[i + k];
                sn += take;
            }
            i += n;
        }
    }
    // 3 + 4 + 5 in one scan.  The reference first rewrites the string with a space on either side of every ASCII punctuation
    // byte and every 3-byte CJK character (byte-wise walk, bert.cpp:317-339), then splits on isspace (341-358) and matches
    // each word greedily (360-416).  The same words come out of one scan that treats a punctuation byte and a qualifying
    // CJK character as words of their own: the isolation test of the reference looks at one byte position at a time and
    // keeps no state, so it can be applied where the word scan stands.
    static const struct ByteClass {
        uint8_t c[256];
        ByteClass() {
            for (int v = 0; v < 256; ++v) {
                c[v] = is_space((unsigned char) v) ? 1 : is_punct((unsigned char) v) ? 2 : (v & 0xF0) == 0xE0 ? 3 : 0;
            }
        }
    } cls;
    int32_t nt = 0;
    out[nt++] = kCls;
    if (n_max == 1) return nt;                                        // no room even for [SEP]
    const unsigned char * p = reinterpret_cast<const unsigned char *>(sd);
    // greedy longest match of one word (bert.cpp:360-416)
    auto match_word = [&](const char * w, size_t n) {
        const int32_t prev = nt;
        size_t i = 0;
        const VocabTable * map = &token_to_id_;
        size_t maxlen = max_tok_len_;
        while (i < n) {
            if (nt >= n_max - 1) break;
            size_t j = n - i > maxlen ? i + maxlen : n;               // longer pieces cannot match
            bool found = false;
            for (; j > i; --j) {
                const int32_t id = map->find(std::string_view(w + i, j - i));
                if (id >= 0) {
                    out[nt++] = id;
                    i = j;
                    found = true;
                    break;
                }
            }
            if (!found) ++i;                                          // skip ONE byte (bert.cpp:401-406)
            map = &subword_to_id_;
            maxlen = max_sub_len_;
        }
        if (prev == nt && nt < n_max - 1) out[nt++] = kUnk;           // bounded, unlike bert.cpp:408-412
    };
    for (size_t i = 0; i < sn;) {
        const uint8_t k = cls.c[p[i]];
        if (k == 1) { ++i; continue; }                                // white space
        if (k == 2) { match_word(sd + i, 1); ++i; continue; }         // a punctuation byte is a word
        if (k == 3 && is_chinese3(p + i, sn - i)) { match_word(sd + i, 3); i += 3; continue; }   // so is a CJK character
        size_t j = i + 1;                                             // an ordinary word: up to the next space / punctuation / CJK
        for (;;) {
            while (j < sn && cls.c[p[j]] == 0) ++j;
            if (j < sn && cls.c[p[j]] == 3 && !is_chinese3(p + j, sn - j)) { ++j; continue; }   // a 0xE? byte that stays in the word
            break;
        }
        match_word(sd + i, j - i);
        i = j;
    }
    out[nt++] = kSep;
    return nt;
}

}  // namespace b200
