"""TEST INFRASTRUCTURE ONLY -- pure-Python restatement of the reference tokenizer.

Follows /root/reference/bert.cpp byte for byte (it works on raw bytes, not code points):
  utf8_len ................. bert.cpp:199-204
  stripAccents ............. bert.cpp:206-238   (52 two-byte Latin-1 letters -> ASCII)
  bert_normalize_prompt .... bert.cpp:240-251   (ASCII lower-casing at char starts only)
  is_Chinese_char .......... bert.cpp:253-295
  bert_tokenize ............ bert.cpp:297-417
  vocab maps ............... bert.cpp:471-495   (token_to_id: first occurrence wins;
                                                 subword map: '##x' -> 'x', last wins)
Known reference UB that is NOT restated: inputs that need more than n_max_tokens ids
overrun the caller's buffer (bert.cpp:408-415 append [UNK]/[SEP] without a bound).
`tokenize(..., safe=True)` (the product's documented behaviour) stops appending at
n_max_tokens-1 and then appends [SEP]; with safe=False the function returns what the
reference WOULD write (possibly more than n_max_tokens ids), for differential tests.
"""

CLS, SEP, UNK = 101, 102, 100

_UTF8_LEN = (1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 3, 4)

_ACCENTS = {}
for _chars, _ascii in (("ÀÁÂÃÄÅ", "A"), ("àáâãäå", "a"), ("ÈÉÊË", "E"), ("èéêë", "e"),
                       ("ÌÍÎÏ", "I"), ("ìíîï", "i"), ("ÒÓÔÕÖ", "O"), ("òóôõö", "o"),
                       ("ÙÚÛÜ", "U"), ("ùúûü", "u"), ("Ý", "Y"), ("ý", "y"),
                       ("Ç", "C"), ("ç", "c"), ("Ñ", "N"), ("ñ", "n")):
    for _c in _chars:
        _ACCENTS[_c.encode("utf-8")] = ord(_ascii)
assert len(_ACCENTS) == 52

_PUNCT = frozenset(range(33, 48)) | frozenset(range(58, 65)) | frozenset(range(91, 97)) | frozenset(range(123, 127))
_SPACE = frozenset(b" \t\n\v\f\r")


def utf8_len(b):
    return _UTF8_LEN[b >> 4]


def strip_accents(s: bytes) -> bytes:
    out = bytearray()
    i = 0
    while i < len(s):
        n = utf8_len(s[i])
        cur = s[i:i + n]                      # substr clips at the end
        a = _ACCENTS.get(bytes(cur))
        if a is not None:
            out.append(a)
        else:
            out += cur
        i += n
    return bytes(out)


def normalize(s: bytes) -> bytes:
    t = bytearray(strip_accents(s))
    i = 0
    while i < len(t):
        c = t[i]
        if 0x41 <= c <= 0x5A:
            t[i] = c + 32
        i += utf8_len(t[i])
    return bytes(t)


def is_chinese_char(s: bytes) -> bool:
    """s = up to 3 bytes starting at a byte whose utf8_len is 3 (bert.cpp:327)."""
    ch = s[0]
    if ch <= 0x7F:
        cp, n = ch, 1
    elif (ch >> 5) == 0x06:
        cp, n = ch & 0x1F, 2
    elif (ch >> 4) == 0x0E:
        cp, n = ch & 0x0F, 3
    elif (ch >> 3) == 0x1E:
        cp, n = ch & 0x07, 4
    else:
        cp, n = 0, 0
    for j in range(1, n):
        if j >= len(s):
            return False
        nb = s[j]
        if (nb >> 6) != 0x02:
            return False
        cp = (cp << 6) | (nb & 0x3F)
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF or
            0x2A700 <= cp <= 0x2B73F or 0x2B740 <= cp <= 0x2B81F or 0x2B920 <= cp <= 0x2CEAF or
            0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F or 0x3000 <= cp <= 0x303F or
            0xFF00 <= cp <= 0xFFEF)


class Vocab:
    def __init__(self, tokens):
        """tokens: list[bytes] in id order (as stored in the model file)."""
        self.tokens = list(tokens)
        self.token_to_id = {}
        self.subword_to_id = {}
        for i, w in enumerate(self.tokens):
            if len(w) >= 2 and w[0:2] == b"##":
                self.subword_to_id[w[2:]] = i          # operator[] assignment: last wins
            elif len(w) == 1 and w == b"#":
                pass                                   # word[1] is the NUL terminator
            if w not in self.token_to_id:              # count()==0: first wins
                self.token_to_id[w] = i


def split_words(text: bytes):
    s = normalize(text)
    new = bytearray()
    i = 0
    while i < len(s):
        n = utf8_len(s[i])
        if n == 1 and s[i] in _PUNCT:
            new += b" " + s[i:i + 1] + b" "
            i += 1
        elif n == 3 and is_chinese_char(s[i:i + 3]):
            new += b" " + s[i:i + 3] + b" "
            i += 3
        else:
            new.append(s[i])
            i += 1
    words, l, r = [], 0, 0
    while r < len(new):
        if new[r] in _SPACE:
            if r > l:
                words.append(bytes(new[l:r]))
            l = r + 1
            r = l
        else:
            r += 1
    if r > l:
        words.append(bytes(new[l:r]))
    return words


def tokenize(vocab: Vocab, text: bytes, n_max_tokens: int, safe: bool = True):
    # the C API takes a NUL-terminated string: std::string(text) stops at the first NUL
    z = text.find(b"\0")
    if z >= 0:
        text = text[:z]
    toks = [CLS]
    for word in split_words(text):
        if not word:
            continue
        prev_t = len(toks)
        i, n = 0, len(word)
        tmap = vocab.token_to_id
        while i < n:
            if len(toks) >= n_max_tokens - 1:
                break
            j = n
            found = False
            while j > i:
                tid = tmap.get(word[i:j])
                if tid is not None:
                    toks.append(tid)
                    i = j
                    tmap = vocab.subword_to_id
                    found = True
                    break
                j -= 1
            if not found:
                tmap = vocab.subword_to_id
                i += 1
        if prev_t == len(toks):
            if safe and len(toks) >= n_max_tokens - 1:
                continue
            toks.append(UNK)
    toks.append(SEP)
    return toks
