// Native byte-level BPE tokenizer (RoBERTa / GPT-2 family; SURVEY.md §2.3 N6, reference
// scripts/train.py:69,75-77 with a roberta-* checkpoint).
//
// Same pipeline as the HF `tokenizers` ByteLevel + BPE stack the reference reaches through
// AutoTokenizer: GPT-2 pre-tokenisation (contractions | optional space + letters | optional space +
// digits | optional space + other | whitespace), every byte of a piece mapped to a printable code
// point, lowest-rank-first pair merging from merges.txt, ids from vocab.json, <s> ... </s> framing,
// truncation and padding into dense int32 [n, max_len] arrays.  Batches are split over std::threads;
// every thread keeps its own piece -> ids cache.
#include <algorithm>
#include <climits>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "hstd_b200.h"

namespace {

inline void put_utf8(std::string& out, uint32_t cp) {
    if (cp < 0x80) {
        out.push_back(static_cast<char>(cp));
    } else if (cp < 0x800) {
        out.push_back(static_cast<char>(0xC0 | (cp >> 6)));
        out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else if (cp < 0x10000) {
        out.push_back(static_cast<char>(0xE0 | (cp >> 12)));
        out.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
        out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else {
        out.push_back(static_cast<char>(0xF0 | (cp >> 18)));
        out.push_back(static_cast<char>(0x80 | ((cp >> 12) & 0x3F)));
        out.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
        out.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    }
}

// code point + byte length at s[i]; malformed bytes come back as single "other" characters
inline uint32_t get_utf8(const unsigned char* s, size_t n, size_t i, int& len) {
    unsigned char c = s[i];
    auto cont = [&](size_t k) { return i + k < n && (s[i + k] & 0xC0) == 0x80; };
    if (c < 0x80) { len = 1; return c; }
    if ((c >> 5) == 0x6 && cont(1)) { len = 2; return ((c & 0x1Fu) << 6) | (s[i + 1] & 0x3Fu); }
    if ((c >> 4) == 0xE && cont(1) && cont(2)) {
        len = 3;
        return ((c & 0x0Fu) << 12) | ((s[i + 1] & 0x3Fu) << 6) | (s[i + 2] & 0x3Fu);
    }
    if ((c >> 3) == 0x1E && cont(1) && cont(2) && cont(3)) {
        len = 4;
        return ((c & 0x07u) << 18) | ((s[i + 1] & 0x3Fu) << 12) | ((s[i + 2] & 0x3Fu) << 6) | (s[i + 3] & 0x3Fu);
    }
    len = 1;
    return 0xFFFD;
}

enum CharClass { CC_SPACE, CC_LETTER, CC_NUMBER, CC_OTHER };

// \s, \p{L}, \p{N} of the GPT-2 pattern.  Exact for ASCII and Latin-1; beyond that the blocks that
// are letters in running text (Latin Extended, Greek, Cyrillic, Hebrew, Arabic, Indic, kana, CJK,
// Hangul) are letters, digit blocks are numbers, and the punctuation / symbol / emoji blocks are "other".
CharClass classify(uint32_t c) {
    if (c < 0x80) {
        if (c == ' ' || (c >= 0x09 && c <= 0x0D)) return CC_SPACE;
        if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z')) return CC_LETTER;
        if (c >= '0' && c <= '9') return CC_NUMBER;
        return CC_OTHER;
    }
    if (c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) || c == 0x2028 || c == 0x2029 || c == 0x202F ||
        c == 0x205F || c == 0x3000)
        return CC_SPACE;
    if (c < 0x100) {
        if (c == 0xAA || c == 0xB5 || c == 0xBA) return CC_LETTER;
        if (c == 0xB2 || c == 0xB3 || c == 0xB9 || (c >= 0xBC && c <= 0xBE)) return CC_NUMBER;
        if (c >= 0xC0 && c != 0xD7 && c != 0xF7) return CC_LETTER;
        return CC_OTHER;
    }
    if (c < 0x2000) {
        if ((c >= 0x0660 && c <= 0x0669) || (c >= 0x06F0 && c <= 0x06F9) || (c >= 0x0966 && c <= 0x096F)) return CC_NUMBER;
        if ((c >= 0x02C2 && c <= 0x02C5) || (c >= 0x02D2 && c <= 0x02DF) || (c >= 0x0300 && c <= 0x036F) || c == 0x037E ||
            c == 0x0387 || (c >= 0x0482 && c <= 0x0489) || (c >= 0x055A && c <= 0x055F) || c == 0x0589 || c == 0x058A ||
            (c >= 0x0591 && c <= 0x05C7) || (c >= 0x0600 && c <= 0x061F) || (c >= 0x064B && c <= 0x065F) || c == 0x06D4)
            return CC_OTHER;
        return CC_LETTER;
    }
    if (c < 0x2E80) {
        if ((c >= 0x2070 && c <= 0x2079) || (c >= 0x2080 && c <= 0x2089) || (c >= 0x2150 && c <= 0x2189) ||
            (c >= 0x2460 && c <= 0x249B) || (c >= 0x24EA && c <= 0x24FF) || (c >= 0x2776 && c <= 0x2793))
            return CC_NUMBER;
        if (c == 0x2071 || c == 0x207F || (c >= 0x2090 && c <= 0x209C) || c == 0x2102 || c == 0x2107 || (c >= 0x210A && c <= 0x2113) ||
            c == 0x2115 || (c >= 0x2119 && c <= 0x211D) || c == 0x2124 || c == 0x2126 || c == 0x2128 || (c >= 0x212A && c <= 0x212D) ||
            (c >= 0x212F && c <= 0x2139) || (c >= 0x2C00 && c <= 0x2DFF))
            return CC_LETTER;
        return CC_OTHER;
    }
    if (c >= 0x3001 && c <= 0x3004) return CC_OTHER;
    if (c >= 0x3008 && c <= 0x3020) return CC_OTHER;
    if (c == 0x3030 || c == 0x303D || c == 0x30A0 || c == 0x30FB) return CC_OTHER;
    if (c >= 0x3021 && c <= 0x3029) return CC_NUMBER;
    if (c >= 0x3040 && c <= 0x318F) return CC_LETTER;
    if (c >= 0x3190 && c <= 0x33FF) return CC_OTHER;
    if (c >= 0x3400 && c <= 0x4DBF) return CC_LETTER;
    if (c >= 0x4DC0 && c <= 0x4DFF) return CC_OTHER;
    if (c >= 0x4E00 && c <= 0xA4CF) return CC_LETTER;
    if (c >= 0xAC00 && c <= 0xD7AF) return CC_LETTER;
    if (c >= 0xF900 && c <= 0xFAFF) return CC_LETTER;
    if (c >= 0xFF10 && c <= 0xFF19) return CC_NUMBER;
    if ((c >= 0xFF21 && c <= 0xFF3A) || (c >= 0xFF41 && c <= 0xFF5A) || (c >= 0xFF66 && c <= 0xFFDC)) return CC_LETTER;
    if (c >= 0x20000 && c <= 0x3134F) return CC_LETTER;
    return CC_OTHER;  // private use, specials, emoji, malformed bytes
}

// Minimal reader for the flat {"token": id, ...} object of vocab.json.
bool parse_vocab_json(const std::string& text, std::unordered_map<std::string, int32_t>& vocab, std::string& err) {
    size_t i = 0;
    const size_t n = text.size();
    auto ws = [&]() { while (i < n && (text[i] == ' ' || text[i] == '\n' || text[i] == '\r' || text[i] == '\t')) ++i; };
    ws();
    if (i >= n || text[i] != '{') { err = "vocab.json: expected '{'"; return false; }
    ++i;
    for (;;) {
        ws();
        if (i < n && text[i] == '}') return true;
        if (i >= n || text[i] != '"') { err = "vocab.json: expected a string key"; return false; }
        ++i;
        std::string key;
        while (i < n && text[i] != '"') {
            char c = text[i++];
            if (c != '\\') { key.push_back(c); continue; }
            if (i >= n) break;
            char e = text[i++];
            switch (e) {
                case 'n': key.push_back('\n'); break;
                case 't': key.push_back('\t'); break;
                case 'r': key.push_back('\r'); break;
                case 'b': key.push_back('\b'); break;
                case 'f': key.push_back('\f'); break;
                case 'u': {
                    auto hex4 = [&](uint32_t& v) {
                        if (i + 4 > n) return false;
                        v = 0;
                        for (int k = 0; k < 4; ++k) {
                            char h = text[i++];
                            v = v * 16 + (h >= '0' && h <= '9' ? h - '0' : (h >= 'a' && h <= 'f') ? h - 'a' + 10 : (h >= 'A' && h <= 'F') ? h - 'A' + 10 : 0);
                        }
                        return true;
                    };
                    uint32_t cp = 0;
                    if (!hex4(cp)) { err = "vocab.json: bad \\u escape"; return false; }
                    if (cp >= 0xD800 && cp <= 0xDBFF && i + 6 <= n && text[i] == '\\' && text[i + 1] == 'u') {
                        i += 2;
                        uint32_t lo = 0;
                        if (!hex4(lo)) { err = "vocab.json: bad surrogate"; return false; }
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    }
                    put_utf8(key, cp);
                    break;
                }
                default: key.push_back(e);  // \" \\ \/
            }
        }
        if (i >= n) { err = "vocab.json: unterminated string"; return false; }
        ++i;  // closing quote
        ws();
        if (i >= n || text[i] != ':') { err = "vocab.json: expected ':'"; return false; }
        ++i;
        ws();
        bool neg = false;
        if (i < n && text[i] == '-') { neg = true; ++i; }
        if (i >= n || text[i] < '0' || text[i] > '9') { err = "vocab.json: expected an integer id"; return false; }
        long v = 0;
        while (i < n && text[i] >= '0' && text[i] <= '9') v = v * 10 + (text[i++] - '0');
        vocab[key] = static_cast<int32_t>(neg ? -v : v);
        ws();
        if (i < n && text[i] == ',') { ++i; continue; }
        if (i < n && text[i] == '}') return true;
        err = "vocab.json: expected ',' or '}'";
        return false;
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Tokenizer object
// ---------------------------------------------------------------------------------------------
struct hstd_bpe {
    std::unordered_map<std::string, int32_t> vocab;
    std::vector<std::string> id_to_token;
    std::unordered_map<std::string, int32_t> ranks;  // "left right" -> merge priority
    std::string byte_char[256];                       // GPT-2 bytes_to_unicode, UTF-8 encoded
    int32_t bos_id = 0, pad_id = 1, eos_id = 2, unk_id = 3;

    typedef std::unordered_map<std::string, std::vector<int32_t>> Cache;

    int32_t lookup(const std::string& t) const {
        auto it = vocab.find(t);
        return it == vocab.end() ? -1 : it->second;
    }

    void init_byte_table() {
        // printable bytes map to themselves, the rest to U+0100.. in order (Ġ = space, Ċ = newline)
        int extra = 0;
        for (int b = 0; b < 256; ++b) {
            const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
            put_utf8(byte_char[b], keep ? static_cast<uint32_t>(b) : 256u + static_cast<uint32_t>(extra++));
        }
    }

    // GPT-2 pattern: 's|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+
    void pretokenize(const char* text, std::vector<std::pair<size_t, size_t>>& pieces) const {
        const unsigned char* s = reinterpret_cast<const unsigned char*>(text);
        const size_t n = std::strlen(text);
        size_t pos = 0;
        int len = 0;
        while (pos < n) {
            const uint32_t c = get_utf8(s, n, pos, len);
            if (c == '\'' && pos + 1 < n) {
                const unsigned char a = s[pos + 1], b = pos + 2 < n ? s[pos + 2] : 0;
                size_t m = 0;
                if (a == 's' || a == 't') m = 2;
                else if (a == 'r' && b == 'e') m = 3;
                else if (a == 'v' && b == 'e') m = 3;
                else if (a == 'm') m = 2;
                else if (a == 'l' && b == 'l') m = 3;
                else if (a == 'd') m = 2;
                if (m) {
                    pieces.emplace_back(pos, pos + m);
                    pos += m;
                    continue;
                }
            }
            // optional single space, then a run of one class (letters | numbers | other)
            size_t q = pos;
            uint32_t cq = c;
            int lq = len;
            if (c == ' ' && pos + 1 < n) {
                int l2 = 0;
                const uint32_t c2 = get_utf8(s, n, pos + 1, l2);
                if (classify(c2) != CC_SPACE) {
                    q = pos + 1;
                    cq = c2;
                    lq = l2;
                }
            }
            const CharClass cls = classify(cq);
            if (cls != CC_SPACE) {
                size_t e = q + lq;
                while (e < n) {
                    int l3 = 0;
                    const uint32_t c3 = get_utf8(s, n, e, l3);
                    if (classify(c3) != cls) break;
                    e += l3;
                }
                pieces.emplace_back(pos, e);
                pos = e;
                continue;
            }
            // whitespace run [pos, e); `last` = start of its final character
            size_t e = pos, last = pos;
            while (e < n) {
                int l3 = 0;
                const uint32_t c3 = get_utf8(s, n, e, l3);
                if (classify(c3) != CC_SPACE) break;
                last = e;
                e += l3;
            }
            if (e == n || last == pos) {  // \s+(?!\S) at the end of the text, or a single character (\s+)
                pieces.emplace_back(pos, e);
                pos = e;
            } else {  // \s+(?!\S): everything but the last whitespace character, which joins what follows
                pieces.emplace_back(pos, last);
                pos = last;
            }
        }
    }

    void bpe(const std::string& piece, std::vector<int32_t>& out, Cache& cache) const {
        auto hit = cache.find(piece);
        if (hit != cache.end()) {
            out.insert(out.end(), hit->second.begin(), hit->second.end());
            return;
        }
        std::vector<std::string> sym;
        sym.reserve(piece.size());
        for (unsigned char b : piece) sym.push_back(byte_char[b]);
        std::string key;
        while (sym.size() > 1) {
            int best = INT_MAX;
            for (size_t i = 0; i + 1 < sym.size(); ++i) {
                key.assign(sym[i]);
                key.push_back(' ');
                key.append(sym[i + 1]);
                auto it = ranks.find(key);
                if (it != ranks.end() && it->second < best) best = it->second;
            }
            if (best == INT_MAX) break;
            // merge every occurrence of the best-ranked pair, left to right
            std::vector<std::string> next;
            next.reserve(sym.size());
            for (size_t i = 0; i < sym.size();) {
                if (i + 1 < sym.size()) {
                    key.assign(sym[i]);
                    key.push_back(' ');
                    key.append(sym[i + 1]);
                    auto it = ranks.find(key);
                    if (it != ranks.end() && it->second == best) {
                        next.push_back(sym[i] + sym[i + 1]);
                        i += 2;
                        continue;
                    }
                }
                next.push_back(sym[i]);
                ++i;
            }
            sym.swap(next);
        }
        std::vector<int32_t> ids;
        ids.reserve(sym.size());
        for (const auto& t : sym) {
            const int32_t id = lookup(t);
            ids.push_back(id >= 0 ? id : unk_id);
        }
        out.insert(out.end(), ids.begin(), ids.end());
        if (cache.size() < 200000) cache.emplace(piece, std::move(ids));
    }

    int32_t encode_row(const char* text, int max_len, bool add_special, int32_t* ids, int32_t* mask, Cache& cache) const {
        std::vector<std::pair<size_t, size_t>> pieces;
        std::vector<int32_t> toks;
        pretokenize(text, pieces);
        std::string piece;
        for (const auto& pr : pieces) {
            piece.assign(text + pr.first, pr.second - pr.first);
            bpe(piece, toks, cache);
        }
        const int specials = add_special ? 2 : 0;
        const int32_t full = static_cast<int32_t>(toks.size()) + specials;
        const int keep = std::min<int>(static_cast<int>(toks.size()), std::max(0, max_len - specials));
        int p = 0;
        if (add_special && p < max_len) ids[p++] = bos_id;
        for (int k = 0; k < keep && p < max_len; ++k) ids[p++] = toks[k];
        if (add_special && p < max_len) ids[p++] = eos_id;
        for (int k = 0; k < p; ++k) mask[k] = 1;
        for (int k = p; k < max_len; ++k) {
            ids[k] = pad_id;
            mask[k] = 0;
        }
        return full;
    }
};

extern "C" HSTD_API hstd_bpe* hstd_bpe_create(const char* vocab_json_path, const char* merges_path) {
    std::ifstream vf(vocab_json_path, std::ios::binary);
    if (!vf) {
        hstd::set_error(-2, std::string("hstd_bpe_create: cannot open ") + vocab_json_path);
        return nullptr;
    }
    std::stringstream buf;
    buf << vf.rdbuf();
    auto* t = new hstd_bpe();
    std::string err;
    if (!parse_vocab_json(buf.str(), t->vocab, err) || t->vocab.empty()) {
        delete t;
        hstd::set_error(-2, "hstd_bpe_create: " + (err.empty() ? std::string("empty vocabulary") : err));
        return nullptr;
    }
    int32_t max_id = -1;
    for (const auto& kv : t->vocab) max_id = std::max(max_id, kv.second);
    t->id_to_token.assign(static_cast<size_t>(max_id) + 1, std::string());
    for (const auto& kv : t->vocab)
        if (kv.second >= 0) t->id_to_token[kv.second] = kv.first;
    std::ifstream mf(merges_path);
    if (!mf) {
        delete t;
        hstd::set_error(-2, std::string("hstd_bpe_create: cannot open ") + merges_path);
        return nullptr;
    }
    std::string line;
    int32_t rank = 0;
    while (std::getline(mf, line)) {
        while (!line.empty() && (line.back() == '\r' || line.back() == '\n')) line.pop_back();
        if (line.empty() || line.rfind("#version", 0) == 0) continue;
        if (line.find(' ') == std::string::npos) continue;
        t->ranks.emplace(line, rank++);
    }
    t->init_byte_table();
    auto special = [&](const char* name, int32_t fallback) {
        const int32_t id = t->lookup(name);
        return id >= 0 ? id : fallback;
    };
    t->bos_id = special("<s>", 0);
    t->pad_id = special("<pad>", 1);
    t->eos_id = special("</s>", 2);
    t->unk_id = special("<unk>", 3);
    return t;
}

extern "C" HSTD_API void hstd_bpe_destroy(hstd_bpe* t) { delete t; }

extern "C" HSTD_API int hstd_bpe_vocab_size(const hstd_bpe* t) { return t ? static_cast<int>(t->id_to_token.size()) : 0; }

extern "C" HSTD_API int hstd_bpe_token_to_id(const hstd_bpe* t, const char* token) { return t ? t->lookup(token) : -1; }

extern "C" HSTD_API const char* hstd_bpe_id_to_token(const hstd_bpe* t, int id) {
    if (!t || id < 0 || id >= static_cast<int>(t->id_to_token.size())) return nullptr;
    return t->id_to_token[id].c_str();
}

extern "C" HSTD_API int hstd_bpe_special_ids(const hstd_bpe* t, int32_t* out4) {
    if (!t) return hstd::set_error(-1, "hstd_bpe_special_ids: null tokenizer");
    out4[0] = t->pad_id;
    out4[1] = t->unk_id;
    out4[2] = t->bos_id;
    out4[3] = t->eos_id;
    return 0;
}

extern "C" HSTD_API int hstd_bpe_encode_batch(const hstd_bpe* t, const char* const* texts, int n, int max_len,
                                              int add_special_tokens, int32_t* ids, int32_t* mask, int32_t* lengths,
                                              int num_threads) {
    if (!t) return hstd::set_error(-1, "hstd_bpe_encode_batch: null tokenizer");
    if (n < 0 || max_len <= 0) return hstd::set_error(-1, "hstd_bpe_encode_batch: n >= 0 and max_len > 0 required");
    if (add_special_tokens && max_len < 2) return hstd::set_error(-1, "hstd_bpe_encode_batch: max_len < 2 with specials");
    if (num_threads <= 0) num_threads = static_cast<int>(std::min<unsigned>(16u, std::max(1u, std::thread::hardware_concurrency())));
    num_threads = std::max(1, std::min(num_threads, (n + 63) / 64));
    auto work = [&](int lo, int hi) {
        hstd_bpe::Cache cache;
        for (int r = lo; r < hi; ++r) {
            const int32_t full = t->encode_row(texts[r] ? texts[r] : "", max_len, add_special_tokens != 0,
                                               ids + static_cast<int64_t>(r) * max_len, mask + static_cast<int64_t>(r) * max_len, cache);
            if (lengths) lengths[r] = full;
        }
    };
    if (num_threads == 1) {
        work(0, n);
        return 0;
    }
    std::vector<std::thread> pool;
    const int per = (n + num_threads - 1) / num_threads;
    for (int k = 0; k < num_threads; ++k) {
        const int lo = k * per, hi = std::min(n, lo + per);
        if (lo < hi) pool.emplace_back(work, lo, hi);
    }
    for (auto& th : pool) th.join();
    return 0;
}
