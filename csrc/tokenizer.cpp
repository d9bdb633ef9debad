// Native WordPiece tokenizer (reference scripts/train.py:69,75-77,89-92).
//
// The reference calls the HF `tokenizers` Rust library through AutoTokenizer; here the host side
// of the data pipeline is C++ like the rest of the runtime: BERT "basic" pre-tokenisation
// (clean-up, optional lower-casing + accent stripping, CJK isolation, punctuation split) followed
// by greedy longest-match WordPiece, [CLS]/[SEP] framing, truncation and padding into dense
// int32 [n, max_len] id / mask arrays that go straight into pinned staging buffers.
// A batch is encoded by a small pool of std::threads (one contiguous slice of rows each).
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "hstd_b200.h"

namespace {

// ---------------------------------------------------------------------------------------------
// UTF-8 helpers
// ---------------------------------------------------------------------------------------------
inline uint32_t decode_utf8(const unsigned char* s, size_t n, size_t& i) {
    unsigned char c = s[i];
    if (c < 0x80) { i += 1; return c; }
    auto cont = [&](size_t k) { return i + k < n && (s[i + k] & 0xC0) == 0x80; };
    if ((c >> 5) == 0x6 && cont(1)) {
        uint32_t cp = ((c & 0x1Fu) << 6) | (s[i + 1] & 0x3Fu);
        i += 2;
        return cp;
    }
    if ((c >> 4) == 0xE && cont(1) && cont(2)) {
        uint32_t cp = ((c & 0x0Fu) << 12) | ((s[i + 1] & 0x3Fu) << 6) | (s[i + 2] & 0x3Fu);
        i += 3;
        return cp;
    }
    if ((c >> 3) == 0x1E && cont(1) && cont(2) && cont(3)) {
        uint32_t cp = ((c & 0x07u) << 18) | ((s[i + 1] & 0x3Fu) << 12) | ((s[i + 2] & 0x3Fu) << 6) | (s[i + 3] & 0x3Fu);
        i += 4;
        return cp;
    }
    i += 1;
    return 0xFFFD;  // malformed byte
}

inline void append_utf8(std::string& out, uint32_t cp) {
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

inline bool is_whitespace(uint32_t c) {
    return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == 0x00A0 || c == 0x1680 ||
           (c >= 0x2000 && c <= 0x200A) || c == 0x2028 || c == 0x2029 || c == 0x202F || c == 0x205F || c == 0x3000;
}

inline bool is_control(uint32_t c) {
    if (c == '\t' || c == '\n' || c == '\r') return false;
    return c < 0x20 || (c >= 0x7F && c < 0xA0) || c == 0x00AD || (c >= 0x200B && c <= 0x200F) ||
           (c >= 0x202A && c <= 0x202E) || (c >= 0x2060 && c <= 0x2064) || c == 0xFEFF;
}

inline bool is_punctuation(uint32_t c) {
    if ((c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126)) return true;
    // Unicode P* blocks that matter for running text
    if (c == 0x00A1 || c == 0x00A7 || c == 0x00AB || c == 0x00B6 || c == 0x00B7 || c == 0x00BB || c == 0x00BF) return true;
    if (c >= 0x2010 && c <= 0x2027) return true;
    if (c >= 0x2030 && c <= 0x205E && c != 0x2044 && c != 0x2052) return true;
    if (c >= 0x3001 && c <= 0x3003) return true;
    if (c >= 0x3008 && c <= 0x3011) return true;
    if (c >= 0x3014 && c <= 0x301F) return true;
    if (c >= 0xFF01 && c <= 0xFF0F && c != 0xFF04 && c != 0xFF0B) return true;
    if (c >= 0xFF1A && c <= 0xFF20 && !(c >= 0xFF1C && c <= 0xFF1E)) return true;
    return false;
}

inline bool is_cjk(uint32_t c) {
    return (c >= 0x4E00 && c <= 0x9FFF) || (c >= 0x3400 && c <= 0x4DBF) || (c >= 0x20000 && c <= 0x2A6DF) ||
           (c >= 0x2A700 && c <= 0x2B73F) || (c >= 0x2B740 && c <= 0x2B81F) || (c >= 0x2B820 && c <= 0x2CEAF) ||
           (c >= 0xF900 && c <= 0xFAFF) || (c >= 0x2F800 && c <= 0x2FA1F);
}

inline bool is_combining_mark(uint32_t c) { return c >= 0x0300 && c <= 0x036F; }

// Base letter of a precomposed Latin-1 / Latin Extended-A character (what NFD followed by
// dropping the Mn marks leaves), or 0 if the character has no decomposition.
uint32_t strip_accent(uint32_t c) {
    static const char* latin1 =  // U+00C0 .. U+00FF
        "AAAAAA\0CEEEEIIII\0NOOOOO\0\0UUUUY\0\0aaaaaa\0ceeeeiiii\0nooooo\0\0uuuuy\0y";
    if (c >= 0x00C0 && c <= 0x00FF) {
        char b = latin1[c - 0x00C0];
        return b ? static_cast<uint32_t>(b) : 0;
    }
    static const char* ext_a =  // U+0100 .. U+017F
        "AaAaAa" "CcCcCcCc" "Dd" "\0\0" "EeEeEeEeEe" "GgGgGgGg" "Hh" "\0\0" "IiIiIiIi" "I\0" "\0\0" "Jj" "Kk" "\0"
        "LlLlLl" "\0\0" "\0\0" "NnNnNn" "\0" "\0\0" "OoOoOo" "\0\0" "RrRrRr" "SsSsSsSs" "TtTt" "\0\0"
        "UuUuUuUuUuUu" "Ww" "Yy" "Y" "ZzZzZz" "\0";
    if (c >= 0x0100 && c <= 0x017F) {
        char b = ext_a[c - 0x0100];
        return b ? static_cast<uint32_t>(b) : 0;
    }
    return 0;
}

inline uint32_t to_lower(uint32_t c) {
    if (c >= 'A' && c <= 'Z') return c + 32;
    if (c >= 0x00C0 && c <= 0x00DE && c != 0x00D7) return c + 32;           // Latin-1
    if (c >= 0x0100 && c <= 0x017F) {                                        // Latin Extended-A pairs
        if (c == 0x0130) return 'i';
        if (c == 0x0178) return 0x00FF;
        if ((c >= 0x0139 && c <= 0x0148) || (c >= 0x0179 && c <= 0x017E)) return (c & 1) ? c + 1 : c;
        if (c == 0x0138 || c == 0x0149 || c == 0x017F) return c;
        return (c & 1) ? c : c + 1;
    }
    if (c >= 0x0391 && c <= 0x03A9 && c != 0x03A2) return c + 32;           // Greek
    if (c >= 0x0410 && c <= 0x042F) return c + 32;                           // Cyrillic
    if (c >= 0x0400 && c <= 0x040F) return c + 80;
    return c;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Tokenizer object
// ---------------------------------------------------------------------------------------------
struct hstd_tokenizer {
    std::unordered_map<std::string, int32_t> vocab;
    std::vector<std::string> id_to_token;
    bool lower = true;
    bool strip = true;
    int32_t pad_id = 0, unk_id = 1, cls_id = 2, sep_id = 3;
    int max_chars_per_word = 100;
    std::string prefix = "##";

    int32_t lookup(const std::string& t) const {
        auto it = vocab.find(t);
        return it == vocab.end() ? -1 : it->second;
    }

    // clean-up + lower-casing + accent stripping + CJK / punctuation isolation -> words
    void basic_tokenize(const char* text, std::vector<std::string>& words) const {
        const unsigned char* s = reinterpret_cast<const unsigned char*>(text);
        const size_t n = std::strlen(text);
        std::string cur;
        auto flush = [&]() {
            if (!cur.empty()) {
                words.push_back(cur);
                cur.clear();
            }
        };
        size_t i = 0;
        while (i < n) {
            uint32_t c = decode_utf8(s, n, i);
            if (c == 0 || c == 0xFFFD || is_control(c)) continue;
            if (is_whitespace(c)) {
                flush();
                continue;
            }
            if (is_cjk(c) || is_punctuation(c)) {
                flush();
                append_utf8(cur, c);
                flush();
                continue;
            }
            if (strip) {
                if (is_combining_mark(c)) continue;
                uint32_t b = strip_accent(c);
                if (b) c = b;
            }
            if (lower) c = to_lower(c);
            if (strip) {  // lower-casing can expose another precomposed letter (e.g. U+00C0 -> U+00E0)
                uint32_t b = strip_accent(c);
                if (b) c = b;
            }
            append_utf8(cur, c);
        }
        flush();
    }

    // greedy longest-match-first over UTF-8 character boundaries
    void wordpiece(const std::string& w, std::vector<int32_t>& out) const {
        std::vector<size_t> bounds;  // byte offset of every character start, plus the end
        for (size_t i = 0; i < w.size();) {
            bounds.push_back(i);
            unsigned char c = static_cast<unsigned char>(w[i]);
            i += c < 0x80 ? 1 : (c >> 5) == 0x6 ? 2 : (c >> 4) == 0xE ? 3 : 4;
        }
        bounds.push_back(w.size());
        const size_t nchar = bounds.size() - 1;
        if (static_cast<int>(nchar) > max_chars_per_word) {
            out.push_back(unk_id);
            return;
        }
        const size_t mark = out.size();
        size_t start = 0;
        std::string piece;
        while (start < nchar) {
            size_t end = nchar;
            int32_t found = -1;
            while (start < end) {
                piece.clear();
                if (start > 0) piece = prefix;
                piece.append(w, bounds[start], bounds[end] - bounds[start]);
                found = lookup(piece);
                if (found >= 0) break;
                --end;
            }
            if (found < 0) {  // the whole word becomes a single [UNK]
                out.resize(mark);
                out.push_back(unk_id);
                return;
            }
            out.push_back(found);
            start = end;
        }
    }

    // one row of the dense output; returns the untruncated length (specials included)
    int32_t encode_row(const char* text, int max_len, bool add_special, int32_t* ids, int32_t* mask) const {
        std::vector<std::string> words;
        std::vector<int32_t> toks;
        basic_tokenize(text, words);
        for (const auto& w : words) wordpiece(w, toks);
        const int specials = add_special ? 2 : 0;
        const int32_t full = static_cast<int32_t>(toks.size()) + specials;
        int keep = std::min<int>(static_cast<int>(toks.size()), std::max(0, max_len - specials));
        int p = 0;
        if (add_special && p < max_len) ids[p++] = cls_id;
        for (int k = 0; k < keep && p < max_len; ++k) ids[p++] = toks[k];
        if (add_special && p < max_len) ids[p++] = sep_id;
        for (int k = 0; k < p; ++k) mask[k] = 1;
        for (int k = p; k < max_len; ++k) {
            ids[k] = pad_id;
            mask[k] = 0;
        }
        return full;
    }
};

extern "C" HSTD_API hstd_tokenizer* hstd_tok_create(const char* vocab_path, int do_lower_case, int strip_accents) {
    std::ifstream in(vocab_path);
    if (!in) {
        hstd::set_error(-2, std::string("hstd_tok_create: cannot open ") + vocab_path);
        return nullptr;
    }
    auto* t = new hstd_tokenizer();
    t->lower = do_lower_case != 0;
    t->strip = strip_accents != 0;
    std::string line;
    while (std::getline(in, line)) {
        while (!line.empty() && (line.back() == '\r' || line.back() == '\n')) line.pop_back();
        const int32_t id = static_cast<int32_t>(t->id_to_token.size());
        t->id_to_token.push_back(line);
        t->vocab.emplace(line, id);  // first occurrence wins, like the HF loader
    }
    if (t->id_to_token.empty()) {
        delete t;
        hstd::set_error(-2, std::string("hstd_tok_create: empty vocabulary ") + vocab_path);
        return nullptr;
    }
    auto special = [&](const char* name, int32_t fallback) {
        int32_t id = t->lookup(name);
        return id >= 0 ? id : fallback;
    };
    t->pad_id = special("[PAD]", 0);
    t->unk_id = special("[UNK]", 1);
    t->cls_id = special("[CLS]", 2);
    t->sep_id = special("[SEP]", 3);
    return t;
}

extern "C" HSTD_API void hstd_tok_destroy(hstd_tokenizer* t) { delete t; }

extern "C" HSTD_API int hstd_tok_vocab_size(const hstd_tokenizer* t) { return t ? static_cast<int>(t->id_to_token.size()) : 0; }

extern "C" HSTD_API int hstd_tok_token_to_id(const hstd_tokenizer* t, const char* token) {
    return t ? t->lookup(token) : -1;
}

extern "C" HSTD_API const char* hstd_tok_id_to_token(const hstd_tokenizer* t, int id) {
    if (!t || id < 0 || id >= static_cast<int>(t->id_to_token.size())) return nullptr;
    return t->id_to_token[id].c_str();
}

extern "C" HSTD_API int hstd_tok_special_ids(const hstd_tokenizer* t, int32_t* out4) {
    if (!t) return hstd::set_error(-1, "hstd_tok_special_ids: null tokenizer");
    out4[0] = t->pad_id;
    out4[1] = t->unk_id;
    out4[2] = t->cls_id;
    out4[3] = t->sep_id;
    return 0;
}

extern "C" HSTD_API int hstd_tok_encode_batch(const hstd_tokenizer* t, const char* const* texts, int n, int max_len,
                                              int add_special_tokens, int32_t* ids, int32_t* mask, int32_t* lengths,
                                              int num_threads) {
    if (!t) return hstd::set_error(-1, "hstd_tok_encode_batch: null tokenizer");
    if (n < 0 || max_len <= 0) return hstd::set_error(-1, "hstd_tok_encode_batch: n >= 0 and max_len > 0 required");
    if (add_special_tokens && max_len < 2) return hstd::set_error(-1, "hstd_tok_encode_batch: max_len < 2 with specials");
    if (num_threads <= 0) num_threads = static_cast<int>(std::min<unsigned>(16u, std::max(1u, std::thread::hardware_concurrency())));
    num_threads = std::max(1, std::min(num_threads, (n + 63) / 64));
    auto work = [&](int lo, int hi) {
        for (int r = lo; r < hi; ++r) {
            int32_t full = t->encode_row(texts[r] ? texts[r] : "", max_len, add_special_tokens != 0,
                                         ids + static_cast<int64_t>(r) * max_len, mask + static_cast<int64_t>(r) * max_len);
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
        int lo = k * per, hi = std::min(n, lo + per);
        if (lo < hi) pool.emplace_back(work, lo, hi);
    }
    for (auto& th : pool) th.join();
    return 0;
}
