// Host-side WordPiece tokenizer (at ~3x10^5 embeddings/s one thread of tokenisation
// does not keep up with the encoder).  Restates what the reference runs before the model: DJL's
// HuggingFaceTokenizer, i.e. the HuggingFace `tokenizers` library with a BERT pipeline, installed by
// TextEmbeddingTranslatorFactory (AbstractHuggingFaceEmbeddingService.java:20,155) -
//   BertNormalizer    clean_text, handle_chinese_chars, strip_accents (= lowercase when unset), lowercase
//   BertPreTokenizer  split on whitespace, isolate punctuation
//   WordPiece         greedy longest match, "##" continuation prefix, [UNK] for a word that cannot be split
//                     or has more than 100 characters
//   template          [CLS] A [SEP], truncated to max_length
// and writes token ids directly in the packed layout lsc_encoder_embed takes (ids + cu_seqlens), on a pool of
// threads.  No GPU work here; it lives in this library because it feeds the C ABI of the hot path.
// Unicode properties come from generated tables (tools/gen_unicode_tables.py), not from a Unicode library.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "common.h"

namespace lsc {
namespace {

struct UniRange { uint32_t lo, hi; uint8_t flags; };
struct UniDecomp { uint32_t cp; uint16_t off; uint8_t n; };
struct UniCcc { uint32_t cp; uint8_t ccc; };
struct UniLower { uint32_t cp; uint32_t to[3]; };
constexpr uint8_t F_OTHER = 1, F_MN = 2, F_PUNCT = 4, F_SPACE = 8, F_CJK = 16;
#include "unicode_tables.inc"

uint8_t uni_flags(uint32_t cp) {
    const UniRange* b = kUniRanges;
    const UniRange* e = kUniRanges + sizeof(kUniRanges) / sizeof(kUniRanges[0]);
    const UniRange* it = std::upper_bound(b, e, cp, [](uint32_t v, const UniRange& r) { return v < r.lo; });
    if (it == b) return 0;
    --it;
    return cp <= it->hi ? it->flags : 0;
}

// The classes below are what the library's BertNormalizer / BertPreTokenizer do with each code point, as
// probed by tools/gen_unicode_tables.py (F_SPACE: mapped to a blank / split on; F_OTHER: removed by clean_text;
// F_CJK: padded with blanks; F_PUNCT: isolated; F_MN: dropped by strip_accents).
// ASCII classes as a table: bit 0 white space, bit 1 punctuation, bit 2 removed by clean_text (the comparisons
// below, evaluated once)
struct AsciiClass {
    uint8_t t[128];
    constexpr AsciiClass() : t() {
        for (int c = 0; c < 128; ++c) {
            const bool ws = c == ' ' || (c >= 0x09 && c <= 0x0D);
            const bool pu = (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
            const bool rm = c < 0x20 ? !(c == '\t' || c == '\n' || c == '\r') : c == 0x7F;  // clean_text removes these ...
            t[c] = static_cast<uint8_t>((ws && !rm ? 1 : 0) | (pu ? 2 : 0) | (rm ? 4 : 0));  // ... before blanks are looked at
        }
    }
};
constexpr AsciiClass kAscii;

inline bool is_white_space(uint32_t c) {
    if (c < 0x80) return (kAscii.t[c] & 1) != 0;
    return (uni_flags(c) & F_SPACE) != 0;
}
bool bert_is_removed(uint32_t c) {
    if (c < 0x80) return c < 0x20 ? !(c == '\t' || c == '\n' || c == '\r') : c == 0x7F;
    return (uni_flags(c) & F_OTHER) != 0;
}
bool is_chinese_char(uint32_t c) { return c >= 0x2E80 && (uni_flags(c) & F_CJK) != 0; }
inline bool bert_is_punctuation(uint32_t c) {
    if (c < 0x80) return (kAscii.t[c] & 2) != 0;
    return (uni_flags(c) & F_PUNCT) != 0;
}

uint8_t combining_class(uint32_t cp) {
    const UniCcc* b = kUniCcc;
    const UniCcc* e = kUniCcc + sizeof(kUniCcc) / sizeof(kUniCcc[0]);
    const UniCcc* it = std::lower_bound(b, e, cp, [](const UniCcc& r, uint32_t v) { return r.cp < v; });
    return (it != e && it->cp == cp) ? it->ccc : 0;
}

// canonical decomposition of one code point, appended to out (Hangul syllables algorithmically)
void decompose(uint32_t cp, std::vector<uint32_t>& out) {
    if (cp >= 0xAC00 && cp <= 0xD7A3) {
        const uint32_t s = cp - 0xAC00;
        out.push_back(0x1100 + s / 588);
        out.push_back(0x1161 + (s % 588) / 28);
        if (s % 28) out.push_back(0x11A7 + s % 28);
        return;
    }
    if (cp < 0xC0) {
        out.push_back(cp);
        return;
    }
    const UniDecomp* b = kUniDecomp;
    const UniDecomp* e = kUniDecomp + sizeof(kUniDecomp) / sizeof(kUniDecomp[0]);
    const UniDecomp* it = std::lower_bound(b, e, cp, [](const UniDecomp& r, uint32_t v) { return r.cp < v; });
    if (it != e && it->cp == cp) {
        for (int i = 0; i < it->n; ++i) out.push_back(kUniDecompPool[it->off + i]);
    } else {
        out.push_back(cp);
    }
}

// NFD: decompose, then order every run of combining marks by combining class (stable)
void nfd(const std::vector<uint32_t>& in, std::vector<uint32_t>& out) {
    out.clear();
    out.reserve(in.size() + 8);
    for (uint32_t c : in) decompose(c, out);
    size_t i = 0;
    while (i < out.size()) {
        if (out[i] < 0x300 || combining_class(out[i]) == 0) {
            ++i;
            continue;
        }
        size_t j = i;
        while (j < out.size() && out[j] >= 0x300 && combining_class(out[j]) != 0) ++j;
        if (j - i > 1)
            std::stable_sort(out.begin() + i, out.begin() + j,
                             [](uint32_t a, uint32_t b) { return combining_class(a) < combining_class(b); });
        i = j;
    }
}

void lowercase_append(uint32_t cp, std::vector<uint32_t>& out) {
    if (cp < 0x80) {
        out.push_back((cp >= 'A' && cp <= 'Z') ? cp + 32 : cp);
        return;
    }
    const UniLower* b = kUniLower;
    const UniLower* e = kUniLower + sizeof(kUniLower) / sizeof(kUniLower[0]);
    const UniLower* it = std::lower_bound(b, e, cp, [](const UniLower& r, uint32_t v) { return r.cp < v; });
    if (it != e && it->cp == cp) {
        for (int i = 0; i < 3 && it->to[i]; ++i) out.push_back(it->to[i]);
    } else {
        out.push_back(cp);
    }
}

// UTF-8 -> code points; malformed bytes become U+FFFD (which clean_text then removes)
void utf8_decode(const char* s, size_t n, std::vector<uint32_t>& out) {
    out.clear();
    out.reserve(n);
    const unsigned char* p = reinterpret_cast<const unsigned char*>(s);
    size_t i = 0;
    while (i < n) {
        const unsigned char c = p[i];
        uint32_t cp = 0xFFFD;
        int len = 1;
        if (c < 0x80) {
            cp = c;
        } else if (c >= 0xC2 && c <= 0xDF && i + 1 < n && (p[i + 1] & 0xC0) == 0x80) {
            cp = ((c & 0x1Fu) << 6) | (p[i + 1] & 0x3Fu);
            len = 2;
        } else if (c >= 0xE0 && c <= 0xEF && i + 2 < n && (p[i + 1] & 0xC0) == 0x80 && (p[i + 2] & 0xC0) == 0x80) {
            const uint32_t v = ((c & 0x0Fu) << 12) | ((p[i + 1] & 0x3Fu) << 6) | (p[i + 2] & 0x3Fu);
            if (v >= 0x800 && !(v >= 0xD800 && v <= 0xDFFF)) {
                cp = v;
                len = 3;
            }
        } else if (c >= 0xF0 && c <= 0xF4 && i + 3 < n && (p[i + 1] & 0xC0) == 0x80 && (p[i + 2] & 0xC0) == 0x80 &&
                   (p[i + 3] & 0xC0) == 0x80) {
            const uint32_t v = ((c & 0x07u) << 18) | ((p[i + 1] & 0x3Fu) << 12) | ((p[i + 2] & 0x3Fu) << 6) | (p[i + 3] & 0x3Fu);
            if (v >= 0x10000 && v <= 0x10FFFF) {
                cp = v;
                len = 4;
            }
        }
        out.push_back(cp);
        i += len;
    }
}

void utf8_append(uint32_t cp, std::string& s) {
    if (cp < 0x80) {
        s.push_back(static_cast<char>(cp));
    } else if (cp < 0x800) {
        s.push_back(static_cast<char>(0xC0 | (cp >> 6)));
        s.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else if (cp < 0x10000) {
        s.push_back(static_cast<char>(0xE0 | (cp >> 12)));
        s.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
        s.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    } else {
        s.push_back(static_cast<char>(0xF0 | (cp >> 18)));
        s.push_back(static_cast<char>(0x80 | ((cp >> 12) & 0x3F)));
        s.push_back(static_cast<char>(0x80 | ((cp >> 6) & 0x3F)));
        s.push_back(static_cast<char>(0x80 | (cp & 0x3F)));
    }
}

}  // namespace
}  // namespace lsc

using namespace lsc;

namespace lsc {
namespace {

// Vocabulary lookup for WordPiece's greedy longest match: an open-addressing table keyed by a POLYNOMIAL hash of the
// token's bytes, so that the hash of any substring of a word comes from the word's prefix hashes in O(1)
// (h[s, e) = ph[e] - ph[s] * B^(e - s) mod 2^64) and trying "the longest piece that is in the vocabulary" costs one
// probe per candidate length - no std::string built, no bytes re-hashed.  (An unordered_map<std::string> did both
// per candidate: 270 ns per token, which at 8 GPUs made the tokenizer the slowest stage of the text pipeline.)
constexpr uint64_t kHashBase = 0x9E3779B97F4A7C15ull | 1ull;
struct PieceTable {
    struct Entry { uint64_t hash; uint32_t off, len; int32_t id; };
    std::vector<Entry> slots;   // id < 0: empty
    std::string arena;
    uint64_t mask = 0;
    static uint64_t hash_bytes(const char* p, size_t n) {
        uint64_t h = 0;
        for (size_t i = 0; i < n; ++i) h = h * kHashBase + (static_cast<unsigned char>(p[i]) + 1u);
        return h;
    }
    static uint64_t mix(uint64_t h) {
        h ^= h >> 32;
        h *= 0xD6E8FEB86659FD93ull;
        return h ^ (h >> 29);
    }
    void reserve(size_t n) {
        size_t cap = 16;
        while (cap < 2 * n + 2) cap <<= 1;
        slots.assign(cap, Entry{0, 0, 0, -1});
        mask = cap - 1;
    }
    void insert(const char* p, size_t n, int32_t id) {  // a repeated token keeps its first id
        const uint64_t h = hash_bytes(p, n);
        for (uint64_t i = mix(h) & mask;; i = (i + 1) & mask) {
            Entry& e = slots[i];
            if (e.id < 0) {
                e = Entry{h, static_cast<uint32_t>(arena.size()), static_cast<uint32_t>(n), id};
                arena.append(p, n);
                return;
            }
            if (e.hash == h && e.len == n && memcmp(arena.data() + e.off, p, n) == 0) return;
        }
    }
    int32_t find(uint64_t h, const char* p, size_t n) const {
        for (uint64_t i = mix(h) & mask;; i = (i + 1) & mask) {
            const Entry& e = slots[i];
            if (e.id < 0) return -1;
            if (e.hash == h && e.len == n && memcmp(arena.data() + e.off, p, n) == 0) return e.id;
        }
    }
    int32_t find(const char* p, size_t n) const { return find(hash_bytes(p, n), p, n); }
};

}  // namespace
}  // namespace lsc

struct lsc_tokenizer {
    lsc::PieceTable first;  // every token: pieces at the start of a word
    lsc::PieceTable cont;   // the "##..." tokens without their prefix: pieces inside a word
    std::vector<uint64_t> pw;  // kHashBase^i
    int32_t n_vocab = 0;
    bool lowercase = true;
    int max_length = 512;
    int n_threads = 1;
    int32_t unk = -1, cls = -1, sep = -1;
};

namespace lsc {
namespace {

constexpr size_t kMaxCharsPerWord = 100;  // WordPiece max_input_chars_per_word

struct Scratch {
    std::vector<uint32_t> a, b;
    std::vector<uint32_t> byte_off;
    std::vector<uint64_t> ph;  // prefix hashes of the word's bytes
    std::string word;
};

// WordPiece on one pre-token (code points [w0, w1) of cps); appends ids to out
void wordpiece_bytes(const lsc_tokenizer* tk, size_t nchar, Scratch& sc, std::vector<int32_t>& out);

void wordpiece(const lsc_tokenizer* tk, const std::vector<uint32_t>& cps, size_t w0, size_t w1, Scratch& sc,
               std::vector<int32_t>& out) {
    const size_t nchar = w1 - w0;
    if (nchar > kMaxCharsPerWord) {
        out.push_back(tk->unk);
        return;
    }
    // UTF-8 of the word and the byte offset of every character
    sc.word.clear();
    sc.byte_off.resize(nchar + 1);
    for (size_t i = w0; i < w1; ++i) {
        sc.byte_off[i - w0] = static_cast<uint32_t>(sc.word.size());
        if (cps[i] < 0x80) sc.word.push_back(static_cast<char>(cps[i])); else utf8_append(cps[i], sc.word);
    }
    sc.byte_off[nchar] = static_cast<uint32_t>(sc.word.size());
    wordpiece_bytes(tk, nchar, sc, out);
}

// the greedy longest match proper: the word is sc.word (UTF-8), character c starts at byte sc.byte_off[c]
void wordpiece_bytes(const lsc_tokenizer* tk, size_t nchar, Scratch& sc, std::vector<int32_t>& out) {
    const size_t nbytes = sc.word.size();
    sc.ph.resize(nbytes + 1);
    sc.ph[0] = 0;
    for (size_t i = 0; i < nbytes; ++i) sc.ph[i + 1] = sc.ph[i] * kHashBase + (static_cast<unsigned char>(sc.word[i]) + 1u);
    const uint64_t* pw = tk->pw.data();
    const size_t first = out.size();
    size_t start = 0;
    while (start < nchar) {
        size_t end = nchar;
        int32_t id = -1;
        const PieceTable& tab = start > 0 ? tk->cont : tk->first;  // inside a word the piece is looked up as "##" + piece
        const uint32_t b0 = sc.byte_off[start];
        while (start < end) {
            const uint32_t b1 = sc.byte_off[end];
            id = tab.find(sc.ph[b1] - sc.ph[b0] * pw[b1 - b0], sc.word.data() + b0, b1 - b0);
            if (id >= 0) break;
            --end;
        }
        if (id < 0) {  // the whole word becomes [UNK]
            out.resize(first);
            out.push_back(tk->unk);
            return;
        }
        out.push_back(id);
        start = end;
    }
}

void encode_one(const lsc_tokenizer* tk, const char* text, size_t nbytes, Scratch& sc, std::vector<int32_t>& out) {
    out.clear();
    out.push_back(tk->cls);
    std::vector<uint32_t>& a = sc.a;
    std::vector<uint32_t>& b = sc.b;
    bool ascii_text = true;
    for (size_t i = 0; i < nbytes; ++i)
        if (static_cast<unsigned char>(text[i]) >= 0x80) {
            ascii_text = false;
            break;
        }
    if (ascii_text) {
        // plain ASCII (the common case): clean_text, lower-casing, the pre-tokenizer and WordPiece straight over the bytes -
        // no code point vectors, no accents, no CJK.  Classes: blank (incl. tab / LF / CR), punctuation, removed (the other
        // control characters: they vanish without splitting a word), everything else.
        const bool lower = tk->lowercase;
        const size_t limit = static_cast<size_t>(tk->max_length > 2 ? tk->max_length - 1 : 1);  // [CLS] + body
        const unsigned char* p = reinterpret_cast<const unsigned char*>(text);
        size_t i = 0;
        while (i < nbytes && out.size() < limit) {
            const unsigned char c = p[i];
            const uint8_t k = kAscii.t[c];
            if (k & 5) {  // blank or removed
                ++i;
                continue;
            }
            sc.word.clear();
            if (k & 2) {
                sc.word.push_back(static_cast<char>(c));
                ++i;
            } else {
                while (i < nbytes) {
                    const unsigned char d = p[i];
                    const uint8_t kd = kAscii.t[d];
                    if (kd & 3) break;
                    if (!(kd & 4)) sc.word.push_back(static_cast<char>((lower && d >= 'A' && d <= 'Z') ? d + 32 : d));
                    ++i;
                }
            }
            const size_t nchar = sc.word.size();
            if (nchar > kMaxCharsPerWord) {
                out.push_back(tk->unk);
                continue;
            }
            sc.byte_off.resize(nchar + 1);
            for (size_t c2 = 0; c2 <= nchar; ++c2) sc.byte_off[c2] = static_cast<uint32_t>(c2);
            wordpiece_bytes(tk, nchar, sc, out);
        }
        if (out.size() > limit) out.resize(limit);
        out.push_back(tk->sep);
        return;
    }
    b.clear();
    b.reserve(nbytes + 16);
    utf8_decode(text, nbytes, a);
    // clean_text + handle_chinese_chars
    for (uint32_t c : a) {
        if (bert_is_removed(c)) continue;  // NUL, U+FFFD, control / format / private-use characters
        if (is_white_space(c)) {
            b.push_back(' ');
        } else if (is_chinese_char(c)) {
            b.push_back(' ');
            b.push_back(c);
            b.push_back(' ');
        } else {
            b.push_back(c);
        }
    }
    if (tk->lowercase) {
        // strip_accents (unset => follows lowercase): NFD, drop nonspacing marks; then lowercase
        bool ascii = true;
        for (uint32_t c : b)
            if (c >= 0x80) {
                ascii = false;
                break;
            }
        if (ascii) {
            for (uint32_t& c : b)
                if (c >= 'A' && c <= 'Z') c += 32;
        } else {
            nfd(b, a);
            b.clear();
            for (uint32_t c : a) {
                if (c >= 0x300 && (uni_flags(c) & F_MN)) continue;
                lowercase_append(c, b);
            }
        }
    }
    // BertPreTokenizer: whitespace splits (removed), punctuation isolated; WordPiece per pre-token
    const size_t limit = static_cast<size_t>(tk->max_length > 2 ? tk->max_length - 1 : 1);  // [CLS] + body
    size_t i = 0;
    const size_t n = b.size();
    while (i < n && out.size() < limit) {
        const uint32_t c = b[i];
        if (is_white_space(c)) {
            ++i;
            continue;
        }
        if (bert_is_punctuation(c)) {
            wordpiece(tk, b, i, i + 1, sc, out);
            ++i;
            continue;
        }
        size_t j = i + 1;
        while (j < n && (b[j] < 0x80 ? kAscii.t[b[j]] == 0 : (!is_white_space(b[j]) && !bert_is_punctuation(b[j])))) ++j;
        wordpiece(tk, b, i, j, sc, out);
        i = j;
    }
    if (out.size() > limit) out.resize(limit);  // truncation keeps the first max_length - 2 tokens
    out.push_back(tk->sep);
}

}  // namespace
}  // namespace lsc

extern "C" int lsc_tokenizer_create(const char* vocab_txt, size_t nbytes, int do_lower_case, int32_t max_length,
                                    int32_t n_threads, lsc_tokenizer** out) {
    if (out == nullptr) return fail(LSC_E_INVALID, "lsc_tokenizer_create: out is null");
    *out = nullptr;
    if (vocab_txt == nullptr || nbytes == 0) return fail(LSC_E_INVALID, "lsc_tokenizer_create: empty vocabulary");
    if (max_length < 3) return fail(LSC_E_INVALID, "lsc_tokenizer_create: max_length must be at least 3");
    lsc_tokenizer* tk = new (std::nothrow) lsc_tokenizer();
    if (tk == nullptr) return fail(LSC_E_NOMEM, "lsc_tokenizer_create: out of memory");
    tk->lowercase = do_lower_case != 0;
    tk->max_length = max_length;
    tk->n_threads = n_threads > 0 ? n_threads : static_cast<int>(std::max(1u, std::thread::hardware_concurrency()));
    // vocab.txt: one token per line, id = line number
    int32_t id = 0;
    size_t pos = 0;
    while (pos < nbytes) {
        size_t eol = pos;
        while (eol < nbytes && vocab_txt[eol] != '\n') ++eol;
        size_t end = eol;
        if (end > pos && vocab_txt[end - 1] == '\r') --end;
        ++id;
        pos = eol + 1;
    }
    tk->n_vocab = id;
    tk->first.reserve(static_cast<size_t>(id));
    tk->cont.reserve(static_cast<size_t>(id));
    id = 0;
    pos = 0;
    while (pos < nbytes) {
        size_t eol = pos;
        while (eol < nbytes && vocab_txt[eol] != '\n') ++eol;
        size_t end = eol;
        if (end > pos && vocab_txt[end - 1] == '\r') --end;
        const char* t = vocab_txt + pos;
        const size_t n = end - pos;
        tk->first.insert(t, n, id);  // a repeated token keeps its first id
        if (n > 2 && t[0] == '#' && t[1] == '#') tk->cont.insert(t + 2, n - 2, id);
        ++id;
        pos = eol + 1;
    }
    tk->pw.resize(4 * kMaxCharsPerWord + 2);
    tk->pw[0] = 1;
    for (size_t i = 1; i < tk->pw.size(); ++i) tk->pw[i] = tk->pw[i - 1] * kHashBase;
    auto find = [&](const char* t) { return tk->first.find(t, strlen(t)); };
    tk->unk = find("[UNK]");
    tk->cls = find("[CLS]");
    tk->sep = find("[SEP]");
    if (tk->unk < 0 || tk->cls < 0 || tk->sep < 0) {
        delete tk;
        return fail(LSC_E_INVALID, "lsc_tokenizer_create: the vocabulary lacks [UNK], [CLS] or [SEP]");
    }
    *out = tk;
    return LSC_OK;
}

extern "C" int lsc_tokenizer_set_special(lsc_tokenizer* tk, const char* cls_token, const char* sep_token, const char* unk_token) {
    if (tk == nullptr) return fail(LSC_E_INVALID, "lsc_tokenizer_set_special: tokenizer is null");
    int32_t ids[3] = {tk->cls, tk->sep, tk->unk};
    const char* names[3] = {cls_token, sep_token, unk_token};
    for (int i = 0; i < 3; ++i) {
        if (names[i] == nullptr) continue;
        const int32_t id = tk->first.find(names[i], strlen(names[i]));
        if (id < 0)
            return fail(LSC_E_INVALID, std::string("lsc_tokenizer_set_special: the vocabulary has no token ") + names[i]);
        ids[i] = id;
    }
    tk->cls = ids[0];
    tk->sep = ids[1];
    tk->unk = ids[2];
    return LSC_OK;
}

extern "C" void lsc_tokenizer_destroy(lsc_tokenizer* tk) { delete tk; }

extern "C" int32_t lsc_tokenizer_vocab_size(const lsc_tokenizer* tk) {
    return tk == nullptr ? 0 : tk->n_vocab;  // ids run from 0 to this - 1 (one per line of vocab.txt)
}

extern "C" int lsc_tokenizer_encode(lsc_tokenizer* tk, const char* const* texts, const int32_t* text_nbytes,
                                    int32_t n_texts, int32_t* token_ids, int64_t capacity, int32_t* cu_seqlens) {
    if (tk == nullptr) return fail(LSC_E_INVALID, "lsc_tokenizer_encode: tokenizer is null");
    if (n_texts < 0 || cu_seqlens == nullptr || (n_texts > 0 && (texts == nullptr || token_ids == nullptr)))
        return fail(LSC_E_INVALID, "lsc_tokenizer_encode: bad argument");
    cu_seqlens[0] = 0;
    if (n_texts == 0) return LSC_OK;
    // Workers take chunks of 32 texts from a shared counter (texts differ in length by orders of magnitude: fixed ranges
    // would leave the worker that drew the long ones alone at the end) and append the ids to ONE buffer of their own,
    // noting where every chunk went; the chunks are then copied to their packed positions.  (A vector per text, grown by
    // push_back, made the workers queue up inside malloc: 4 threads gave 1.7x.)
    constexpr int kChunk = 32;
    const int n_chunks = (n_texts + kChunk - 1) / kChunk;
    const int workers = std::max(1, std::min(tk->n_threads, n_chunks));
    const size_t stride = static_cast<size_t>(tk->max_length);
    std::vector<std::vector<int32_t>> bufs(static_cast<size_t>(workers));
    struct ChunkAt { int worker; size_t off; };
    std::vector<ChunkAt> where(static_cast<size_t>(n_chunks));
    std::atomic<int> next{0};
    auto work = [&](int w) {
        Scratch sc;
        std::vector<int32_t> one;
        one.reserve(stride + 8);
        std::vector<int32_t>& buf = bufs[static_cast<size_t>(w)];
        // an even share of the worst case, at most 16M ids up front (the vector grows geometrically beyond that)
        buf.reserve(std::min<size_t>((static_cast<size_t>(n_chunks) / static_cast<size_t>(workers) + 2) * kChunk * stride, size_t{1} << 24));
        size_t at = 0;
        for (int c = next.fetch_add(1); c < n_chunks; c = next.fetch_add(1)) {
            const int lo = c * kChunk, hi = std::min(n_texts, lo + kChunk);
            if (buf.size() < at + static_cast<size_t>(hi - lo) * stride) buf.resize(at + static_cast<size_t>(hi - lo) * stride);
            where[static_cast<size_t>(c)] = ChunkAt{w, at};
            for (int i = lo; i < hi; ++i) {
                const char* t = texts[i];
                const size_t nb = t == nullptr ? 0 : (text_nbytes != nullptr ? static_cast<size_t>(text_nbytes[i]) : strlen(t));
                encode_one(tk, t == nullptr ? "" : t, nb, sc, one);
                memcpy(buf.data() + at, one.data(), one.size() * sizeof(int32_t));
                at += one.size();
                cu_seqlens[i + 1] = static_cast<int32_t>(one.size());  // the length for now; the prefix sum follows
            }
        }
    };
    auto run = [&](auto&& fn) {
        if (workers == 1) {
            fn(0);
            return;
        }
        std::vector<std::thread> pool;
        pool.reserve(static_cast<size_t>(workers) - 1);
        for (int w = 1; w < workers; ++w) pool.emplace_back(fn, w);
        fn(0);
        for (auto& th : pool) th.join();
    };
    run(work);
    int64_t total = 0;
    for (int i = 0; i < n_texts; ++i) total += cu_seqlens[i + 1];
    if (total > capacity) {
        for (int i = 0; i < n_texts; ++i) cu_seqlens[i + 1] = 0;
        return fail(LSC_E_INVALID, "lsc_tokenizer_encode: token buffer too small (n_texts * max_length always suffices)");
    }
    for (int i = 0; i < n_texts; ++i) cu_seqlens[i + 1] += cu_seqlens[i];
    next.store(0);
    run([&](int) {
        for (int c = next.fetch_add(1); c < n_chunks; c = next.fetch_add(1)) {
            const int lo = c * kChunk, hi = std::min(n_texts, lo + kChunk);
            const ChunkAt& ca = where[static_cast<size_t>(c)];
            memcpy(token_ids + cu_seqlens[lo], bufs[static_cast<size_t>(ca.worker)].data() + ca.off,
                   static_cast<size_t>(cu_seqlens[hi] - cu_seqlens[lo]) * sizeof(int32_t));
        }
    });
    return LSC_OK;
}
