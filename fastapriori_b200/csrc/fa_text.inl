// Host-side text loader: D.dat / U.dat bytes -> CSR over a string vocabulary.
// Restates Utils.scala:21,23: Hadoop TextInputFormat records ('\n', "\r\n" or '\r' end a record, a trailing
// terminator adds none), Java `trim` (strip chars <= U+0020 at both ends), `split("\\s+")` with Java's \s
// = [ \t\n\x0B\f\r]; a blank line is the one-token line [""].  Dense ids are handed out in first-seen order.
// No CUDA here: it is the piece of the host boundary that is too slow in an interpreter at 10^7..10^9 tokens.

#include <string>
#include <unordered_map>

struct fa_text {
    std::vector<int64_t> offsets{0};
    std::vector<int32_t> ids;
    std::vector<std::string> vocab;
    int64_t vocab_bytes = 0;
};

namespace {
struct SvHash {
    size_t operator()(const std::pair<const char*, size_t>& k) const {
        uint64_t h = 0xcbf29ce484222325ull;
        for (size_t i = 0; i < k.second; ++i) { h ^= (unsigned char)k.first[i]; h *= 0x100000001b3ull; }
        return (size_t)h;
    }
};
struct SvEq {
    bool operator()(const std::pair<const char*, size_t>& a, const std::pair<const char*, size_t>& b) const {
        return a.second == b.second && memcmp(a.first, b.first, a.second) == 0;
    }
};
inline bool java_ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == 0x0B || c == '\f' || c == '\r'; }
}  // namespace

extern "C" {

int fa_text_parse(const char* buf, int64_t len, fa_text** out) {
    if (!out || len < 0 || (len > 0 && !buf)) return FA_EINVAL;
    fa_text* t = new (std::nothrow) fa_text();
    if (!t) return FA_ENOMEM;
    // keys point into `buf` (which outlives the call); the vocab keeps its own copies
    std::unordered_map<std::pair<const char*, size_t>, int32_t, SvHash, SvEq> index;
    index.reserve(1 << 16);
    auto emit = [&](const char* p, size_t n) {
        auto key = std::make_pair(p, n);
        auto it = index.find(key);
        int32_t id;
        if (it == index.end()) {
            id = (int32_t)t->vocab.size();
            t->vocab.emplace_back(p, n);
            t->vocab_bytes += (int64_t)n + 1;
            // re-key on the vocab's own storage? not needed: `buf` is valid for the whole call
            index.emplace(key, id);
        } else id = it->second;
        t->ids.push_back(id);
    };
    int64_t i = 0;
    while (i < len) {
        int64_t s = i;
        while (i < len && buf[i] != '\n' && buf[i] != '\r') ++i;
        int64_t e = i;
        if (i < len) { if (buf[i] == '\r' && i + 1 < len && buf[i + 1] == '\n') i += 2; else ++i; }
        while (s < e && (unsigned char)buf[s] <= 0x20) ++s;
        while (e > s && (unsigned char)buf[e - 1] <= 0x20) --e;
        if (s == e) emit(buf + s, 0);  // blank line: the single token ""
        int64_t p = s;
        while (p < e) {
            int64_t q = p;
            while (q < e && !java_ws(buf[q])) ++q;
            emit(buf + p, (size_t)(q - p));
            while (q < e && java_ws(buf[q])) ++q;
            p = q;
        }
        t->offsets.push_back((int64_t)t->ids.size());
    }
    *out = t;
    return FA_OK;
}

int64_t fa_text_rows(const fa_text* t) { return t ? (int64_t)t->offsets.size() - 1 : 0; }
int64_t fa_text_nnz(const fa_text* t) { return t ? (int64_t)t->ids.size() : 0; }
int32_t fa_text_vocab_size(const fa_text* t) { return t ? (int32_t)t->vocab.size() : 0; }
int64_t fa_text_vocab_bytes(const fa_text* t) { return t ? t->vocab_bytes : 0; }

int fa_text_copy(const fa_text* t, int64_t* row_offsets_out, int32_t* ids_out, char* vocab_out) {
    if (!t) return FA_EINVAL;
    if (row_offsets_out) memcpy(row_offsets_out, t->offsets.data(), sizeof(int64_t) * t->offsets.size());
    if (ids_out && !t->ids.empty()) memcpy(ids_out, t->ids.data(), sizeof(int32_t) * t->ids.size());
    if (vocab_out) {  // tokens never contain whitespace, so '\n' separates them unambiguously ("" stays "")
        char* w = vocab_out;
        for (const std::string& s : t->vocab) { memcpy(w, s.data(), s.size()); w += s.size(); *w++ = '\n'; }
    }
    return FA_OK;
}

void fa_text_free(fa_text* t) { delete t; }

}  // extern "C"
