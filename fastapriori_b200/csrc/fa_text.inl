// Host-side text loader: D.dat / U.dat bytes -> CSR over a string vocabulary.
// Restates Utils.scala:21,23: Hadoop TextInputFormat records ('\n', "\r\n" or '\r' end a record, a trailing
// terminator adds none), Java `trim` (strip chars <= U+0020 at both ends), `split("\\s+")` with Java's \s
// = [ \t\n\x0B\f\r]; a blank line is the one-token line [""].  Dense ids are handed out in first-seen order.
// No CUDA here: it is the piece of the host boundary that is too slow in an interpreter at 10^7..10^9 tokens.

#include <chrono>
#include <string>
#include <unordered_map>
#ifdef _OPENMP
#include <omp.h>
#endif

struct fa_text {
    // per chunk: row lengths, chunk-local token ids and the local -> global id map; fa_text_copy writes the merged CSR
    // straight into the caller's arrays (in parallel), so the tokens are never gathered into an intermediate copy
    std::vector<std::vector<int64_t>> rl;
    std::vector<std::vector<int32_t>> ids;
    std::vector<std::vector<int32_t>> remap;
    std::vector<int64_t> row0{0}, id0{0};
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

// tokenises buf[s0, s1) (s0 at a record start) into local ids over a local first-seen vocabulary
// Tokens of up to 8 bytes (every numeric item id) are looked up in an open-addressing table keyed by their bytes
// packed into one 64-bit word plus the length: one multiply, one probe, no string compare, no node allocation.
// Longer tokens take the std::unordered_map path.
namespace {
struct ShortTable {
    struct Slot { uint64_t key; int32_t len; int32_t id; };
    std::vector<Slot> slots;
    uint64_t mask;
    size_t used = 0;
    ShortTable() : slots(1 << 13, Slot{0, -1, -1}), mask((1 << 13) - 1) {}
    static uint64_t mix(uint64_t k, int len) { return (k ^ ((uint64_t)len << 59)) * 0x9E3779B97F4A7C15ull; }
    void grow() {
        std::vector<Slot> old;
        old.swap(slots);
        slots.assign(old.size() * 2, Slot{0, -1, -1});
        mask = slots.size() - 1;
        for (const Slot& e : old)
            if (e.len >= 0) {
                uint64_t h = (mix(e.key, e.len) >> 20) & mask;
                while (slots[h].len >= 0) h = (h + 1) & mask;
                slots[h] = e;
            }
    }
    // returns the slot of (key, len); slot.len < 0 means "not present yet"
    Slot& find(uint64_t key, int len) {
        uint64_t h = (mix(key, len) >> 20) & mask;
        while (slots[h].len >= 0 && !(slots[h].key == key && slots[h].len == len)) h = (h + 1) & mask;
        return slots[h];
    }
};
}  // namespace

// tokenises buf[s0, s1) (s0 at a record start) into local ids over a local first-seen vocabulary
static void parse_chunk(const char* buf, int64_t len_total, int64_t s0, int64_t s1, std::vector<int64_t>& row_len, std::vector<int32_t>& ids,
                        std::vector<std::pair<const char*, size_t>>& vocab) {
    std::unordered_map<std::pair<const char*, size_t>, int32_t, SvHash, SvEq> index;  // tokens longer than 8 bytes
    ShortTable table;
    ids.reserve((size_t)((s1 - s0) / 3 + 16));
    row_len.reserve((size_t)((s1 - s0) / 64 + 16));
    auto emit = [&](const char* p, size_t n) {
        int32_t id;
        if (n <= 8) {
            uint64_t key = 0;
            if (p + 8 <= buf + len_total) { memcpy(&key, p, 8); if (n < 8) key &= (1ull << (8 * n)) - 1; }  // one unaligned load
            else memcpy(&key, p, n);
            ShortTable::Slot& sl = table.find(key, (int)n);
            if (sl.len < 0) {
                id = (int32_t)vocab.size();
                vocab.emplace_back(p, n);
                sl = ShortTable::Slot{key, (int32_t)n, id};
                if (++table.used * 2 > table.slots.size()) table.grow();
            } else id = sl.id;
        } else {
            auto key = std::make_pair(p, n);
            auto it = index.find(key);
            if (it == index.end()) { id = (int32_t)vocab.size(); vocab.push_back(key); index.emplace(key, id); }
            else id = it->second;
        }
        ids.push_back(id);
    };
    int64_t i = s0;
    while (i < s1) {
        int64_t s = i;
        while (i < s1 && buf[i] != '\n' && buf[i] != '\r') ++i;
        int64_t e = i;
        if (i < s1) { if (buf[i] == '\r' && i + 1 < s1 && buf[i + 1] == '\n') i += 2; else ++i; }
        while (s < e && (unsigned char)buf[s] <= 0x20) ++s;
        while (e > s && (unsigned char)buf[e - 1] <= 0x20) --e;
        const size_t before = ids.size();
        if (s == e) emit(buf + s, 0);  // blank line: the single token ""
        int64_t p = s;
        while (p < e) {
            int64_t q = p;
            while (q < e && !java_ws(buf[q])) ++q;
            emit(buf + p, (size_t)(q - p));
            while (q < e && java_ws(buf[q])) ++q;
            p = q;
        }
        row_len.push_back((int64_t)(ids.size() - before));
    }
}

extern "C" {

int fa_text_parse(const char* buf, int64_t len, fa_text** out) {
    if (!out || len < 0 || (len > 0 && !buf)) return FA_EINVAL;
    fa_text* t = new (std::nothrow) fa_text();
    if (!t) return FA_ENOMEM;
    // split at record boundaries; every chunk is tokenised independently with its own first-seen vocabulary, then the
    // vocabularies are merged in chunk order, which reproduces the sequential first-seen numbering
    int n_chunks = 1;
#ifdef _OPENMP
    n_chunks = (int)std::max<int64_t>(1, std::min<int64_t>(omp_get_max_threads(), len / (1 << 20)));
#endif
    std::vector<int64_t> cut(n_chunks + 1, len);
    cut[0] = 0;
    for (int c = 1; c < n_chunks; ++c) {
        int64_t p = std::max(cut[c - 1], len * c / n_chunks);
        while (p < len && buf[p] != '\n' && buf[p] != '\r') ++p;     // end of the record that straddles the cut
        if (p < len) { if (buf[p] == '\r' && p + 1 < len && buf[p + 1] == '\n') p += 2; else ++p; }
        cut[c] = p;
    }
    std::vector<std::vector<int64_t>>& rl = t->rl;
    std::vector<std::vector<int32_t>>& ids = t->ids;
    rl.resize(n_chunks);
    ids.resize(n_chunks);
    std::vector<std::vector<std::pair<const char*, size_t>>> voc(n_chunks);
    const auto dbg_t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(static, 1)
    for (int c = 0; c < n_chunks; ++c) parse_chunk(buf, len, cut[c], cut[c + 1], rl[c], ids[c], voc[c]);
    const auto dbg_t1 = std::chrono::steady_clock::now();
    std::unordered_map<std::pair<const char*, size_t>, int32_t, SvHash, SvEq> index;
    std::vector<std::vector<int32_t>>& remap = t->remap;
    remap.resize(n_chunks);
    for (int c = 0; c < n_chunks; ++c) {
        remap[c].resize(voc[c].size());
        for (size_t v = 0; v < voc[c].size(); ++v) {
            auto it = index.find(voc[c][v]);
            if (it == index.end()) {
                const int32_t id = (int32_t)t->vocab.size();
                t->vocab.emplace_back(voc[c][v].first, voc[c][v].second);
                t->vocab_bytes += (int64_t)voc[c][v].second + 1;
                index.emplace(voc[c][v], id);  // keyed by the bytes in `buf`, which outlive this loop
                remap[c][v] = id;
            } else remap[c][v] = it->second;
        }
    }
    t->row0.assign(n_chunks + 1, 0);
    t->id0.assign(n_chunks + 1, 0);
    for (int c = 0; c < n_chunks; ++c) { t->row0[c + 1] = t->row0[c] + (int64_t)rl[c].size(); t->id0[c + 1] = t->id0[c] + (int64_t)ids[c].size(); }
    if (getenv("FA_DEBUG_PARSE"))
        fprintf(stderr, "[parse] chunks %d: tokenise %.3f s, merge %.3f s\n", n_chunks, std::chrono::duration<double>(dbg_t1 - dbg_t0).count(),
                std::chrono::duration<double>(std::chrono::steady_clock::now() - dbg_t1).count());
    *out = t;
    return FA_OK;
}

int64_t fa_text_rows(const fa_text* t) { return t ? t->row0.back() : 0; }
int64_t fa_text_nnz(const fa_text* t) { return t ? t->id0.back() : 0; }
int32_t fa_text_vocab_size(const fa_text* t) { return t ? (int32_t)t->vocab.size() : 0; }
int64_t fa_text_vocab_bytes(const fa_text* t) { return t ? t->vocab_bytes : 0; }

int fa_text_copy(const fa_text* t, int64_t* row_offsets_out, int32_t* ids_out, char* vocab_out) {
    if (!t) return FA_EINVAL;
    const int n_chunks = (int)t->rl.size();
    if (row_offsets_out) row_offsets_out[0] = 0;
#pragma omp parallel for schedule(static, 1)
    for (int c = 0; c < n_chunks; ++c) {
        if (row_offsets_out) {
            int64_t o = t->id0[c];
            for (size_t r = 0; r < t->rl[c].size(); ++r) { o += t->rl[c][r]; row_offsets_out[(size_t)t->row0[c] + r + 1] = o; }
        }
        if (ids_out) {
            const int32_t* map = t->remap[c].data();
            int32_t* dst = ids_out + t->id0[c];
            for (size_t i = 0; i < t->ids[c].size(); ++i) dst[i] = map[(size_t)t->ids[c][i]];
        }
    }
    if (vocab_out) {  // tokens never contain whitespace, so '\n' separates them unambiguously ("" stays "")
        char* w = vocab_out;
        for (const std::string& s : t->vocab) { memcpy(w, s.data(), s.size()); w += s.size(); *w++ = '\n'; }
    }
    return FA_OK;
}

void fa_text_free(fa_text* t) { delete t; }

}  // extern "C"

// ---- output formatter (Utils.scala:38-39 / :60-61): "items by rank descending joined by a space" per itemset,
// optional "[count]", lines sorted as byte strings (== Java String order for ASCII items), joined by '\n'.
struct fa_lines {
    std::vector<std::string> items;   // rank -> item string
    std::vector<char> arena;          // formatted lines, each terminated by '\n'
    std::vector<std::pair<int64_t, int32_t>> lines;  // (offset, length without the terminator)
    std::vector<char> out;
};

extern "C" {

int fa_lines_new(const char* items_joined, int64_t bytes, int32_t n_items, fa_lines** out) {
    if (!out || n_items < 0 || bytes < 0) return FA_EINVAL;
    fa_lines* L = new (std::nothrow) fa_lines();
    if (!L) return FA_ENOMEM;
    int64_t p = 0;
    for (int32_t i = 0; i < n_items; ++i) {
        int64_t q = p;
        while (q < bytes && items_joined[q] != '\n') ++q;
        L->items.emplace_back(items_joined + p, (size_t)(q - p));
        p = q + 1;
    }
    *out = L;
    return FA_OK;
}

int fa_lines_add_level(fa_lines* L, const int32_t* sets, const int32_t* counts_or_null, int64_t n, int32_t k) {
    if (!L || n < 0 || k < 1 || (n && !sets)) return FA_EINVAL;
    const int F = (int)L->items.size();
    // size pass, then parallel fill
    std::vector<int64_t> off((size_t)n + 1, 0);
#pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < n; ++s) {
        int64_t len = k - 1;
        for (int j = 0; j < k; ++j) { const int r = sets[s * k + j]; len += (r >= 0 && r < F) ? (int64_t)L->items[r].size() : 0; }
        if (counts_or_null) { char tmp[16]; len += snprintf(tmp, sizeof tmp, "[%d]", counts_or_null[s]); }
        off[(size_t)s + 1] = len + 1;
    }
    for (int64_t s = 0; s < n; ++s) off[(size_t)s + 1] += off[(size_t)s];
    const int64_t base = (int64_t)L->arena.size(), l0 = (int64_t)L->lines.size();
    L->arena.resize((size_t)(base + off[(size_t)n]));
    L->lines.resize((size_t)(l0 + n));
#pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < n; ++s) {
        char* w = L->arena.data() + base + off[(size_t)s];
        char* w0 = w;
        for (int j = k - 1; j >= 0; --j) {  // sortBy(-_): rank descending
            const int r = sets[s * k + j];
            if (r >= 0 && r < F) { memcpy(w, L->items[r].data(), L->items[r].size()); w += L->items[r].size(); }
            if (j) *w++ = ' ';
        }
        if (counts_or_null) w += snprintf(w, 16, "[%d]", counts_or_null[s]);
        L->lines[(size_t)(l0 + s)] = std::make_pair(base + off[(size_t)s], (int32_t)(w - w0));
        *w = '\n';
    }
    return FA_OK;
}

int fa_lines_finish(fa_lines* L, int32_t sort_lines, int64_t* total_bytes_out) {
    if (!L || !total_bytes_out) return FA_EINVAL;
    const char* A = L->arena.data();
    const int64_t n = (int64_t)L->lines.size();
    if (sort_lines && n > 1) {
        // Byte-string order without a serial merge (sample sort): every line gets its first 8 bytes as a big-endian key
        // (zero padded: a proper prefix sorts first, and no line contains a NUL); splitters taken from a sorted sample
        // cut the lines into buckets of about equal size whatever the distribution of the leading items is (real
        // outputs are skewed: every line starts with its rarest item); the buckets - independent ranges of the final
        // order - are then sorted concurrently, comparing keys first and the tails in the arena only on a tie.
        struct Ent { uint64_t key; int64_t off; int32_t len; };
        std::unique_ptr<Ent[]> ent(new Ent[(size_t)n]), tmp(new Ent[(size_t)n]);  // uninitialised: every entry is written below
        auto less = [A](const Ent& a, const Ent& b) {
            if (a.key != b.key) return a.key < b.key;
            const int la = std::max(a.len - 8, 0), lb = std::max(b.len - 8, 0);
            const int c = memcmp(A + a.off + 8, A + b.off + 8, (size_t)std::min(la, lb));  // equal keys: both >= 8 bytes, or equal short lines
            return c < 0 || (c == 0 && a.len < b.len);
        };
        int T = 1;
#ifdef _OPENMP
        T = std::min(omp_get_max_threads(), 32);
#endif
#pragma omp parallel for schedule(static) num_threads(T)
        for (int64_t i = 0; i < n; ++i) {
            const auto& ln = L->lines[(size_t)i];
            uint64_t key = 0;
            const int m = std::min(ln.second, 8);
            for (int b = 0; b < m; ++b) key |= (uint64_t)(unsigned char)A[ln.first + b] << (56 - 8 * b);
            ent[(size_t)i] = Ent{key, ln.first, ln.second};
        }
        const int NB = (int)std::max<int64_t>(1, std::min<int64_t>(T * 8, n / 2048));  // buckets
        std::vector<Ent> split;  // NB - 1 splitters: bucket b holds the lines in [split[b-1], split[b])
        if (NB > 1) {
            const int64_t n_sample = std::min<int64_t>(n, (int64_t)NB * 64);
            std::vector<Ent> sample((size_t)n_sample);
            for (int64_t j = 0; j < n_sample; ++j) sample[(size_t)j] = ent[(size_t)(n * j / n_sample)];
            std::sort(sample.begin(), sample.end(), less);
            for (int b = 1; b < NB; ++b) split.push_back(sample[(size_t)(n_sample * b / NB)]);
        }
        std::vector<int64_t> cnt((size_t)T * NB, 0);
        std::unique_ptr<uint16_t[]> bucket(new uint16_t[(size_t)n]);
#pragma omp parallel num_threads(T)
        {
            int t = 0;
#ifdef _OPENMP
            t = omp_get_thread_num();
#endif
            int64_t* c = cnt.data() + (size_t)t * NB;
            const int64_t lo = n * t / T, hi = n * (t + 1) / T;
            for (int64_t i = lo; i < hi; ++i) {
                const int b = (int)(std::upper_bound(split.begin(), split.end(), ent[(size_t)i], less) - split.begin());
                bucket[(size_t)i] = (uint16_t)b;
                ++c[b];
            }
        }
        std::vector<int64_t> start((size_t)NB + 1, 0);
        {   // bucket-major, thread-minor prefix sums: thread t writes its lines of bucket b from cnt[t][b] on
            int64_t run = 0;
            for (int b = 0; b < NB; ++b) {
                start[(size_t)b] = run;
                for (int t = 0; t < T; ++t) { const int64_t c = cnt[(size_t)t * NB + b]; cnt[(size_t)t * NB + b] = run; run += c; }
            }
            start[(size_t)NB] = run;
        }
#pragma omp parallel num_threads(T)
        {
            int t = 0;
#ifdef _OPENMP
            t = omp_get_thread_num();
#endif
            int64_t* c = cnt.data() + (size_t)t * NB;
            const int64_t lo = n * t / T, hi = n * (t + 1) / T;
            for (int64_t i = lo; i < hi; ++i) tmp[(size_t)c[bucket[(size_t)i]]++] = ent[(size_t)i];
        }
#pragma omp parallel for schedule(dynamic, 1)
        for (int b = 0; b < NB; ++b) std::sort(tmp.get() + start[(size_t)b], tmp.get() + start[(size_t)b + 1], less);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) L->lines[(size_t)i] = std::make_pair(tmp[(size_t)i].off, tmp[(size_t)i].len);
    }
    std::vector<int64_t> off((size_t)n + 1, 0);
    for (int64_t i = 0; i < n; ++i) off[(size_t)i + 1] = off[(size_t)i] + L->lines[(size_t)i].second + 1;
    L->out.resize((size_t)off[(size_t)n]);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) memcpy(L->out.data() + off[(size_t)i], A + L->lines[(size_t)i].first, (size_t)L->lines[(size_t)i].second + 1);
    *total_bytes_out = off[(size_t)n];
    return FA_OK;
}

int fa_lines_copy(const fa_lines* L, char* out) {
    if (!L || (!out && !L->out.empty())) return FA_EINVAL;
    if (!L->out.empty()) memcpy(out, L->out.data(), L->out.size());
    return FA_OK;
}

const char* fa_lines_data(const fa_lines* L) { return (L && !L->out.empty()) ? L->out.data() : ""; }

void fa_lines_free(fa_lines* L) { delete L; }

}  // extern "C"
