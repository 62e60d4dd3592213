// Seeded IBM-Quest-style (Agrawal & Srikant, VLDB'94 §4.1) synthetic transaction generator.
//
// This is workload tooling for tests and bench.py, not part of the mining hot path. The
// reference (relife957/FastApriori) ships no data at all; its input contract
// is "one transaction per line, items separated by whitespace" (Utils.scala:21).
//
// Determinism and sharding: the pattern table depends only on (seed, params). Rows are
// produced in blocks of QUEST_BLOCK rows; block b draws from its own RNG stream
// splitmix(seed, b), so rows [r0, r1) of the virtual dataset are the same no matter which
// rank or thread generates them.  Item ids are 1..n_items ("0" is the reference's
// "no recommendation" sentinel, AssociationRules.scala:49,105). Items are distinct within a
// row, so level-1 occurrence counts equal presence counts.
//
// No std:: random distributions are used (they are implementation-defined); everything is
// derived from a splitmix64/xoshiro256** stream with explicit transforms.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr int64_t QUEST_BLOCK = 16384;

struct Rng {
    uint64_t s[4];
    static uint64_t splitmix(uint64_t& x) {
        uint64_t z = (x += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    Rng(uint64_t seed, uint64_t stream) {
        uint64_t x = seed * 0xD1342543DE82EF95ull + stream * 0x2545F4914F6CDD1Dull + 0x1234567ull;
        for (int i = 0; i < 4; ++i) s[i] = splitmix(x);
    }
    static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next() {
        uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }  // [0,1)
    uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
    double expo() { return -std::log(1.0 - uniform()); }
    double normal() {  // Box-Muller, one value
        double u1 = 1.0 - uniform(), u2 = uniform();
        return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
    }
    int poisson(double mean) {  // Knuth for small means, normal approx above 64
        if (mean > 64.0) {
            int v = (int)std::floor(mean + std::sqrt(mean) * normal() + 0.5);
            return v < 0 ? 0 : v;
        }
        double l = std::exp(-mean), p = 1.0;
        int k = 0;
        do { ++k; p *= uniform(); } while (p > l);
        return k - 1;
    }
};

struct Params {
    int64_t n_rows;     // D
    double avg_len;     // T
    double avg_pat;     // I
    int32_t n_pat;      // L
    int32_t n_items;    // N
    double corr;        // fraction of a pattern inherited from its predecessor (mean)
    uint64_t seed;
};

struct Patterns {
    std::vector<int32_t> off, items;
    std::vector<double> cum;      // cumulative weights for pattern picking
    std::vector<float> corrupt;   // per-pattern corruption level
};

static void build_patterns(const Params& p, Patterns& P) {
    Rng rng(p.seed, 0xA11CE);
    P.off.assign(1, 0);
    P.cum.resize(p.n_pat);
    P.corrupt.resize(p.n_pat);
    std::vector<int32_t> prev, cur;
    std::vector<char> used(p.n_items + 1, 0);
    double total = 0;
    for (int l = 0; l < p.n_pat; ++l) {
        int sz = rng.poisson(p.avg_pat - 1.0) + 1;
        if (sz > p.n_items) sz = p.n_items;
        cur.clear();
        if (!prev.empty()) {
            double frac = p.corr * rng.expo();
            if (frac > 1.0) frac = 1.0;
            int take = (int)(frac * sz + 0.5);
            if (take > (int)prev.size()) take = (int)prev.size();
            // partial Fisher-Yates over a copy of prev
            std::vector<int32_t> pv(prev);
            for (int i = 0; i < take; ++i) {
                int j = i + (int)rng.below((uint32_t)(pv.size() - i));
                std::swap(pv[i], pv[j]);
                cur.push_back(pv[i]);
                used[pv[i]] = 1;
            }
        }
        while ((int)cur.size() < sz) {
            int32_t it = 1 + (int32_t)rng.below((uint32_t)p.n_items);
            if (!used[it]) { used[it] = 1; cur.push_back(it); }
        }
        for (int32_t it : cur) used[it] = 0;
        P.items.insert(P.items.end(), cur.begin(), cur.end());
        P.off.push_back((int32_t)P.items.size());
        double w = rng.expo();
        total += w;
        P.cum[l] = total;
        double c = 0.5 + 0.1 * rng.normal();
        P.corrupt[l] = (float)(c < 0 ? 0 : (c > 1 ? 1 : c));
        prev = cur;
    }
    for (auto& c : P.cum) c /= total;
    P.cum.back() = 1.0;
}

// Generates one block of rows into `out_items` (appended) with per-row lengths in `out_len`.
static void gen_block(const Params& p, const Patterns& P, int64_t block, int64_t r0, int64_t r1,
                      std::vector<int32_t>& out_items, std::vector<int32_t>& out_len) {
    Rng rng(p.seed, 0xB10C0000ull + (uint64_t)block);
    std::vector<int32_t> stamp(p.n_items + 1, -1);
    std::vector<int32_t> row, pend;
    for (int64_t r = r0; r < r1; ++r) {
        int target = rng.poisson(p.avg_len - 1.0) + 1;
        if (target > p.n_items) target = p.n_items;
        row.clear();
        int32_t tag = (int32_t)(r - r0);
        auto add = [&](int32_t it) {
            if (stamp[it] != tag) { stamp[it] = tag; row.push_back(it); }
        };
        for (int32_t it : pend) add(it);
        pend.clear();
        int guard = 0;
        while ((int)row.size() < target && guard++ < 64) {
            double u = rng.uniform();
            int l = (int)(std::lower_bound(P.cum.begin(), P.cum.end(), u) - P.cum.begin());
            if (l >= p.n_pat) l = p.n_pat - 1;
            // corrupt: keep dropping one random item while uniform < c
            int b = P.off[l], e = P.off[l + 1];
            static thread_local std::vector<int32_t> pat;
            pat.assign(P.items.begin() + b, P.items.begin() + e);
            while (!pat.empty() && rng.uniform() < P.corrupt[l]) {
                int j = (int)rng.below((uint32_t)pat.size());
                pat[j] = pat.back();
                pat.pop_back();
            }
            int fresh = 0;
            for (int32_t it : pat) fresh += (stamp[it] != tag);
            if ((int)row.size() + fresh > target && !row.empty()) {
                // does not fit: half the time add anyway, otherwise defer to the next row
                if (rng.uniform() < 0.5) { for (int32_t it : pat) add(it); }
                else { pend = pat; }
                break;
            }
            for (int32_t it : pat) add(it);
        }
        if (row.empty()) add(1 + (int32_t)rng.below((uint32_t)p.n_items));
        out_len.push_back((int32_t)row.size());
        out_items.insert(out_items.end(), row.begin(), row.end());
    }
}

}  // namespace

extern "C" {

// Generates rows [row_begin, row_end) of the virtual dataset described by the parameters.
// Returns malloc'd CSR arrays (free with quest_free): *offsets_out has (row_end-row_begin+1)
// int64 entries, *items_out has offsets[last] int32 item ids in 1..n_items.
int quest_generate(int64_t n_rows, double avg_len, double avg_pat, int32_t n_pat, int32_t n_items,
                   double corr, uint64_t seed, int64_t row_begin, int64_t row_end,
                   int64_t** offsets_out, int32_t** items_out) {
    if (n_rows <= 0 || n_items <= 0 || n_pat <= 0 || row_begin < 0 || row_end > n_rows ||
        row_begin > row_end || avg_len < 1 || avg_pat < 1)
        return 1;
    Params p{n_rows, avg_len, avg_pat, n_pat, n_items, corr, seed};
    Patterns P;
    build_patterns(p, P);
    int64_t b0 = row_begin / QUEST_BLOCK, b1 = (row_end + QUEST_BLOCK - 1) / QUEST_BLOCK;
    int64_t nb = b1 - b0;
    std::vector<std::vector<int32_t>> bit(nb), blen(nb);
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t bi = 0; bi < nb; ++bi) {
        int64_t b = b0 + bi;
        int64_t s = b * QUEST_BLOCK, e = std::min<int64_t>(n_rows, s + QUEST_BLOCK);
        // A block is always generated whole so its rows do not depend on the requested range.
        gen_block(p, P, b, s, e, bit[bi], blen[bi]);
    }
    int64_t nr = row_end - row_begin;
    int64_t* off = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nr + 1));
    if (!off) return 2;
    int64_t nnz = 0, ri = 0;
    off[0] = 0;
    for (int64_t bi = 0; bi < nb; ++bi) {
        int64_t s = (b0 + bi) * QUEST_BLOCK;
        for (size_t j = 0; j < blen[bi].size(); ++j) {
            int64_t r = s + (int64_t)j;
            if (r >= row_begin && r < row_end) { nnz += blen[bi][j]; off[++ri] = nnz; }
        }
    }
    int32_t* items = (int32_t*)malloc(sizeof(int32_t) * (size_t)std::max<int64_t>(nnz, 1));
    if (!items) { free(off); return 2; }
    int64_t w = 0;
    for (int64_t bi = 0; bi < nb; ++bi) {
        int64_t s = (b0 + bi) * QUEST_BLOCK;
        int64_t pos = 0;
        for (size_t j = 0; j < blen[bi].size(); ++j) {
            int64_t r = s + (int64_t)j;
            int32_t len = blen[bi][j];
            if (r >= row_begin && r < row_end) {
                memcpy(items + w, bit[bi].data() + pos, sizeof(int32_t) * (size_t)len);
                w += len;
            }
            pos += len;
        }
    }
    *offsets_out = off;
    *items_out = items;
    return 0;
}

void quest_free(void* p) { free(p); }

// Writes CSR rows as the reference's text format: items separated by one space, '\n' per row.
// Returns 0 on success.
int quest_write_text(const char* path, const int64_t* offsets, const int32_t* items, int64_t n_rows) {
    FILE* f = fopen(path, "wb");
    if (!f) return 1;
    std::vector<char> buf;
    buf.reserve(1 << 22);
    char tmp[16];
    for (int64_t r = 0; r < n_rows; ++r) {
        for (int64_t j = offsets[r]; j < offsets[r + 1]; ++j) {
            int n = snprintf(tmp, sizeof tmp, j == offsets[r] ? "%d" : " %d", items[j]);
            buf.insert(buf.end(), tmp, tmp + n);
        }
        buf.push_back('\n');
        if (buf.size() > (1u << 22) - 4096) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
    }
    fwrite(buf.data(), 1, buf.size(), f);
    return fclose(f) != 0;
}

}  // extern "C"
