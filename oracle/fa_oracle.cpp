// fa_oracle — CPU restatement of relife957/FastApriori's mining + rule/recommendation path.
//
// TEST INFRASTRUCTURE ONLY.  This file is the parity checker and the timed CPU baseline
// ("port").  Nothing under fastapriori_b200/ may include, link, import or execute it; only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
//
// PARITY UNPINNED BY THE REFERENCE: the reference ships no tests, fixtures, sample data or
// golden output (pom.xml:55 names src/test/scala, which does not exist) and cannot run here
// (no JVM/Scala/Spark in the image).  The pins this oracle is held to are ours: the
// hand-derived known-answer vector of SURVEY.md Appendix B (tests/golden/appendix_b/) and an
// independent brute-force enumerator (oracle/brute.py) on random small inputs.
//
// Each function cites the reference lines it restates (paths under
// /root/reference/src/main/scala/com/hazzacheng/AR/).  The data structures are deliberately
// the reference's in spirit: string items, a rank map, weighted de-duplicated rows, one BYTE
// per boolean bitmap entry, a full scan per pair, a set-of-sets candidate prune, linear
// subset scans in rule generation, a stable sort.  OpenMP parallelises the axes Spark
// parallelises (candidate pairs, prefix groups, (k-1)-itemsets, supersets, rules, baskets).
//
// Two switches exist only so that the checker finishes at the sizes BASELINE.json names; both
// are held to the naive mode, output for output, by tests/test_oracle.py:
//   flags bit3 (8)  "packed": the per-item Array[Boolean] becomes one bit per row (uint64
//                   words); a support is still the weighted sum over rows where the AND holds.
//   flags bit4 (16) "no dedup": identical baskets are not merged (every row weighs 1), which
//                   yields the same supports; only totalCount (a counter) differs.
// Lines are interned first (token string -> dense id, CSR of ids): what the reference does on
// strings per line is done here on ids, with the strings kept for ranking ties and output.
//
// The one place Spark itself decides a result — the order of equal level-1 counts after
// `collect()` (FastApriori.scala:55-60) — is replaced by a declared
// canonical rule: count descending, then items that parse as Int ascending by value, then
// remaining items in byte order.  It changes only the intra-line item order of freqItemset.

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using Items = std::vector<int32_t>;  // a Set[Int] kept sorted ascending by rank

struct ItemsHash {
    size_t operator()(const Items& v) const {
        uint64_t h = 0xcbf29ce484222325ull;
        for (int32_t x : v) { h ^= (uint32_t)x; h *= 0x100000001b3ull; }
        return (size_t)h;
    }
};

struct Rule { Items ant; int32_t cons; double conf; };

struct Result {
    int status = 0;            // 0 ok; 1 = reference would throw empty.min (AssociationRules.scala:147)
    int number_format = 0;     // 1 = reference would throw NumberFormatException (AssociationRules.scala:119)
    int64_t n_lines = 0;
    int32_t min_count = 0;
    int32_t total_count = 0;   // distinct weighted rows (FastApriori.scala:79)
    int level1_ties = 0;       // equal counts among frequent items (Appendix C detector)
    std::vector<std::string> freq_items;          // rank -> item string
    std::vector<int32_t> level1_counts;           // rank -> occurrence count
    std::vector<std::vector<std::pair<Items, int32_t>>> levels;  // levels[k] = frequent k-itemsets
    std::vector<int64_t> n_groups, n_cands;       // per level k: candidate groups / candidates
    std::vector<Rule> raw_rules, rules;           // before prune / kept + sorted
    std::vector<int64_t> before_cut, after_cut;   // per antecedent size (println counters :155,:177)
    std::vector<std::string> recommends;          // per user line
    std::string freqitemset_text, recommends_text;
    double t_parse = 0, t_mine = 0, t_rules = 0;
    // finer split: row-dependent (level-1 + dedup, bitmaps, pair pass, k-level counting), row-independent
    // (candidate generation, rule generation + prune + sort) and user-dependent (matching)
    double t_level1 = 0, t_bitmaps = 0, t_pairs = 0, t_cand = 0, t_count = 0, t_rulegen = 0, t_sort = 0, t_match = 0;
    std::string scratch;
};

// FAO_VERBOSE=1: phase timings on stderr (the long runs of tests/golden/make_digests.py)
bool verbose() { static const bool v = getenv("FAO_VERBOSE") != nullptr; return v; }
#define VLOG(...) do { if (verbose()) { fprintf(stderr, "[oracle] " __VA_ARGS__); fputc('\n', stderr); } } while (0)

double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// A file as the reference sees it after `_.trim().split("\\s+")`: lines of tokens.  Tokens are interned
// (vocab[id] is the string), lines are CSR over ids; a duplicate token in a line stays a duplicate.
struct Corpus {
    std::vector<std::string> vocab;
    std::vector<int64_t> off{0};
    std::vector<int32_t> ids;
    std::unordered_map<std::string, int32_t> index;
    int64_t n_lines() const { return (int64_t)off.size() - 1; }
    int32_t intern(const std::string& t) {
        auto it = index.find(t);
        if (it != index.end()) return it->second;
        int32_t id = (int32_t)vocab.size();
        index.emplace(t, id);
        vocab.push_back(t);
        return id;
    }
};

// Utils.scala:21,23 — Hadoop TextInputFormat records ('\n', '\r\n', '\r' terminate; a final
// terminator adds no record), then Java `trim` (strip chars <= U+0020 at both ends) and
// `split("\\s+")` (Java \s = [ \t\n\x0B\f\r]).  A blank line yields the single token "".
void tokenize(const char* buf, int64_t len, Corpus& C) {
    int64_t i = 0;
    std::string tok;
    while (i < len) {
        int64_t s = i;
        while (i < len && buf[i] != '\n' && buf[i] != '\r') ++i;
        int64_t e = i;
        if (i < len) { if (buf[i] == '\r' && i + 1 < len && buf[i + 1] == '\n') i += 2; else ++i; }
        while (s < e && (unsigned char)buf[s] <= 0x20) ++s;
        while (e > s && (unsigned char)buf[e - 1] <= 0x20) --e;
        auto is_ws = [](char c) { return c == ' ' || c == '\t' || c == '\n' || c == 0x0B || c == '\f' || c == '\r'; };
        int64_t p = s, n_tok = 0;
        while (p < e) {
            int64_t q = p;
            while (q < e && !is_ws(buf[q])) ++q;
            tok.assign(buf + p, buf + q);
            C.ids.push_back(C.intern(tok));
            ++n_tok;
            while (q < e && is_ws(buf[q])) ++q;
            p = q;
        }
        if (n_tok == 0) C.ids.push_back(C.intern(std::string()));
        C.off.push_back((int64_t)C.ids.size());
    }
}

// The same corpus from lines that are already integer item ids (the synthetic workloads): the token of id v is
// its decimal string, exactly what tokenize() would have produced from the text form of the same rows.
void corpus_from_csr(const int64_t* off, const int32_t* items, int64_t n_rows, Corpus& C) {
    std::unordered_map<int32_t, int32_t> seen;
    C.off.reserve((size_t)n_rows + 1);
    C.ids.reserve((size_t)(off[n_rows] - off[0]));
    for (int64_t r = 0; r < n_rows; ++r) {
        for (int64_t e = off[r]; e < off[r + 1]; ++e) {
            const int32_t v = items[e];
            auto it = seen.find(v);
            if (it == seen.end()) it = seen.emplace(v, C.intern(std::to_string(v))).first;
            C.ids.push_back(it->second);
        }
        if (off[r + 1] == off[r]) C.ids.push_back(C.intern(std::string()));  // an empty line is the token ""
        C.off.push_back((int64_t)C.ids.size());
    }
}

bool parse_int(const std::string& s, long long& v) {  // Scala String.toInt (base 10, optional sign)
    if (s.empty()) return false;
    size_t i = 0;
    bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; if (s.size() == 1) return false; }
    long long x = 0;
    for (; i < s.size(); ++i) {
        if (s[i] < '0' || s[i] > '9') return false;
        x = x * 10 + (s[i] - '0');
        if (x > 2147483648ll) return false;
    }
    v = neg ? -x : x;
    return v >= -2147483648ll && v <= 2147483647ll;
}

struct Miner {
    double min_support;
    int flags;  // bit0: hashed subset lookups in genRules; bit2: sibling join in genCandidates; bit3: packed bitmaps; bit4: no dedup
    Result& R;
    std::vector<int32_t> id_to_rank;    // itemToRank over interned token ids (-1 = not frequent)
    std::vector<Items> rows;            // newData: distinct filtered baskets (FastApriori.scala:66-78)
    std::vector<int32_t> weights;       // countMap
    std::vector<std::vector<uint8_t>> bitmaps;  // freqItemsTrans: Array[Boolean] per rank
    // packed mode (flags bit3): one bit per (rank, row); rows are ordered by weight descending, so only the words
    // below heavy_words hold rows that weigh more than 1 and every other word is summed with a popcount
    std::vector<std::vector<uint64_t>> pbits;
    int64_t n_words = 0, heavy_words = 0;
    bool packed() const { return flags & 8; }
    const Corpus* corpus = nullptr;     // packed + no dedup: the bits are set straight from the lines, `rows` stays empty
    bool lean() const { return (flags & 8) && (flags & 16); }

    // FastApriori.scala:46-86 genFreqItems
    void gen_freq_items(const Corpus& data) {
        // :55-58 occurrence counts (a duplicate token in one line counts twice), filter >= minCount
        std::vector<int32_t> occ(data.vocab.size(), 0);
        for (int32_t t : data.ids) ++occ[t];
        struct Ent { int32_t id; int32_t count; bool is_int; long long iv; };
        std::vector<Ent> tmp;
        for (size_t t = 0; t < occ.size(); ++t)
            if (occ[t] >= R.min_count) { long long v = 0; bool ok = parse_int(data.vocab[t], v); tmp.push_back({(int32_t)t, occ[t], ok, v}); }
        // :60 stable sortBy(-count); Spark's collect order among ties replaced by the canonical rule
        std::sort(tmp.begin(), tmp.end(), [&](const Ent& a, const Ent& b) {
            if (a.count != b.count) return a.count > b.count;
            if (a.is_int != b.is_int) return a.is_int;
            if (a.is_int && a.iv != b.iv) return a.iv < b.iv;
            return data.vocab[a.id] < data.vocab[b.id];
        });
        id_to_rank.assign(data.vocab.size(), -1);
        for (size_t i = 0; i < tmp.size(); ++i) {
            if (i && tmp[i].count == tmp[i - 1].count) R.level1_ties = 1;
            R.freq_items.push_back(data.vocab[tmp[i].id]);
            R.level1_counts.push_back(tmp[i].count);
            id_to_rank[tmp[i].id] = (int32_t)i;  // :62
        }
        // :66-75 filter frequent -> rank -> toSet, keep size > 1, merge identical baskets
        const bool dedup = !(flags & 16);
        std::unordered_map<Items, int32_t, ItemsHash> merged;
        Items s;
        for (int64_t l = 0; l < data.n_lines(); ++l) {
            s.clear();
            for (int64_t e = data.off[l]; e < data.off[l + 1]; ++e) { int32_t r = id_to_rank[data.ids[e]]; if (r >= 0) s.push_back(r); }
            std::sort(s.begin(), s.end());
            s.erase(std::unique(s.begin(), s.end()), s.end());
            if (s.size() > 1) {
                if (lean()) { ++R.total_count; continue; }
                if (!dedup) { rows.push_back(s); weights.push_back(1); continue; }
                auto it = merged.find(s);
                if (it == merged.end()) { merged.emplace(s, (int32_t)rows.size()); rows.push_back(s); weights.push_back(1); }
                else ++weights[it->second];
            }
        }
        if (!lean()) R.total_count = (int32_t)rows.size();  // :79
        corpus = &data;
    }

    // FastApriori.scala:195-210 getFreqItemsTrans — one byte per (item, distinct row); packed: one bit
    void get_freq_items_trans() {
        size_t F = R.freq_items.size();
        if (!packed()) {
            bitmaps.assign(F, std::vector<uint8_t>((size_t)R.total_count, 0));
            for (size_t r = 0; r < rows.size(); ++r)
                for (int32_t x : rows[r]) bitmaps[x][r] = 1;
            return;
        }
        if (lean()) {  // every kept line is a row of weight 1, in line order
            n_words = ((int64_t)R.total_count + 63) / 64;
            pbits.assign(F, std::vector<uint64_t>((size_t)n_words, 0));
            heavy_words = 0;
            Items s;
            int64_t p = 0;
            for (int64_t l = 0; l < corpus->n_lines(); ++l) {
                s.clear();
                for (int64_t e = corpus->off[l]; e < corpus->off[l + 1]; ++e) { int32_t r = id_to_rank[corpus->ids[e]]; if (r >= 0) s.push_back(r); }
                std::sort(s.begin(), s.end());
                s.erase(std::unique(s.begin(), s.end()), s.end());
                if (s.size() > 1) {
                    for (int32_t x : s) pbits[x][p >> 6] |= 1ull << (p & 63);
                    ++p;
                }
            }
            return;
        }
        // row ids are arbitrary in the reference (zipWithIndex of a shuffle, :73-74): order them by weight
        std::vector<int32_t> order(rows.size());
        for (size_t r = 0; r < rows.size(); ++r) order[r] = (int32_t)r;
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return weights[a] > weights[b]; });
        std::vector<int32_t> w2(rows.size());
        int64_t heavy = 0;
        n_words = ((int64_t)rows.size() + 63) / 64;
        pbits.assign(F, std::vector<uint64_t>((size_t)n_words, 0));
        for (size_t p = 0; p < order.size(); ++p) {
            const int32_t r = order[p];
            w2[p] = weights[r];
            if (weights[r] > 1) heavy = (int64_t)p + 1;
            for (int32_t x : rows[r]) pbits[x][p >> 6] |= 1ull << (p & 63);
            Items().swap(rows[r]);  // the row lists are not needed any more
        }
        weights.swap(w2);
        heavy_words = (heavy + 63) / 64;
    }

    // weighted number of rows whose bit is set in both packed operands
    int32_t weighted_and_count(const uint64_t* a, const uint64_t* b) const {
        int64_t count = 0;
        for (int64_t w = 0; w < heavy_words; ++w)
            for (uint64_t m = a[w] & b[w]; m; m &= m - 1) count += weights[(size_t)(w * 64 + __builtin_ctzll(m))];
        for (int64_t w = heavy_words; w < n_words; ++w) count += __builtin_popcountll(a[w] & b[w]);
        return (int32_t)count;
    }

    // FastApriori.scala:212-241 genTwoFreqItems — every i<j pair, full scan, weighted sum, >= minCount
    std::vector<std::pair<Items, int32_t>> gen_two_freq_items() {
        int F = (int)R.freq_items.size();
        int tc = R.total_count;
        std::vector<std::vector<std::pair<Items, int32_t>>> per_i((size_t)std::max(F, 1));
        if (lean() && F > 1) {
            // every row weighs 1: the same C(F,2) sums, accumulated block of words by block of words so that the rows'
            // slices stay in cache (a pair-at-a-time scan of 10^8-row bitmaps is bound by main memory)
            std::vector<int32_t> cnt((size_t)F * F, 0);
            const int64_t BW = 1024;
            for (int64_t w0 = 0; w0 < n_words; w0 += BW) {
                const int64_t w1 = std::min(n_words, w0 + BW);
#pragma omp parallel for schedule(dynamic, 4)
                for (int i = 0; i < F - 1; ++i) {
                    const uint64_t* x = pbits[i].data();
                    for (int j = i + 1; j < F; ++j) {
                        const uint64_t* y = pbits[j].data();
                        int64_t c = 0;
                        for (int64_t w = w0; w < w1; ++w) c += __builtin_popcountll(x[w] & y[w]);
                        cnt[(size_t)i * F + j] += (int32_t)c;
                    }
                }
            }
            std::vector<std::pair<Items, int32_t>> out;
            for (int i = 0; i < F - 1; ++i)
                for (int j = i + 1; j < F; ++j)
                    if (cnt[(size_t)i * F + j] >= R.min_count) out.push_back({Items{i, j}, cnt[(size_t)i * F + j]});
            return out;
        }
#pragma omp parallel for schedule(dynamic, 1)
        for (int i = 0; i < F - 1; ++i) {
            for (int j = i + 1; j < F; ++j) {
                int32_t count = 0;
                if (packed()) count = weighted_and_count(pbits[i].data(), pbits[j].data());
                else {
                    const uint8_t* x = bitmaps[i].data();
                    const uint8_t* y = bitmaps[j].data();
                    for (int r = 0; r < tc; ++r) if (x[r] && y[r]) count += weights[r];
                }
                if (count >= R.min_count) per_i[i].push_back({Items{i, j}, count});
            }
        }
        std::vector<std::pair<Items, int32_t>> out;
        for (auto& v : per_i) out.insert(out.end(), v.begin(), v.end());
        return out;
    }

    // FastApriori.scala:167-193 genCandidates
    std::vector<std::pair<Items, Items>> gen_candidates(const std::vector<Items>& k_items) {
        int F = (int)R.freq_items.size();
        std::unordered_set<Items, ItemsHash> k_set(k_items.begin(), k_items.end());  // :172
        std::vector<Items> ys_of(k_items.size());
        // flags bit2: only ys for which prefix(x)+y is itself frequent can survive the t = max(x) round of
        // the prune below, so start from those instead of the whole range (same result, far fewer probes)
        const bool fast = flags & 4;
        std::unordered_map<Items, Items, ItemsHash> by_prefix;
        if (fast)
            for (const Items& x : k_items) { Items p(x.begin(), x.end() - 1); by_prefix[p].push_back(x.back()); }
#pragma omp parallel for schedule(dynamic, 64)
        for (int64_t xi = 0; xi < (int64_t)k_items.size(); ++xi) {
            const Items& x = k_items[xi];
            std::vector<int32_t> cand;
            if (fast) {
                Items p(x.begin(), x.end() - 1);
                for (int32_t y : by_prefix.find(p)->second) if (y > x.back()) cand.push_back(y);
                std::sort(cand.begin(), cand.end());
            } else
            for (int y = x.back() + 1; y < F; ++y) cand.push_back(y);  // :176 (x is sorted: max = back)
            for (size_t i = 0; i < x.size() && !cand.empty(); ++i) {   // :181
                Items sub;
                for (size_t j = 0; j < x.size(); ++j) if (j != i) sub.push_back(x[j]);  // :182 x - temp(i)
                std::vector<int32_t> keep;
                for (int32_t y : cand) {                                // :183-186
                    Items probe(sub);
                    probe.push_back(y);  // y > max(x) so the probe stays sorted
                    if (k_set.count(probe)) keep.push_back(y);
                }
                cand.swap(keep);
            }
            ys_of[xi] = cand;
        }
        std::vector<std::pair<Items, Items>> out;  // :189-190 keep non-empty groups
        for (size_t xi = 0; xi < k_items.size(); ++xi)
            if (!ys_of[xi].empty()) out.push_back({k_items[xi], ys_of[xi]});
        return out;
    }

    // FastApriori.scala:132-165 genNextFreqItemsets + logicalAnd
    std::vector<std::pair<Items, int32_t>> gen_next(const std::vector<std::pair<Items, Items>>& cands) {
        int tc = R.total_count;
        std::vector<std::vector<std::pair<Items, int32_t>>> per_g(cands.size());
#pragma omp parallel
        {
            std::vector<uint8_t> common(packed() ? 0 : (size_t)tc);
            std::vector<uint64_t> pcommon(packed() ? (size_t)n_words : 0);
#pragma omp for schedule(dynamic, 4)
            for (int64_t g = 0; g < (int64_t)cands.size(); ++g) {
                const Items& sub = cands[g].first;
                if (packed()) {                 // :143-145,162-165 on words
                    for (int64_t w = 0; w < n_words; ++w) {
                        uint64_t v = ~0ull;
                        for (int32_t t : sub) v &= pbits[t][w];
                        pcommon[w] = v;
                    }
                } else
                for (int r = 0; r < tc; ++r) {  // :143-145,162-165
                    uint8_t v = 1;
                    for (int32_t t : sub) if (!bitmaps[t][r]) { v = 0; break; }
                    common[r] = v;
                }
                for (int32_t y : cands[g].second) {  // :147-152
                    int32_t count = 0;
                    if (packed()) count = weighted_and_count(pcommon.data(), pbits[y].data());
                    else {
                        const uint8_t* ya = bitmaps[y].data();
                        for (int r = 0; r < tc; ++r) if (common[r] && ya[r]) count += weights[r];
                    }
                    if (count >= R.min_count) {
                        Items s(sub);
                        s.push_back(y);
                        per_g[g].push_back({s, count});
                    }
                }
            }
        }
        std::vector<std::pair<Items, int32_t>> out;
        for (auto& v : per_g) out.insert(out.end(), v.begin(), v.end());
        return out;
    }

    // FastApriori.scala:31-44 run + :88-130 genFreqItemsets
    void run(const Corpus& data) {
        R.n_lines = data.n_lines();                                         // :38
        R.min_count = (int32_t)std::ceil(min_support * (double)R.n_lines);  // :39 (IEEE double)
        double t0 = now();
        gen_freq_items(data);                                               // :40
        R.t_level1 = now() - t0;
        int F = (int)R.freq_items.size();
        R.levels.assign(2, {});
        for (int i = 0; i < F; ++i) R.levels[1].push_back({Items{i}, R.level1_counts[i]});  // :83
        R.n_groups.assign(3, 0);
        R.n_cands.assign(3, 0);
        t0 = now();
        get_freq_items_trans();                                             // :97
        R.t_bitmaps = now() - t0; t0 = now();
        auto two = gen_two_freq_items();                                    // :104
        R.t_pairs = now() - t0;
        VLOG("F = %d, rows = %d, level1 %.1f s, bitmaps %.1f s, pairs %.1f s -> %zu", F, R.total_count, R.t_level1, R.t_bitmaps, R.t_pairs, two.size());
        R.n_groups[2] = R.n_cands[2] = (int64_t)F * (F - 1) / 2;            // :226
        R.levels.push_back(two);
        std::vector<Items> k_items;
        for (auto& p : two) k_items.push_back(p.first);
        size_t k = 3;
        while (k_items.size() >= k) {                                       // :110-111
            t0 = now();
            auto cands = gen_candidates(k_items);                           // :113
            R.t_cand += now() - t0;
            int64_t nc = 0;
            for (auto& c : cands) nc += (int64_t)c.second.size();
            R.n_groups.push_back((int64_t)cands.size());                    // :114 prints the group count
            R.n_cands.push_back(nc);
            t0 = now();
            auto next = gen_next(cands);                                    // :115
            R.t_count += now() - t0;
            VLOG("level %zu: %zu groups, %lld candidates, %zu frequent (cand %.1f s, count %.1f s so far)", k, cands.size(), (long long)nc, next.size(), R.t_cand, R.t_count);
            R.levels.push_back(next);
            k_items.clear();
            for (auto& p : next) k_items.push_back(p.first);                // :117
            ++k;
        }
    }
};

struct Ruler {
    Result& R;
    int flags;
    const std::unordered_map<std::string, int32_t>& item_to_rank;  // freqItems string -> rank

    // AssociationRules.scala:122-188 genRules
    bool gen_rules() {
        size_t maxk = R.levels.size() - 1;
        std::vector<std::vector<Rule>> by_len(maxk + 1);  // index = antecedent size
        bool fast = flags & 1;
        for (size_t k = 2; k <= maxk; ++k) {
            const auto& sups = R.levels[k];
            const auto& subs = R.levels[k - 1];
            std::unordered_map<Items, int32_t, ItemsHash> sub_index;
            if (fast) for (size_t i = 0; i < subs.size(); ++i) sub_index.emplace(subs[i].first, (int32_t)i);
            std::vector<std::vector<Rule>> per(sups.size());
#pragma omp parallel for schedule(dynamic, 16)
            for (int64_t si = 0; si < (int64_t)sups.size(); ++si) {
                const Items& S = sups[si].first;
                int32_t count = sups[si].second;
                if (fast) {
                    for (size_t i = 0; i < S.size(); ++i) {
                        Items sub;
                        for (size_t j = 0; j < S.size(); ++j) if (j != i) sub.push_back(S[j]);
                        auto it = sub_index.find(sub);
                        if (it != sub_index.end())
                            per[si].push_back({sub, S[i], (double)count / (double)subs[it->second].second});
                    }
                } else {
                    // :132-143 targets = {S - i}; scan the (k-1)-itemsets until all are found
                    std::vector<Items> targets;
                    std::vector<int32_t> missing;
                    for (size_t i = 0; i < S.size(); ++i) {
                        Items sub;
                        for (size_t j = 0; j < S.size(); ++j) if (j != i) sub.push_back(S[j]);
                        targets.push_back(sub);
                        missing.push_back(S[i]);
                    }
                    size_t left = targets.size();
                    for (size_t i = 0; i < subs.size() && left; ++i) {
                        for (size_t t = 0; t < targets.size(); ++t)
                            if (missing[t] >= 0 && targets[t] == subs[i].first) {
                                per[si].push_back({targets[t], missing[t], (double)count / (double)subs[i].second});  // :139
                                missing[t] = -1;
                                --left;
                                break;
                            }
                    }
                }
            }
            for (auto& v : per) for (auto& r : v) by_len[k - 1].push_back(r);
        }
        for (auto& v : by_len) R.raw_rules.insert(R.raw_rules.end(), v.begin(), v.end());
        // :147-148 rules.keys.min on an empty map throws
        size_t min_len = 0, max_len = 0;
        for (size_t l = 1; l <= maxk; ++l) if (!by_len[l].empty()) { if (!min_len) min_len = l; max_len = l; }
        if (!min_len) return false;
        R.before_cut.assign(max_len + 1, 0);
        R.after_cut.assign(max_len + 1, 0);
        std::vector<Rule> real(by_len[min_len]);   // :149-151
        std::vector<Rule> low(by_len[min_len]);
        for (size_t lvl = min_len + 1; lvl <= max_len; ++lvl) {  // :152
            // :154 lowLevel.groupBy(consequent)
            std::unordered_map<int32_t, std::vector<const Rule*>> by_cons;
            if (!fast) for (auto& r : low) by_cons[r.cons].push_back(&r);
            const auto& cur = by_len[lvl];
            R.before_cut[lvl] = (int64_t)cur.size();
            std::vector<uint8_t> keep(cur.size(), 0);
            // flags bit0: the kept lower-level rules hashed by (antecedent, consequent) instead of a scan of the
            // consequent's list per rule — the same three conditions (:168 strictly smaller, :173 all found, :174 any)
            std::unordered_map<Items, double, ItemsHash> low_conf;
            if (fast) {
                low_conf.reserve(low.size() * 2);
                for (auto& r : low) { Items key(r.ant); key.push_back(r.cons); low_conf.emplace(std::move(key), r.conf); }
            }
#pragma omp parallel for schedule(dynamic, 64)
            for (int64_t ri = 0; ri < (int64_t)cur.size(); ++ri) {
                const Rule& r = cur[ri];
                bool flag = true;
                if (fast) {
                    Items key;
                    for (size_t i = 0; flag && i < r.ant.size(); ++i) {
                        key.clear();
                        for (size_t j = 0; j < r.ant.size(); ++j) if (j != i) key.push_back(r.ant[j]);
                        key.push_back(r.cons);
                        auto f = low_conf.find(key);
                        if (f == low_conf.end() || f->second >= r.conf) flag = false;
                    }
                    keep[ri] = flag;
                    continue;
                }
                auto it = by_cons.find(r.cons);
                if (it != by_cons.end()) {             // :159
                    std::vector<Items> targets;        // :161-162
                    for (size_t i = 0; i < r.ant.size(); ++i) {
                        Items sub;
                        for (size_t j = 0; j < r.ant.size(); ++j) if (j != i) sub.push_back(r.ant[j]);
                        targets.push_back(sub);
                    }
                    std::vector<uint8_t> gone(targets.size(), 0);
                    size_t left = targets.size();
                    const auto& subsets = it->second;
                    for (size_t i = 0; flag && left && i < subsets.size(); ++i) {  // :165
                        for (size_t t = 0; t < targets.size(); ++t)
                            if (!gone[t] && targets[t] == subsets[i]->ant) {
                                if (subsets[i]->conf >= r.conf) flag = false;       // :168
                                gone[t] = 1;
                                --left;
                                break;
                            }
                    }
                    if (left) flag = false;            // :173
                } else flag = false;                   // :174
                keep[ri] = flag;
            }
            std::vector<Rule> filtered;
            for (size_t ri = 0; ri < cur.size(); ++ri) if (keep[ri]) filtered.push_back(cur[ri]);
            R.after_cut[lvl] = (int64_t)filtered.size();
            VLOG("cut level %zu: %zu -> %zu", lvl, cur.size(), filtered.size());
            real.insert(real.end(), filtered.begin(), filtered.end());  // :178
            low.swap(filtered);                                         // :179
        }
        R.rules.swap(real);
        return true;
    }

    // AssociationRules.scala:66-113 genAssociationRules (+ :116-120 sort, :33-64 removeRedundancy)
    void run(const Corpus& users) {
        size_t n_users = (size_t)users.n_lines();
        // the users' tokens -> ranks of frequent items (AssociationRules.scala:45 looks the strings up in itemToRank)
        std::vector<int32_t> tok_rank(users.vocab.size(), -1);
        for (size_t t = 0; t < users.vocab.size(); ++t) { auto it = item_to_rank.find(users.vocab[t]); if (it != item_to_rank.end()) tok_rank[t] = it->second; }
        R.recommends.assign(n_users, "0");
        // :42-58 filter frequent -> rank -> toSet; empty -> "0"; group identical baskets
        std::unordered_map<Items, int32_t, ItemsHash> group_of;
        std::vector<Items> baskets;
        std::vector<std::vector<int32_t>> members;
        for (size_t u = 0; u < n_users; ++u) {
            Items s;
            for (int64_t e = users.off[u]; e < users.off[u + 1]; ++e) { int32_t r = tok_rank[users.ids[e]]; if (r >= 0) s.push_back(r); }
            std::sort(s.begin(), s.end());
            s.erase(std::unique(s.begin(), s.end()), s.end());
            if (s.empty()) continue;  // :49
            auto it = group_of.find(s);
            if (it == group_of.end()) { group_of.emplace(s, (int32_t)baskets.size()); baskets.push_back(s); members.push_back({(int32_t)u}); }
            else members[it->second].push_back((int32_t)u);
        }
        double t0 = now();
        const bool have = gen_rules();
        R.t_rulegen = now() - t0;
        VLOG("rules: %zu raw, %zu kept, %.1f s", R.raw_rules.size(), R.rules.size(), R.t_rulegen);
        if (!have) { R.status = 1; R.recommends.clear(); return; }  // :147 empty.min
        t0 = now();
        // :74 sortWith(associationRulesSort): conf desc, tie -> freqItems(consequent).toInt asc; stable
        std::vector<long long> item_int(R.freq_items.size(), 0);
        std::vector<uint8_t> item_ok(R.freq_items.size(), 0);
        for (size_t i = 0; i < R.freq_items.size(); ++i) item_ok[i] = parse_int(R.freq_items[i], item_int[i]);
        int nf = 0;
        std::stable_sort(R.rules.begin(), R.rules.end(), [&](const Rule& x, const Rule& y) {
            if (x.conf > y.conf) return true;
            if (x.conf < y.conf) return false;
            if (!item_ok[x.cons] || !item_ok[y.cons]) nf = 1;  // .toInt would throw here (:119)
            return item_int[x.cons] < item_int[y.cons];
        });
        R.number_format = nf;
        R.t_sort = now() - t0;
        t0 = now();
        int F = (int)R.freq_items.size();
        // flags bit0: the same scan over the same sorted rules, on flat arrays, with a 64-bit signature of the antecedent
        // tested first (a rule whose signature has a bit outside the basket's cannot have its antecedent inside it)
        const bool fast = flags & 1;
        std::vector<uint64_t> sig;
        std::vector<int32_t> flat_items, flat_off, flat_cons;
        if (fast) {
            sig.resize(R.rules.size());
            flat_off.assign(R.rules.size() + 1, 0);
            flat_cons.resize(R.rules.size());
            for (size_t i = 0; i < R.rules.size(); ++i) {
                uint64_t sg = 0;
                for (int32_t a : R.rules[i].ant) { sg |= 1ull << (a & 63); flat_items.push_back(a); }
                sig[i] = sg;
                flat_off[i + 1] = (int32_t)flat_items.size();
                flat_cons[i] = R.rules[i].cons;
            }
        }
        // :80-106 first matching rule per distinct basket
#pragma omp parallel
        {
            std::vector<uint8_t> temp((size_t)F);
#pragma omp for schedule(dynamic, 16)
            for (int64_t b = 0; b < (int64_t)baskets.size(); ++b) {
                const Items& user = baskets[b];
                std::fill(temp.begin(), temp.end(), 0);
                for (int32_t x : user) temp[x] = 1;   // :88-89
                int32_t rec = -1;
                if (fast) {
                    uint64_t usig = 0;
                    for (int32_t x : user) usig |= 1ull << (x & 63);
                    const size_t nr = sig.size();
                    for (size_t i = 0; i < nr; ++i) {
                        if (sig[i] & ~usig) continue;
                        const int32_t lo = flat_off[i], hi = flat_off[i + 1];
                        if ((size_t)(hi - lo) > user.size() || temp[flat_cons[i]]) continue;  // :90
                        bool all = true;
                        for (int32_t e = lo; e < hi && all; ++e) all = temp[flat_items[e]];    // :97
                        if (all) { rec = flat_cons[i]; break; }
                    }
                } else
                for (const Rule& r : R.rules) {
                    if (r.ant.size() > user.size() || temp[r.cons]) continue;  // :90
                    size_t hit = 0;
                    for (int32_t a : r.ant) hit += temp[a];                    // :97 (A & user).size == A.size
                    if (hit == r.ant.size()) { rec = r.cons; break; }
                }
                if (rec >= 0) for (int32_t u : members[b]) R.recommends[u] = R.freq_items[rec];  // :104
            }
        }
        R.t_match = now() - t0;
        VLOG("sort %.1f s, match %.1f s (%zu distinct baskets)", R.t_sort, R.t_match, baskets.size());
    }
};

// Utils.scala:29-41 saveFreqItemset and :43-49 saveRecommends — the bytes of part-00000
void render_outputs(Result& R) {
    std::vector<std::string> lines;
    for (size_t k = 1; k < R.levels.size(); ++k)
        for (auto& p : R.levels[k]) {
            std::string s;
            for (size_t i = p.first.size(); i-- > 0;) {  // sortBy(-_): rank descending
                s += R.freq_items[p.first[i]];
                if (i) s += ' ';
            }
            lines.push_back(std::move(s));
        }
    std::sort(lines.begin(), lines.end());  // Java String order == byte order for ASCII item ids
    for (auto& l : lines) { R.freqitemset_text += l; R.freqitemset_text += '\n'; }
    for (auto& r : R.recommends) { R.recommends_text += r; R.recommends_text += '\n'; }
}

}  // namespace

extern "C" {

// Main.scala:23-37 sequencing on two corpora (D.dat, U.dat).  flags: bit0 = hashed subset lookup in genRules
// instead of the reference's linear scan; bit1 = skip the rule/recommendation stage; bit2 = sibling join in
// genCandidates instead of probing every y in (max(x), F); bit3 = packed bitmaps; bit4 = no row dedup.
// Every flag leaves every output unchanged (bit4: except the totalCount counter) — tests/test_oracle.py.
static void* run_corpora(const Corpus& data, const Corpus& users, double min_support, int threads, int flags, double t_parse) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    Result* R = new Result();
    double t1 = now();
    Miner m{min_support, flags, *R};
    m.run(data);
    for (size_t k = 2; k < R->levels.size(); ++k)  // canonical order for comparisons
        std::sort(R->levels[k].begin(), R->levels[k].end());
    double t2 = now();
    if (!(flags & 2)) {
        std::unordered_map<std::string, int32_t> item_to_rank;  // FastApriori.scala:62
        for (size_t i = 0; i < R->freq_items.size(); ++i) item_to_rank.emplace(R->freq_items[i], (int32_t)i);
        Ruler r{*R, flags, item_to_rank};
        r.run(users);
    }
    double t3 = now();
    if (!(flags & 32)) render_outputs(*R);
    else for (auto& r : R->recommends) { R->recommends_text += r; R->recommends_text += '\n'; }
    R->t_parse = t_parse; R->t_mine = t2 - t1; R->t_rules = t3 - t2;
    return R;
}

// d/u are the raw bytes of D.dat / U.dat.
void* fao_run(const char* d, int64_t dlen, const char* u, int64_t ulen, double min_support, int threads, int flags) {
    double t0 = now();
    Corpus data, users;
    tokenize(d, dlen, data);
    tokenize(u, ulen, users);
    return run_corpora(data, users, min_support, threads, flags, now() - t0);
}

// The same run on rows that are already integer item ids (CSR): identical to fao_run on the text form of the
// same rows (decimal ids separated by one space), without building that text.  flags bit5 (32): do not render
// the freqItemset text (large runs that are compared through levels and digests).
void* fao_run_csr(const int64_t* d_off, const int32_t* d_items, int64_t n_rows, const int64_t* u_off, const int32_t* u_items,
                  int64_t n_users, double min_support, int threads, int flags) {
    double t0 = now();
    Corpus data, users;
    corpus_from_csr(d_off, d_items, n_rows, data);
    if (n_users > 0) corpus_from_csr(u_off, u_items, n_users, users);
    return run_corpora(data, users, min_support, threads, flags, now() - t0);
}

// Order-independent digest of the kept rules and a position-dependent digest of their (confidence, consequent
// item id) sequence — what bench.py and the parity tests compare at sizes where shipping 10^7 rules is wasteful.
// The same arithmetic runs in the product's fa_rules_digest (csrc/fa_kernels.cuh: k_rule_digest).
static inline uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
    return x;
}
void fao_rules_digest(void* h, uint64_t* out /* [4]: n kept, multiset sum, multiset xor, order sum */) {
    Result* R = (Result*)h;
    uint64_t sum = 0, x = 0, ord = 0;
    std::vector<long long> item_int(R->freq_items.size(), 0);
    for (size_t i = 0; i < R->freq_items.size(); ++i) { long long v = 0; if (parse_int(R->freq_items[i], v)) item_int[i] = v; }
    for (size_t p = 0; p < R->rules.size(); ++p) {
        const Rule& r = R->rules[p];
        uint64_t bits;
        memcpy(&bits, &r.conf, 8);
        uint64_t hsh = mix64((uint64_t)(uint32_t)r.cons + 0x9E3779B97F4A7C15ull);
        for (int32_t a : r.ant) hsh = mix64(hsh ^ ((uint64_t)(uint32_t)(a + 1) * 0xD6E8FEB86659FD93ull));
        hsh = mix64(hsh ^ bits);
        sum += hsh;
        x ^= hsh;
        ord += mix64(mix64((uint64_t)p + 1) ^ bits) ^ mix64(((uint64_t)p << 32) ^ (uint64_t)(uint32_t)(int32_t)item_int[r.cons]);
    }
    out[0] = (uint64_t)R->rules.size(); out[1] = sum; out[2] = x; out[3] = ord;
}

void fao_free(void* h) { delete (Result*)h; }

int fao_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// scalar getters: 0 status, 1 n_lines, 2 min_count, 3 total_count, 4 F, 5 max level,
// 6 level1_ties, 7 number_format, 8 n_raw_rules, 9 n_rules, 10 n_users
int64_t fao_scalar(void* h, int which) {
    Result* R = (Result*)h;
    switch (which) {
        case 0: return R->status;
        case 1: return R->n_lines;
        case 2: return R->min_count;
        case 3: return R->total_count;
        case 4: return (int64_t)R->freq_items.size();
        case 5: return (int64_t)R->levels.size() - 1;
        case 6: return R->level1_ties;
        case 7: return R->number_format;
        case 8: return (int64_t)R->raw_rules.size();
        case 9: return (int64_t)R->rules.size();
        case 10: return (int64_t)R->recommends.size();
    }
    return -1;
}

double fao_time(void* h, int which) {
    Result* R = (Result*)h;
    const double v[] = {R->t_parse, R->t_mine, R->t_rules, R->t_level1, R->t_bitmaps, R->t_pairs, R->t_cand,
                        R->t_count, R->t_rulegen, R->t_sort, R->t_match};
    return (which >= 0 && which < 11) ? v[which] : -1.0;
}

int64_t fao_level_size(void* h, int k) {
    Result* R = (Result*)h;
    return (k >= 1 && k < (int)R->levels.size()) ? (int64_t)R->levels[k].size() : 0;
}
int64_t fao_level_groups(void* h, int k) { Result* R = (Result*)h; return k < (int)R->n_groups.size() ? R->n_groups[k] : 0; }
int64_t fao_level_cands(void* h, int k) { Result* R = (Result*)h; return k < (int)R->n_cands.size() ? R->n_cands[k] : 0; }
int64_t fao_cut(void* h, int lvl, int after) {
    Result* R = (Result*)h;
    auto& v = after ? R->after_cut : R->before_cut;
    return lvl < (int)v.size() ? v[lvl] : 0;
}

// level k itemsets, lexicographically sorted by rank: sets[n*k], counts[n]
void fao_level(void* h, int k, int32_t* sets, int32_t* counts) {
    Result* R = (Result*)h;
    auto& L = R->levels[k];
    for (size_t i = 0; i < L.size(); ++i) {
        for (int j = 0; j < k; ++j) sets[i * k + j] = L[i].first[j];
        counts[i] = L[i].second;
    }
}

// rules: which = 0 raw (unsorted), 1 kept + sorted.  ant is padded to `stride` with -1.
int fao_rules_max_len(void* h, int which) {
    Result* R = (Result*)h;
    size_t m = 0;
    for (auto& r : (which ? R->rules : R->raw_rules)) m = std::max(m, r.ant.size());
    return (int)m;
}
void fao_rules(void* h, int which, int stride, int32_t* ant, int32_t* cons, double* conf) {
    Result* R = (Result*)h;
    auto& V = which ? R->rules : R->raw_rules;
    for (size_t i = 0; i < V.size(); ++i) {
        for (int j = 0; j < stride; ++j) ant[i * stride + j] = j < (int)V[i].ant.size() ? V[i].ant[j] : -1;
        cons[i] = V[i].cons;
        conf[i] = V[i].conf;
    }
}

// text getters: 0 freq items (one per line, rank order), 1 freqItemset/part-00000,
// 2 recommends/part-00000
const char* fao_text(void* h, int which, int64_t* len) {
    Result* R = (Result*)h;
    if (which == 0) {
        R->scratch.clear();
        for (auto& s : R->freq_items) { R->scratch += s; R->scratch += '\n'; }
        *len = (int64_t)R->scratch.size();
        return R->scratch.data();
    }
    std::string& s = which == 1 ? R->freqitemset_text : R->recommends_text;
    *len = (int64_t)s.size();
    return s.data();
}

void fao_level1_counts(void* h, int32_t* out) {
    Result* R = (Result*)h;
    std::copy(R->level1_counts.begin(), R->level1_counts.end(), out);
}

}  // extern "C"
