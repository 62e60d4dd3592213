// fa_main — C++ twin of com.hazzacheng.AR.Main on the C-ABI of include/fastapriori_b200.h.
//
//   fa_main <input prefix> <output prefix> <temp (ignored)> [minSupport]
//
// Same contract as the reference's Main (Main.scala:15-41, README.md:19-41): reads <input>D.dat and
// <input>U.dat, writes <output>freqItemset/part-00000 and <output>recommends/part-00000 (+ _SUCCESS),
// minSupport defaults to 0.092 (Main.scala:23).  Only the C-ABI is used: text loading (fa_text_*), the
// kernels (fa_*), nothing else — this is what a JNI/JNA/cgo host does, in a language this image can build.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <sys/stat.h>
#include <unordered_map>
#include <vector>

#include "../../include/fastapriori_b200.h"

static fa_ctx* g_ctx = nullptr;

static void check(int rc, const char* what) {
    if (rc != 0) {
        fprintf(stderr, "fa_main: %s failed with %d: %s\n", what, rc, g_ctx ? fa_last_error(g_ctx) : "");
        exit(2);
    }
}

static std::string slurp(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) { fprintf(stderr, "fa_main: cannot read %s\n", path.c_str()); exit(1); }
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

struct Lines { std::vector<int64_t> off; std::vector<int32_t> ids; std::vector<std::string> vocab; };

static Lines load(const std::string& path) {  // Utils.scala:21,23
    const std::string buf = slurp(path);
    fa_text* t = nullptr;
    check(fa_text_parse(buf.data(), (int64_t)buf.size(), &t), "fa_text_parse");
    Lines L;
    L.off.resize((size_t)fa_text_rows(t) + 1);
    L.ids.resize((size_t)fa_text_nnz(t));
    std::string vb((size_t)fa_text_vocab_bytes(t), '\0');
    check(fa_text_copy(t, L.off.data(), L.ids.empty() ? nullptr : L.ids.data(), vb.empty() ? nullptr : &vb[0]), "fa_text_copy");
    const int nv = fa_text_vocab_size(t);
    size_t p = 0;
    for (int i = 0; i < nv; ++i) { size_t q = vb.find('\n', p); L.vocab.emplace_back(vb.substr(p, q - p)); p = q + 1; }
    fa_text_free(t);
    return L;
}

static bool to_int(const std::string& s, long long& v) {  // Scala String.toInt
    if (s.empty()) return false;
    char* e = nullptr;
    v = strtoll(s.c_str(), &e, 10);
    if (*e != 0 || (s[0] != '-' && s[0] != '+' && (s[0] < '0' || s[0] > '9'))) return false;
    return v >= -2147483648ll && v <= 2147483647ll;
}

static void write_part(const std::string& dir, const std::vector<std::string>& lines) {
    struct stat st;
    if (stat(dir.c_str(), &st) == 0) { fprintf(stderr, "fa_main: Output directory %s already exists\n", dir.c_str()); exit(3); }  // saveAsTextFile
    if (mkdir(dir.c_str(), 0777) != 0) { fprintf(stderr, "fa_main: cannot create %s\n", dir.c_str()); exit(3); }
    std::ofstream f(dir + "/part-00000", std::ios::binary);
    for (auto& l : lines) f << l << '\n';
    std::ofstream(dir + "/_SUCCESS").flush();
}

static long long ms_since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
}

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: fa_main <input prefix> <output prefix> <temp> [minSupport]\n"); return 64; }
    const std::string input = argv[1], output = argv[2];
    const char* env = getenv("FA_MIN_SUPPORT");
    const double minSupport = argc > 4 ? atof(argv[4]) : (env ? atof(env) : 0.092);
    Lines data = load(input + "D.dat"), users = load(input + "U.dat");
    check(fa_create(0, nullptr, &g_ctx), "fa_create");

    auto time1 = std::chrono::steady_clock::now();
    const int64_t count = (int64_t)data.off.size() - 1;
    const int n_items = (int)data.vocab.size();
    check(fa_load_transactions(g_ctx, data.off.data(), data.ids.data(), count, n_items), "fa_load_transactions");
    const int minCount = (int)std::ceil(minSupport * (double)count);                    // FastApriori.scala:39
    std::vector<int32_t> counts((size_t)std::max(n_items, 1));
    check(fa_count_items(g_ctx, counts.data()), "fa_count_items");                      // K1
    // FastApriori.scala:57-62: frequent items by count descending; canonical tie rule (Int value, then bytes)
    std::vector<int> freq;
    for (int i = 0; i < n_items; ++i) if (counts[i] >= minCount) freq.push_back(i);
    std::sort(freq.begin(), freq.end(), [&](int a, int b) {
        if (counts[a] != counts[b]) return counts[a] > counts[b];
        long long va = 0, vb = 0;
        const bool ia = to_int(data.vocab[a], va), ib = to_int(data.vocab[b], vb);
        if (ia != ib) return ia;
        if (ia && va != vb) return va < vb;
        return data.vocab[a] < data.vocab[b];
    });
    const int F = (int)freq.size();
    std::vector<int32_t> id_to_rank((size_t)std::max(n_items, 1), -1), rank_counts((size_t)std::max(F, 1));
    std::unordered_map<std::string, int> item_to_rank;
    std::vector<std::string> freq_items;
    for (int r = 0; r < F; ++r) {
        id_to_rank[freq[r]] = r;
        rank_counts[r] = counts[freq[r]];
        freq_items.push_back(data.vocab[freq[r]]);
        item_to_rank.emplace(freq_items.back(), r);
    }
    check(fa_build_bitmaps(g_ctx, id_to_rank.data(), F, rank_counts.data()), "fa_build_bitmaps");   // K2
    int64_t n = 0, g = 0, c = 0;
    check(fa_count_pairs(g_ctx, minCount, &n), "fa_count_pairs");                       // K3
    printf("==== 2 freq items %lld\n", (long long)n);
    int k = 3;
    while (n >= k) {                                                                    // FastApriori.scala:110-111
        check(fa_gen_candidates(g_ctx, k, &g, &c), "fa_gen_candidates");                // K5
        printf("==== %d candidate items %lld\n", k, (long long)g);
        check(fa_count_candidates(g_ctx, k, minCount, &n), "fa_count_candidates");      // K4
        printf("==== %d freq items %lld\n", k, (long long)n);
        ++k;
    }
    const int max_k = k - 1;
    std::vector<std::string> lines(freq_items);                                         // Utils.scala:29-41
    for (int kk = 2; kk <= max_k; ++kk) {
        check(fa_level_size(g_ctx, kk, &n), "fa_level_size");
        if (n == 0) continue;
        std::vector<int32_t> sets((size_t)n * kk), cnt((size_t)n);
        check(fa_get_level(g_ctx, kk, sets.data(), cnt.data()), "fa_get_level");
        for (int64_t s = 0; s < n; ++s) {
            std::string l;
            for (int j = kk - 1; j >= 0; --j) { l += freq_items[sets[s * kk + j]]; if (j) l += ' '; }
            lines.push_back(std::move(l));
        }
    }
    std::sort(lines.begin(), lines.end());
    write_part(output + "freqItemset", lines);
    printf("==== Total time for get freqItemsets %lld\n", ms_since(time1));

    auto time2 = std::chrono::steady_clock::now();                                      // AssociationRules.scala:23-31
    check(fa_gen_rules(g_ctx, &n), "fa_gen_rules");                                     // K6
    if (n == 0) { fprintf(stderr, "fa_main: empty.min (no frequent itemset of size >= 2, AssociationRules.scala:147)\n"); return 4; }
    check(fa_prune_rules(g_ctx, &n), "fa_prune_rules");                                 // K7
    printf("==== Size association rules %lld\n", (long long)n);
    std::vector<int32_t> item_int((size_t)std::max(F, 1), 0);
    for (int r = 0; r < F; ++r) { long long v = 0; item_int[r] = to_int(freq_items[r], v) ? (int32_t)v : 0; }
    check(fa_sort_rules(g_ctx, item_int.data()), "fa_sort_rules");                      // K8
    const int64_t n_users = (int64_t)users.off.size() - 1;
    std::vector<int64_t> uoff((size_t)n_users + 1, 0);
    std::vector<int32_t> uranks;
    for (int64_t u = 0; u < n_users; ++u) {                                             // AssociationRules.scala:42-47
        for (int64_t e = users.off[u]; e < users.off[u + 1]; ++e) {
            auto it = item_to_rank.find(users.vocab[users.ids[e]]);
            if (it != item_to_rank.end()) uranks.push_back(it->second);
        }
        uoff[u + 1] = (int64_t)uranks.size();
    }
    std::vector<int32_t> rec((size_t)std::max<int64_t>(n_users, 1), -1);
    check(fa_recommend(g_ctx, uoff.data(), uranks.data(), n_users, rec.data()), "fa_recommend");     // K9
    std::vector<std::string> out;
    for (int64_t u = 0; u < n_users; ++u) out.push_back(rec[u] >= 0 ? freq_items[rec[u]] : "0");       // Utils.scala:43-49
    write_part(output + "recommends", out);
    printf("==== Total time for get recommends %lld\n", ms_since(time2));
    fa_destroy(g_ctx);
    return 0;
}
