/* fastapriori_b200 — C-ABI of the B200-native mining hot path of relife957/FastApriori.
 *
 * The reference has no FFI of its own: its operator interface is three Scala calls made by
 * Main (Main.scala:26-37).  Each entry point below replaces one step of those calls; the
 * citation names the reference lines it stands in for (paths under
 * src/main/scala/com/hazzacheng/AR/).  A host (the Python mirror in fastapriori_b200/, the
 * Scala/JNA Main in fastapriori_b200/scala/, or any other FFI) tokenises text, maps item
 * strings to dense ids, decides minCount and the rank order, drives the level loop and
 * writes the output files; everything that touches transactions, bitmaps, candidates,
 * rules or baskets runs in the sm_100a kernels behind these calls.
 *
 * Conventions: plain C types only; every function returns 0 on success or a non-zero
 * FA_E* code, with a message available from fa_last_error(); buffers named *_out are
 * caller-owned host memory unless stated otherwise; one fa_ctx per host thread (calls on
 * one ctx must not overlap); all device memory is owned by the ctx and released by
 * fa_destroy().  There is no CPU fallback: fa_create() fails if no CUDA device is usable.
 */
#ifndef FASTAPRIORI_B200_H
#define FASTAPRIORI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fa_ctx fa_ctx;

enum {
    FA_OK = 0,
    FA_EINVAL = 1,   /* bad argument or call order */
    FA_ECUDA = 2,    /* CUDA runtime error (message has the CUDA string) */
    FA_ENOMEM = 3,   /* device or host allocation failed */
    FA_ELIMIT = 4,   /* size beyond a documented limit */
    FA_EXCHANGE = 5  /* the exchange callback reported failure */
};

/* Sum-reduction hook for row-sharded multi-GPU runs.  When set, every
 * support-count vector is handed to it on the device, in place, right after the local
 * kernel: `dev_int32` holds `count` int32 partial sums of this rank's rows and must hold
 * the sums over all ranks on return.  `cuda_stream` is the ctx stream; work enqueued on it
 * is ordered after the counting kernel.  Return 0 on success.  Replaces Spark's
 * reduceByKey/collect of partial counts (FastApriori.scala:56,238,157). */
typedef int (*fa_allreduce_fn)(void* user, void* dev_int32, int64_t count, void* cuda_stream);

/* Lifecycle.  `cuda_stream` may be NULL (the ctx creates its own) or an existing
 * cudaStream_t the host wants all work ordered on.  Replaces `new SparkContext(conf)`
 * as the owner of executor state (Main.scala:17-22). */
int fa_create(int device, void* cuda_stream, fa_ctx** ctx_out);
void fa_destroy(fa_ctx* ctx);
const char* fa_last_error(const fa_ctx* ctx);
int fa_set_exchange(fa_ctx* ctx, fa_allreduce_fn fn, void* user);

/* Built-in exchange over NCCL (NVLink / NVSwitch): the library opens libnccl.so.2 at run time and
 * all-reduces every support-count vector itself, on the ctx stream, right behind the counting
 * kernel - no host callback per level.  One rank calls fa_comm_unique_id and ships the FA_COMM_ID_BYTES
 * to the others by any means (the Python host uses its torch.distributed group, a Spark driver
 * would broadcast them); then every rank calls fa_comm_init, which is collective.  It takes
 * precedence over a callback set with fa_set_exchange.  The communicator is destroyed by
 * fa_destroy.  Same role as the callback: Spark's reduceByKey/collect of partial counts
 * (FastApriori.scala:56,238,157). */
#define FA_COMM_ID_BYTES 128
int fa_comm_unique_id(uint8_t* id_out, char* err_out, int64_t err_cap);
int fa_comm_init(fa_ctx* ctx, const uint8_t* id, int32_t rank, int32_t world_size);

/* Transactions as CSR over dense item ids in [0, n_items): row r holds
 * item_ids[row_offsets[r] .. row_offsets[r+1]).  Duplicates inside a row are allowed and
 * count twice at level 1 only (FastApriori.scala:55 vs :69).  This is the tokenised form of
 * the RDD[Array[String]] that Utils.readAsRDD yields (Utils.scala:21-22).
 * fa_load_transactions copies from host memory; fa_attach_transactions takes DEVICE
 * pointers that must stay valid until the next load/attach or fa_destroy. */
int fa_load_transactions(fa_ctx* ctx, const int64_t* row_offsets, const int32_t* item_ids,
                         int64_t n_rows, int32_t n_items);
int fa_attach_transactions(fa_ctx* ctx, const int64_t* dev_row_offsets, const int32_t* dev_item_ids,
                           int64_t n_rows, int32_t n_items);
/* fa_load_transactions for hosts that keep dense ids in 16 bits (n_items <= 65536): half the bytes cross PCIe, the ids
 * are widened to int32 on the device.  Everything downstream is unchanged. */
int fa_load_transactions_u16(fa_ctx* ctx, const int64_t* row_offsets, const uint16_t* item_ids,
                             int64_t n_rows, int32_t n_items);

/* K1 — level-1 occurrence counts per dense id (then exchange).
 * Replaces flatMap/reduceByKey in genFreqItems (FastApriori.scala:55-56). */
int fa_count_items(fa_ctx* ctx, int32_t* counts_out /* [n_items] */);

/* data.count() (FastApriori.scala:38): the rows of ALL shards.  In a sharded run the shards' row counts are summed by the
 * same exchange that sums the level-1 counts, so this needs no collective of its own; valid after fa_count_items. */
int fa_global_rows(fa_ctx* ctx, int64_t* n_rows_out);

/* K2 — vertical tidset bitmaps, one bit per (rank, row).  `id_to_rank[id]` is the dense
 * rank 0..n_frequent-1 of a frequent item (count-descending order decided by the host) or
 * -1; `rank_counts[rank]` is that item's level-1 count.  Replaces the per-row filter/remap
 * (FastApriori.scala:66-70) and getFreqItemsTrans (FastApriori.scala:195-210).  When at
 * most 64 items are frequent, identical filtered baskets are merged into weighted rows as the reference
 * does (FastApriori.scala:66-79; bit-sliced weight planes, exact), provided that at least halves the row
 * axis (FA_WEIGHTED=0 / 1 in the environment of fa_create: never / whenever F <= 64); otherwise every row
 * weighs 1, which yields the same supports.  The level loop later trims rows that can no longer support any
 * candidate (exact; supports are unchanged), so a new pair pass needs a new fa_build_bitmaps call. */
int fa_build_bitmaps(fa_ctx* ctx, const int32_t* id_to_rank /* [n_items] */, int32_t n_frequent,
                     const int32_t* rank_counts /* [n_frequent] */);

/* K3 — supports of all C(F,2) pairs, exchange, keep >= min_count: frequent 2-itemsets.
 * Replaces genTwoFreqItems (FastApriori.scala:212-241). */
int fa_count_pairs(fa_ctx* ctx, int32_t min_count, int64_t* n_frequent_out);

/* K5 — candidates of size k from the frequent (k-1)-itemsets: for each set x the ys > max(x)
 * such that every (k-1)-subset of x+y is frequent.  n_groups_out = sets x with at least one y
 * (what the reference prints as "candidate items"), n_candidates_out = total (x, y) pairs.
 * Replaces genCandidates (FastApriori.scala:167-193).
 * Pass BOTH outputs or NEITHER.  With outputs the call waits for the device and returns the totals.  With two null
 * pointers nothing is waited for: the totals stay on the device, fa_count_candidates(k) sizes its work by host-side
 * bounds (candidates <= sibling pairs, groups <= frequent (k-1)-itemsets) and the level costs ONE host/device round trip
 * instead of two; read the totals afterwards with fa_level_counters.  The results are the same either way. */
int fa_gen_candidates(fa_ctx* ctx, int32_t k, int64_t* n_groups_out, int64_t* n_candidates_out);

/* The two totals of fa_gen_candidates(k), valid once fa_count_candidates(k) has returned (0, 0 for a level that was
 * never counted): the "candidate items" figure the reference prints per level (FastApriori.scala:114). */
int fa_level_counters(fa_ctx* ctx, int32_t k, int64_t* n_groups_out, int64_t* n_candidates_out);

/* Host-only: the bit range [lo, hi) of the K8 confidence key ~bits(conf) that can differ between two rules when every
 * support lies in [sup_min, sup_max] (fa_sort_rules sorts on that range only; 0, 64 when the bounds are unusable).
 * Exported for tests: the claim is that sorting on the range alone orders any two different confidences correctly. */
int fa_conf_key_bits(int64_t sup_min, int64_t sup_max, int32_t* lo_bit_out, int32_t* hi_bit_out);

/* K4 — support of every candidate of the last fa_gen_candidates call (prefix AND, then
 * AND + popcount per extension), exchange, keep >= min_count: frequent k-itemsets.
 * Replaces genNextFreqItemsets + logicalAnd (FastApriori.scala:132-165). */
int fa_count_candidates(fa_ctx* ctx, int32_t k, int32_t min_count, int64_t* n_frequent_out);

/* Frequent k-itemsets, k >= 1, as rank rows in lexicographic order with their supports:
 * the Array[(Set[Int], Int)] of FastApriori.run (FastApriori.scala:41-43). */
int fa_level_size(fa_ctx* ctx, int32_t k, int64_t* n_out);
int fa_get_level(fa_ctx* ctx, int32_t k, int32_t* sets_out /* [n*k] */, int32_t* counts_out /* [n] */);
/* Same copy, enqueued on the ctx's copy stream behind the kernels issued so far; it overlaps later calls
 * (the next levels, the rule stage) when the host buffers are page-locked.  The buffers may be read after
 * fa_wait_copies() returns.  Same reference step as fa_get_level (the driver-side `collect()`,
 * FastApriori.scala:157,238). */
int fa_get_level_async(fa_ctx* ctx, int32_t k, int32_t* sets_out /* [n*k] */, int32_t* counts_out /* [n] */);
int fa_wait_copies(fa_ctx* ctx);

/* Resume at the rule stage: installs the frequent itemsets of a finished run (for instance read back
 * from the `items[count]` dump of Utils.saveFreqItemsetWithCount, Utils.scala:51-63) instead of mining
 * them, so that fa_gen_rules ... fa_recommend can run without transactions.  This is the constructor
 * argument `freqItemset` of AssociationRules (AssociationRules.scala:17-21) crossing the boundary.
 * Call for k = 1, 2, ... in order.  k = 1: `sets` is ignored (rank r is row r), `counts[r]` is the
 * level-1 count of rank r, n = F.  k >= 2: `sets` is [n x k] ranks, ascending inside a row, rows in
 * strictly ascending lexicographic order; every (k-1)-subset of every row must be a row of level k-1
 * (FA_EINVAL otherwise: the rule kernels index subsets, they do not search).  Host pointers. */
int fa_put_level(fa_ctx* ctx, int32_t k, const int32_t* sets, const int32_t* counts, int64_t n);

/* K6 — for every frequent S (|S| >= 2) and i in S the rule (S-{i} -> i) with confidence
 * (double)count(S) / (double)count(S-{i}), one IEEE-754 division.
 * Replaces the first half of genRules (AssociationRules.scala:126-145). */
int fa_gen_rules(fa_ctx* ctx, int64_t* n_raw_out);

/* K7 — level-wise prune: a rule with |A| >= 2 survives iff every (A-{a} -> c) survived the
 * level below with strictly smaller confidence.  Replaces AssociationRules.scala:147-182.
 * fa_rule_cut reports the reference's "Before/After cut level" counters
 * (AssociationRules.scala:155,177) for antecedent size `level` >= 2. */
int fa_prune_rules(fa_ctx* ctx, int64_t* n_kept_out);
int fa_rule_cut(fa_ctx* ctx, int32_t level, int64_t* before_out, int64_t* after_out);

/* K8 — order kept rules by confidence descending, then by the consequent's original item id
 * as an integer ascending (`rank_to_item_int[rank]`), then by (size, set, position) so the
 * order is total.  Replaces sortWith(associationRulesSort) (AssociationRules.scala:74,116-120). */
int fa_sort_rules(fa_ctx* ctx, const int32_t* rank_to_item_int /* [n_frequent] */);

/* Parity dump of rules: which = 0 all raw rules (set order), 1 kept rules in sorted order.
 * `ant_out` rows are padded to `stride` with -1. */
int fa_rules_max_len(fa_ctx* ctx, int32_t* max_antecedent_out);
int fa_get_rules(fa_ctx* ctx, int32_t which, int32_t stride, int32_t* ant_out /* [n*stride] */,
                 int32_t* cons_out /* [n] */, double* conf_out /* [n] */);
/* Digest of the kept, sorted rules computed on the device (for audits at sizes where copying 10^7 rules out is
 * wasteful; bench.py prints it, the parity tests compare it with the same arithmetic over the oracle's rules):
 * out[0] = number of kept rules; out[1], out[2] = sum and xor over rules of h(antecedent ranks ascending,
 * consequent rank, confidence bits) - independent of the rule order; out[3] = sum over positions p of
 * g(p, confidence bits, consequent's item id as Int) - pins the sorted (confidence, item id) sequence of
 * associationRulesSort (AssociationRules.scala:116-120).  Call after fa_sort_rules. */
int fa_rules_digest(fa_ctx* ctx, uint64_t* out /* [4] */);

/* K9 — per user basket (CSR over RANKS of frequent items, duplicates allowed) the consequent
 * rank of the first rule in sorted order with antecedent inside the basket and consequent
 * outside it, or -1 (the reference's "0").  Replaces removeRedundancy + the matching loop of
 * genAssociationRules (AssociationRules.scala:33-64,80-106). */
int fa_recommend(fa_ctx* ctx, const int64_t* user_offsets, const int32_t* user_ranks, int64_t n_users,
                 int32_t* consequent_rank_out /* [n_users] */);
/* Same, from the users' dense item ids: `id_to_rank[id]` (host, n_ids entries, -1 = not frequent) is applied on
 * the device, which is the filter + rank map of removeRedundancy (AssociationRules.scala:45) moved behind the
 * boundary — the host no longer touches every token of U.dat. */
int fa_recommend_ids(fa_ctx* ctx, const int64_t* user_offsets, const int32_t* user_item_ids, int64_t n_users,
                     const int32_t* id_to_rank, int32_t n_ids, int32_t* consequent_rank_out);

/* Device-time and traffic accounting since the last fa_reset_stats (CUDA events on the ctx
 * stream around each kernel family).  count_bytes is the algorithmic traffic of K4 as
 * SURVEY.md §8d defines it: W * sum_groups(k-1+|ys|) + 4*C_k per level. */
typedef struct fa_stats {
    double ms_hist, ms_bitmap, ms_pairs, ms_candgen, ms_count, ms_threshold, ms_rules, ms_sort, ms_match;
    double ms_exchange;        /* the support-count all-reduces, including the wait for the slowest rank */
    double count_bytes;        /* K4 algorithmic bytes */
    double count_load_bytes;   /* K4: bytes of item rows the kernel actually requests (after prefix re-use), all L1/L2/HBM levels */
    double pair_ops;           /* K3: 2*N*F*F */
    int64_t count_launches, kernel_launches;
    int64_t bitmap_bytes;      /* F * pitch: resident vertical database (after trimming, if any) */
    int64_t row_trims;         /* times the level loop squeezed dead rows out of the bitmaps */
    int64_t live_rows;         /* rows still represented in the bitmaps (weighted: virtual rows over all weight planes) */
    int64_t total_count;       /* distinct baskets with >= 2 frequent items, the reference's totalCount (FastApriori.scala:79);
                                  computed when at most 64 items are frequent, else -1 */
    int64_t weighted;          /* 1: the bitmaps hold bit-sliced weight planes of the distinct baskets (FastApriori.scala:66-79) */
    int64_t distinct_baskets;  /* user baskets the last fa_recommend matched after merging identical ones
                                  (AssociationRules.scala:51-58), -1 = every line was matched on its own */
    int64_t match_retries;     /* baskets whose trie walk outgrew the first pass's per-warp stack and were matched again with a
                                  stack sized for the worst case (F x depth ranges); results are the same */
} fa_stats;
int fa_get_stats(fa_ctx* ctx, fa_stats* out);
int fa_reset_stats(fa_ctx* ctx);

/* Measurement aids (no counterpart in the reference).  fa_probe_l1_bandwidth times a kernel whose warps re-read a
 * 2 KB-per-warp working set with the same 16-byte read-only loads K4 uses, all L1 hits: the GB/s it reports is the
 * measured ceiling of the unit that bounds K4 (the L1 data pipe), the denominator bench.py reports K4 against.
 * fa_pair_counts_device exposes the F x F int32 pair matrix of the last fa_count_pairs (upper triangle valid,
 * device memory owned by the ctx, valid until the next pair pass) so that tests can compare ALL pair supports with an
 * independent matrix product. */
int fa_probe_l1_bandwidth(fa_ctx* ctx, double* gb_per_s_out);
int fa_pair_counts_device(fa_ctx* ctx, const int32_t** dev_matrix_out, int32_t* n_frequent_out);

/* Host-side text loader (no GPU work): the bytes of D.dat or U.dat -> CSR over dense ids + the id -> string
 * vocabulary, with the reference's tokenisation (`_.trim().split("\\s+")` per Hadoop text record, Utils.scala:21,23;
 * a blank line is the one-token line [""]).  Ids are dense in first-seen order.  Two-step: parse, read the sizes,
 * copy into caller buffers (`vocab_out` receives the strings joined by '\n', fa_text_vocab_bytes() bytes), free. */
typedef struct fa_text fa_text;
int fa_text_parse(const char* buf, int64_t len, fa_text** text_out);
int64_t fa_text_rows(const fa_text* t);
int64_t fa_text_nnz(const fa_text* t);
int32_t fa_text_vocab_size(const fa_text* t);
int64_t fa_text_vocab_bytes(const fa_text* t);
int fa_text_copy(const fa_text* t, int64_t* row_offsets_out, int32_t* ids_out, char* vocab_out);
void fa_text_free(fa_text* t);

/* Host-side output formatter (no GPU work) for saveFreqItemset / saveFreqItemsetWithCount (Utils.scala:38-39, :60-61):
 * one line per itemset, items by rank DESCENDING joined by one space, optionally "[count]", lines sorted as byte strings
 * (Java String order for ASCII item ids) and joined by '\n'.  `items_joined` holds the item strings by rank separated
 * by '\n'.  Add levels in any order, finish (sort_lines != 0 sorts), copy `total_bytes_out` bytes (or read them in
 * place through fa_lines_data, valid until fa_lines_free), free. */
typedef struct fa_lines fa_lines;
int fa_lines_new(const char* items_joined, int64_t bytes, int32_t n_items, fa_lines** lines_out);
int fa_lines_add_level(fa_lines* l, const int32_t* sets /* [n*k] */, const int32_t* counts_or_null /* [n] */, int64_t n, int32_t k);
int fa_lines_finish(fa_lines* l, int32_t sort_lines, int64_t* total_bytes_out);
int fa_lines_copy(const fa_lines* l, char* out);
const char* fa_lines_data(const fa_lines* l);
void fa_lines_free(fa_lines* l);

/* Allocation-free support counting.  This entry point has no counterpart in the reference; it is the
 * K4 kernel of fa_count_candidates for hosts that are not allowed to allocate device memory: a plugin
 * inside an application that owns the whole GPU through its own arena, a serving loop that cannot
 * afford the device-wide stall and driver lock of cudaMalloc/cudaFree, or a process so close to its
 * memory limit that an allocation in the middle of a job might fail.  It takes no fa_ctx and calls no
 * cudaMalloc, cudaFree or allocating library routine: inputs and outputs are the caller's DEVICE
 * buffers, and the one piece of scratch the kernel needs (the bitmap of rows that lie in some group
 * prefix's tidset) lives in a static array inside the library image, FA_NOALLOC_MAX_WORDS 32-bit
 * words = 4 MiB, which every GPU that loads the library carries until the process exits.
 *
 *   dev_bits            uint32 [n_frequent][pitch_words] item bitmaps, 16-byte aligned rows
 *   dev_prefix_sets     int32  [.][prefix_len] rows of item ranks (the frequent (k-1)-itemsets)
 *   dev_group_rows      int32  [n_groups] row of dev_prefix_sets that is the prefix of each group
 *   dev_group_offsets   int32  [n_groups+1] extensions of group g are dev_extensions[off[g]..off[g+1])
 *   dev_extensions      int32  [n_candidates] item rank appended to the group's prefix
 *   dev_counts_out      int32  [n_candidates] support of prefix+extension (cleared by the call)
 *   alive_rows_out      host int64, may be NULL: rows that contain at least one group prefix
 *   err_out/err_cap     optional host buffer for a message when the return value is not FA_OK
 *
 * Limits: pitch_words <= FA_NOALLOC_MAX_WORDS (33,554,432 rows) and a multiple of 4, prefix_len <= 64
 * (FA_NOALLOC_MAX_PREFIX: the kernel stages at most 64 prefix items per group); larger inputs are refused
 * with FA_ELIMIT.  A group may have no extensions (dev_group_offsets[g] == dev_group_offsets[g+1]): its
 * prefix still counts towards alive_rows_out, also when n_candidates is 0.  The caller's current device is
 * the same after the call as before it.  The first call on a device looks up the scratch array's device address
 * (with that device current) and should therefore be made before any stream capture begins.  Every call
 * on a device shares that one array: the call clears the part it uses before launching, keeps nothing
 * from one call to the next, and two calls whose GPU work overlaps (different streams or threads) race
 * on it — serialise them.  `cuda_stream` NULL means the default stream; with alive_rows_out non-NULL the
 * call waits for the stream, otherwise it only enqueues. */
#define FA_NOALLOC_MAX_WORDS (1 << 20)
#define FA_NOALLOC_MAX_PREFIX 64
int fa_count_groups_noalloc(int device, void* cuda_stream, const uint32_t* dev_bits, int64_t pitch_words,
                            int32_t n_frequent, const int32_t* dev_prefix_sets, int32_t prefix_len,
                            const int32_t* dev_group_rows, const int32_t* dev_group_offsets,
                            const int32_t* dev_extensions, int64_t n_groups, int64_t n_candidates,
                            int32_t* dev_counts_out, int64_t* alive_rows_out, char* err_out, int64_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* FASTAPRIORI_B200_H */
