// Shared host/device plumbing for libfastapriori_b200: context, buffers, timers.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <string>
#include <vector>

#include "../../include/fastapriori_b200.h"

#define FA_MAX_LEVELS 64

// Stream-ordered device buffer that only grows (cudaMallocAsync pool keeps freed blocks).
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes, cudaStream_t s) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaError_t e = cudaFreeAsync(p, s); if (e != cudaSuccess) return e; p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMallocAsync(&p, want, s);
        if (e != cudaSuccess) { p = nullptr; return e; }
        cap = want;
        return cudaSuccess;
    }
    void release(cudaStream_t s) { if (p) cudaFreeAsync(p, s); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return (T*)p; }
};

struct Level {
    int32_t k = 0;
    int64_t n = 0;
    DevBuf sets;     // int32 [n * k], rows sorted lexicographically
    DevBuf counts;   // int32 [n]
    DevBuf sub;      // int32 [n * k]: sub[s*k+p] = index in level k-1 of the set without its p-th item (k >= 2)
    DevBuf child;    // int32 [n + 1]: rows of level k+1 that extend row z are [child[z], child[z+1]) (once level k+1 exists)
    int64_t groups = 0, cands = 0;  // K5's totals for this level (the reference's println counters), known once K4 has run
};

// Passed by value to kernels that walk the itemset trie or decode a rule id into (level, set, position).
struct LevelTab {
    const int32_t* sets[FA_MAX_LEVELS];
    const int32_t* sub[FA_MAX_LEVELS];
    const int32_t* child[FA_MAX_LEVELS];  // null when the level above does not exist
    int64_t base[FA_MAX_LEVELS + 1];      // rule-id base of level k (levels 2..max_k); base[max_k+1] = total
    int64_t nodebase[FA_MAX_LEVELS + 1];  // global node id of (level k, row 0): sum of n_q for 1 <= q < k
    int32_t max_k;
};

enum TimerClass { T_HIST = 0, T_BITMAP, T_PAIRS, T_CANDGEN, T_COUNT, T_THRESH, T_RULES, T_SORT, T_MATCH, T_EXCHANGE, T_NCLASS };

struct fa_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaStream_t copy_stream = nullptr;  // device->host copies that overlap the level loop
    cudaEvent_t copy_event = nullptr;
    // Size scalars come back through mapped pinned host memory written by a tiny kernel, not through cudaMemcpy: a
    // 4-byte copy would queue on the copy engine behind the bulk level copies that overlap the level loop.
    uint32_t* h_scalars = nullptr;  // host view
    uint32_t* d_scalars = nullptr;  // device view of the same memory
    int pre_level = 0;              // K5 of this level already has its sibling-slot count queued (h_scalars[PRE_SLOT])
    bool pre_read = false;          // ... and the host has read it already (pre_slots)
    int64_t pre_slots = 0;          // the count
    int sm_count = 148;
    std::string err;

    fa_allreduce_fn xchg = nullptr;
    void* xchg_user = nullptr;
    void* nccl_comm = nullptr;  // ncclComm_t of the built-in exchange (fa_comm_init), else null

    // transactions (CSR, dense ids)
    int64_t n_rows = 0, nnz = 0;
    int64_t global_rows = -1;  // rows of all shards, summed along with the level-1 counts (fa_global_rows); -1 = not counted yet
    int32_t n_items = 0;
    const int64_t* d_off = nullptr;
    const int32_t* d_items = nullptr;
    DevBuf own_off, own_items, own_items16;  // own_items16: staging of 16-bit ids (fa_load_transactions_u16)
    bool have_tx = false;

    // vertical database
    int32_t F = 0;
    DevBuf id2rank, bits;
    int64_t pitch_words = 0;  // uint32 words per item row, multiple of 512 (one 2 KB K4 chunk)
    bool have_bits = false;
    int64_t live_rows = 0;    // rows still represented in the bitmaps (shrinks when the level loop trims rows)
    bool bits_trimmed = false;
    bool no_trim = false, debug_trim = false;  // FA_NO_TRIM / FA_DEBUG_TRIM, read once in fa_create
    DevBuf bits2, t_active, t_alive, t_cnt, t_off;
    // weighted distinct rows: FA_WEIGHTED=0 never, 1 always when F <= 64, unset = when it at least halves the rows
    int weighted_mode = -1;
    bool weighted = false;          // the bitmaps hold weight planes of distinct baskets; chunk_shift[c] = plane of chunk c
    int64_t total_count = -1;       // distinct baskets with >= 2 frequent items (FastApriori.scala:79), -1 = not computed
    DevBuf w_masks, w_sorted, w_umask, w_weights, w_misc, chunk_shift;
    int match_scap = 160;           // shared-memory part of that stack in the first pass (FA_MATCH_SCAP <= 160, tests)
    int match_gcap = 4096;          // spill ranges per warp in K9's first pass (FA_MATCH_GCAP: tests force the second pass)
    bool user_dedup = true;         // FA_USER_DEDUP=0: match every user line even when baskets fit a 64-bit mask
    int64_t trims = 0;

    std::vector<Level> levels;  // index = k; levels[0] unused

    // candidates of the current level
    int32_t cand_k = 0;
    int64_t n_groups = 0, n_cands = 0;   // exact once read back (fa_gen_candidates with outputs, or fa_count_candidates); the device holds them in scalars words 52, 53
    int64_t ng_bound = 0, nc_bound = 0;  // what the host sizes grids and arrays by: the level below, the sibling slots
    int64_t n_slots = 0;
    int k2_smem_kb = 100;    // FA_K2_SMEM_KB: shared memory one K2 CTA may take for its bit tile.  100 KB = two CTAs per SM, so one CTA's
                             // load phase overlaps the other's write-out (measured: 200 -> 100 KB takes K2 from 0.27 to 0.16 ms at configs[1],
                             // 3.4 to 2.4 ms at 20 M rows of configs[3]; 64 and 48 KB are no better)
    int k4_warps = 2;        // FA_K4_WARPS=1|2|4|8: warps per K4 CTA (a CTA's slots are held until its slowest warp is done; measured at
                             // configs[1]: 8 -> 25.8 ms, 4 -> 25.2, 2 -> 25.05, 1 -> 27.6)
    bool k5_wait = false;    // FA_K5_WAIT=1: fa_gen_candidates always reads its totals back (the two-round-trip level loop, for A/B runs)
    DevBuf c_cnt, c_off, c_gflag, c_gpos, c_gx, c_gdesc, c_goff, c_ys, c_cx, c_counts, c_flag, c_pos;
    DevBuf c_sflag, c_spos, c_slotx, c_slotsub, c_csub;
    DevBuf c_tflag, c_tpos, c_tdesc, c_tstream;  // K4 tiles of the current level: flags/positions per group, 3 x int4 per tile, extension stream
    bool k4_groups = true;   // FA_K4=tiles: count with the sibling-tile kernel instead of the group-wise one (measured slower)

    DevBuf pair_counts, pair_tiles, pair_rowcnt, pair_rowoff;
    int pair_tiles_nt = 0;    // key of the tile list in pair_tiles (column blocks x 4 + tile shape; 0 = none yet)
    int64_t pair_tiles_count = 0;
    DevBuf scratch, scalars;  // cub temp storage; small device scalars
    DevBuf errflag;

    // rules
    bool have_rules = false, have_sorted = false;
    int64_t n_raw = 0, n_kept = 0;
    int64_t sup_min = 0, sup_max = 0;  // bounds on every support in the lattice (threshold passed in / largest item count); 0 = unknown
    LevelTab tab;
    DevBuf r_conf, r_kept, r_list, r_key32, r_key64, r_tmp32, r_tmp64, r_val, r_val2, r_order;
    DevBuf m_aoff, m_lp, m_rec, m_stack;
    int64_t n_nodes = 0;
    std::vector<int64_t> cut_before, cut_after;
    DevBuf u_off, u_items, u_out;
    DevBuf ud_idx, ud_user, ud_flag, ud_run, ud_len, ud_off, ud_ranks, ud_rec;  // distinct-basket matching (F <= 64)
    int64_t match_retries = 0;      // baskets K9 matched a second time with the worst-case stack
    int64_t distinct_baskets = -1;  // of the last fa_recommend call, -1 = baskets were not merged
    DevBuf id2rank_u;  // fa_recommend_ids: the users' dense id -> rank table
    DevBuf r_itemint;  // rank -> item id as Int, kept from fa_sort_rules for fa_rules_digest

    // accounting
    double ms[T_NCLASS] = {0};
    double count_bytes = 0, count_load_bytes = 0, pair_ops = 0;
    int64_t count_launches = 0, kernel_launches = 0;
    struct Pending { int cls; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    std::vector<cudaEvent_t> ev_pool;

    int fail(int code, const std::string& msg) { err = msg; return code; }
    int fail_cuda(cudaError_t e, const char* what, int line) {
        char buf[512];
        snprintf(buf, sizeof buf, "CUDA error %d (%s) at fa_lib.cu:%d: %s", (int)e, cudaGetErrorString(e), line, what);
        err = buf;
        return e == cudaErrorMemoryAllocation ? FA_ENOMEM : FA_ECUDA;
    }
    cudaEvent_t get_event() {
        if (!ev_pool.empty()) { cudaEvent_t e = ev_pool.back(); ev_pool.pop_back(); return e; }
        cudaEvent_t e;
        cudaEventCreate(&e);
        return e;
    }
    void tbegin(int cls) {
        if (pending.size() >= 4096) resolve_timers();  // a host that never reads the stats must not grow this without bound
        Pending p{cls, get_event(), get_event()};
        cudaEventRecord(p.a, stream);
        pending.push_back(p);
    }
    void tend() { cudaEventRecord(pending.back().b, stream); }
    void resolve_timers() {
        for (auto& p : pending) {
            float t = 0;
            if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) ms[p.cls] += t;
            ev_pool.push_back(p.a);
            ev_pool.push_back(p.b);
        }
        pending.clear();
        cudaGetLastError();  // an unrecorded event pair (error path) must not poison later launch checks
    }
};

#define CK(call)                                                                 \
    do {                                                                         \
        cudaError_t e_ = (call);                                                 \
        if (e_ != cudaSuccess) return ctx->fail_cuda(e_, #call, __LINE__);       \
    } while (0)

#define LAUNCH_CHECK()                                                           \
    do {                                                                         \
        ctx->kernel_launches++;                                                  \
        cudaError_t e_ = cudaGetLastError();                                     \
        if (e_ != cudaSuccess) return ctx->fail_cuda(e_, "kernel launch", __LINE__); \
    } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
