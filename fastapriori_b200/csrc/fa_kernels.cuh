// Hand-written sm_100a kernels of the mining hot path.  Each kernel names the reference
// lines it replaces (paths under src/main/scala/com/hazzacheng/AR/ of relife957/FastApriori).
#pragma once

#include "fa_common.cuh"

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    // read-only, do not allocate in L1: bitmap chunks are reused across SMs through L2
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ int popc4(uint4 a, uint4 b) {
    return __popc(a.x & b.x) + __popc(a.y & b.y) + __popc(a.z & b.z) + __popc(a.w & b.w);
}

// ------------------------------------------------------------------------------------------
// K1 — level-1 occurrence histogram (FastApriori.scala:55-56).  HBM-bound: reads 4*nnz bytes.
// ------------------------------------------------------------------------------------------
template <bool SMEM>
__global__ void __launch_bounds__(512) k_item_hist(const int32_t* __restrict__ items, int64_t nnz, int32_t n_items,
                                                   int32_t* __restrict__ counts, int32_t* __restrict__ errflag) {
    extern __shared__ int32_t hist[];
    if (SMEM) {
        for (int i = threadIdx.x; i < n_items; i += blockDim.x) hist[i] = 0;
        __syncthreads();
    }
    int32_t* dst = SMEM ? hist : counts;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n4 = nnz >> 2;
    const int4* items4 = reinterpret_cast<const int4*>(items);
    bool bad = false;
    for (int64_t i = tid; i < n4; i += stride) {
        const int4 v = __ldg(items4 + i);
        if ((unsigned)v.x < (unsigned)n_items) atomicAdd(dst + v.x, 1); else bad = true;
        if ((unsigned)v.y < (unsigned)n_items) atomicAdd(dst + v.y, 1); else bad = true;
        if ((unsigned)v.z < (unsigned)n_items) atomicAdd(dst + v.z, 1); else bad = true;
        if ((unsigned)v.w < (unsigned)n_items) atomicAdd(dst + v.w, 1); else bad = true;
    }
    for (int64_t i = (n4 << 2) + tid; i < nnz; i += stride) {
        const int v = items[i];
        if ((unsigned)v < (unsigned)n_items) atomicAdd(dst + v, 1); else bad = true;
    }
    if (bad) *errflag = 1;
    if (SMEM) {
        __syncthreads();
        for (int i = threadIdx.x; i < n_items; i += blockDim.x)
            if (hist[i]) atomicAdd(counts + i, hist[i]);
    }
}

// ------------------------------------------------------------------------------------------
// K2 — vertical bitmap builder (FastApriori.scala:66-70 remap + :195-210 getFreqItemsTrans).
// One CTA transposes a tile of `tile_rows` transactions into a [ranks x tile_rows] bit tile in
// shared memory (atomicOr on XOR-swizzled words), then streams it out with uint4 stores so
// each item row receives tile_rows/8 contiguous bytes.  HBM-bound: reads the CSR once per
// rank pass, writes F * pitch bytes once.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) k_build_bitmaps(const int64_t* __restrict__ off, const int32_t* __restrict__ items,
                                                       const int32_t* __restrict__ id2rank, int64_t n_rows, int tile_rows,
                                                       int rank_chunk, int F, uint32_t* __restrict__ bits,
                                                       int64_t pitch_words, int64_t n_tiles) {
    extern __shared__ uint32_t smem_u32[];
    const int wpt = tile_rows >> 5;  // words per item per tile (power of two >= 4)
    const int rank_lo = blockIdx.y * rank_chunk;
    const int rank_cnt = min(rank_chunk, F - rank_lo);
    uint32_t* tile = smem_u32;
    int32_t* roff = reinterpret_cast<int32_t*>(smem_u32 + (size_t)rank_chunk * wpt);
    const int64_t t = blockIdx.x;
    const int64_t r0 = t * tile_rows;
    const int nr = (int)min((int64_t)tile_rows, n_rows - r0);
    const int64_t base = off[r0];
    for (int i = threadIdx.x; i <= nr; i += blockDim.x) roff[i] = (int32_t)(off[r0 + i] - base);
    for (int i = threadIdx.x; i < rank_cnt * wpt; i += blockDim.x) tile[i] = 0;
    __syncthreads();
    const int nnz = roff[nr];
    // Each warp owns a contiguous slice of the tile's elements and walks it 32 x 4 elements at a time: one binary search
    // for the slice's first row, after that an element's row is found by stepping forward from the row of the step before
    // (rows are ~avg_len elements long, so a lane moves 0-3 rows), instead of a 10-step binary search per element.  The
    // item id and its rank are two dependent global loads; four elements per lane keep four such chains in flight (with
    // one CTA per SM — the bit tile fills shared memory — a single chain left the kernel latency-bound: 35 ms for 100 M
    // rows x 20 items, 8x its HBM time).
    {
        const int n_warps = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const int per_warp = ((nnz + n_warps - 1) / n_warps + 127) & ~127;
        const int e_begin = min(nnz, wid * per_warp), e_end = min(nnz, e_begin + per_warp);
        int row = 0;
        if (e_begin < e_end) {
            int lo = 0, hi = nr;  // roff[lo] <= e_begin < roff[hi]
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (roff[mid] <= e_begin) lo = mid; else hi = mid;
            }
            row = lo;
        }
        for (int e0 = e_begin; e0 < e_end; e0 += 128) {
            int id[4], rk[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int e = e0 + q * 32 + lane;
                id[q] = e < e_end ? __ldg(items + base + e) : -1;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) rk[q] = id[q] >= 0 ? __ldg(id2rank + id[q]) - rank_lo : -1;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int e = e0 + q * 32 + lane;
                int r = row;
                if (e < e_end)
                    while (roff[r + 1] <= e) ++r;   // e < nnz = roff[nr], so r stays below nr
                if ((unsigned)rk[q] < (unsigned)rank_cnt)
                    atomicOr(&tile[rk[q] * wpt + ((r >> 5) ^ (rk[q] & (wpt - 1)))], 1u << (r & 31));
                row = __shfl_sync(0xffffffffu, r, 31);   // the next 32 elements start at or after this lane's row
            }
        }
    }
    __syncthreads();
    const int qpi = wpt >> 2;  // uint4 per item per tile
    const int64_t w0 = t * wpt;
    // the last tile also clears the padding words up to the row pitch
    const int ext_q = (t == n_tiles - 1) ? (int)((pitch_words - w0) >> 2) : qpi;
    const int total = rank_cnt * ext_q;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int rk = i / ext_q, q = i - rk * ext_q;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (q < qpi) {
            const int sw = rk & (wpt - 1);
            const uint32_t* row = tile + rk * wpt;
            v.x = row[(4 * q + 0) ^ sw];
            v.y = row[(4 * q + 1) ^ sw];
            v.z = row[(4 * q + 2) ^ sw];
            v.w = row[(4 * q + 3) ^ sw];
        }
        *reinterpret_cast<uint4*>(bits + (int64_t)(rank_lo + rk) * pitch_words + w0 + 4 * q) = v;
    }
}

// ------------------------------------------------------------------------------------------
// Weighted distinct rows (FastApriori.scala:66-79: identical filtered baskets are merged and every support is a
// weighted sum, :150-151, :234-235).  With at most 64 frequent items a filtered basket IS a 64-bit mask, so the merge is
// exact: one mask per row (rows with fewer than two frequent items drop out, :70), a radix sort, a run-length encode.
// The weights are then bit-sliced: a distinct basket of weight w appears once in every PLANE b for which bit b of w is
// set, planes are laid out one after the other (each padded to whole 2 KB chunks), and a chunk of plane b counts with
// weight 2^b — support = sum_b 2^b * popc(AND restricted to plane b) — so K3/K4 only shift their per-chunk popcount.
// The row axis shrinks from N to sum_d popc(w_d) virtual rows.
// ------------------------------------------------------------------------------------------
__global__ void k_row_masks(const int64_t* __restrict__ off, const int32_t* __restrict__ items, const int32_t* __restrict__ id2rank,
                            int64_t n_rows, int min_items, unsigned long long* __restrict__ masks) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    unsigned long long mk = 0;
    for (int64_t e = off[r]; e < off[r + 1]; ++e) {
        const int it = __ldg(items + e);
        const int rk = id2rank ? __ldg(id2rank + it) : it;  // null table: the items are ranks already
        if ((unsigned)rk < 64u) mk |= 1ull << rk;
    }
    masks[r] = __popcll(mk) >= min_items ? mk : 0ull;
}
// per weight plane: how many distinct baskets have that bit set; slot 32: distinct non-empty baskets
__global__ void k_plane_counts(const unsigned long long* __restrict__ umask, const int32_t* __restrict__ w, const int32_t* __restrict__ n_ptr,
                               unsigned* __restrict__ plane_cnt) {
    __shared__ unsigned cnt[33];
    if (threadIdx.x < 33) cnt[threadIdx.x] = 0;
    __syncthreads();
    const int n = *n_ptr;
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < n; d += gridDim.x * blockDim.x) {
        if (umask[d] == 0) continue;
        atomicAdd(&cnt[32], 1u);
        for (unsigned x = (unsigned)w[d]; x; x &= x - 1) atomicAdd(&cnt[__ffs(x) - 1], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 33 && cnt[threadIdx.x]) atomicAdd(plane_cnt + threadIdx.x, cnt[threadIdx.x]);
}
struct PlaneOffsets { long long row0[33]; };  // first virtual row of plane b (multiples of 16384); row0[32] = total
__global__ void k_weighted_scatter(const unsigned long long* __restrict__ umask, const int32_t* __restrict__ w, int n, PlaneOffsets po,
                                   unsigned* __restrict__ cursor, uint32_t* __restrict__ bits, int64_t pitch_words) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= n) return;
    const unsigned long long mk = umask[d];
    if (mk == 0) return;
    for (unsigned x = (unsigned)w[d]; x; x &= x - 1) {
        const int b = __ffs(x) - 1;
        const long long vr = po.row0[b] + atomicAdd(cursor + b, 1u);  // the order inside a plane is free: supports are sums
        for (unsigned long long y = mk; y; y &= y - 1) {
            const int i = __ffsll((long long)y) - 1;
            atomicOr(bits + (int64_t)i * pitch_words + (vr >> 5), 1u << (vr & 31));
        }
    }
}
__global__ void k_chunk_shifts(PlaneOffsets po, int64_t n_chunks, uint8_t* __restrict__ shift) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_chunks) return;
    const long long r = c * 16384;
    int b = 0;
    while (b < 31 && r >= po.row0[b + 1]) ++b;
    shift[c] = (uint8_t)b;
}

// K9 prologue for at most 64 frequent items (AssociationRules.scala:51-58: identical baskets are matched once):
// distinct basket masks -> CSR of ranks, and the per-user scatter of the distinct baskets' recommendations
__global__ void k_run_heads(const unsigned long long* __restrict__ sorted, const int32_t* __restrict__ flag, const int32_t* __restrict__ run_of,
                            int64_t n, unsigned long long* __restrict__ umask, int64_t* __restrict__ len) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && (i == 0 || flag[i])) { umask[run_of[i]] = sorted[i]; len[run_of[i]] = __popcll(sorted[i]); }
}
__global__ void k_mask_expand(const unsigned long long* __restrict__ umask, const int64_t* __restrict__ off, int n, int32_t* __restrict__ ranks) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= n) return;
    int64_t o = off[d];
    for (unsigned long long y = umask[d]; y; y &= y - 1) ranks[o++] = __ffsll((long long)y) - 1;
}
__global__ void k_run_starts(const unsigned long long* __restrict__ sorted, int64_t n, int32_t* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = i > 0 && sorted[i] != sorted[i - 1];
}
__global__ void k_scatter_runs(const int32_t* __restrict__ user_of, const int32_t* __restrict__ run_of, const int32_t* __restrict__ rec,
                               int64_t n, int32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[user_of[i]] = rec[run_of[i]];
}

// ------------------------------------------------------------------------------------------
// K3 — pair pass as an AND+popcount "binary GEMM" (FastApriori.scala:212-241).
// 64x64 item tile per CTA, upper-triangle tiles only, split over the row axis; 4x4 outputs
// per thread fed by conflict-free LDS.128.  POPC-pipe-bound.
// ------------------------------------------------------------------------------------------
#define PAIR_T 64
#define PAIR_KW 32
#define PAIR_STR 36

__global__ void __launch_bounds__(256) k_pair_popc(const uint32_t* __restrict__ bits, int64_t pitch_words, int F,
                                                   int64_t words_per_split, const int2* __restrict__ tiles,
                                                   int32_t* __restrict__ P, const uint8_t* __restrict__ chunk_shift) {
    __shared__ __align__(16) uint32_t A[PAIR_T * PAIR_STR];
    __shared__ __align__(16) uint32_t B[PAIR_T * PAIR_STR];
    const int2 tl = tiles[blockIdx.x];
    const int i0 = tl.x * PAIR_T, j0 = tl.y * PAIR_T;
    const int64_t w_begin = (int64_t)blockIdx.y * words_per_split;
    const int64_t w_end = min(pitch_words, w_begin + words_per_split);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    int acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 0;
    for (int64_t w = w_begin; w < w_end; w += PAIR_KW) {
        const int sh = chunk_shift ? chunk_shift[w >> 9] : 0;  // weighted rows: this 2 KB chunk counts with weight 2^sh
        for (int l = threadIdx.x; l < PAIR_T * 8; l += 256) {
            const int r = l >> 3, q = l & 7;
            uint4 va = make_uint4(0u, 0u, 0u, 0u), vb = va;
            if (i0 + r < F) va = ldg_stream(reinterpret_cast<const uint4*>(bits + (int64_t)(i0 + r) * pitch_words + w) + q);
            if (j0 + r < F) vb = ldg_stream(reinterpret_cast<const uint4*>(bits + (int64_t)(j0 + r) * pitch_words + w) + q);
            *reinterpret_cast<uint4*>(A + r * PAIR_STR + 4 * q) = va;
            *reinterpret_cast<uint4*>(B + r * PAIR_STR + 4 * q) = vb;
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            uint4 a[4], b[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                a[r] = *reinterpret_cast<const uint4*>(A + (ty + 16 * r) * PAIR_STR + 4 * q);
                b[r] = *reinterpret_cast<const uint4*>(B + (tx + 16 * r) * PAIR_STR + 4 * q);
            }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] += popc4(a[r], b[c]) << sh;
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int i = i0 + ty + 16 * r, j = j0 + tx + 16 * c;
            if (i < j && j < F && acc[r][c]) atomicAdd(P + (int64_t)i * F + j, acc[r][c]);
        }
}

// frequent pairs per row i (j > i), then ordered compaction: one warp per row
template <bool WRITE>
__global__ void __launch_bounds__(256) k_pair_select(const int32_t* __restrict__ P, int F, int min_count,
                                                     int32_t* __restrict__ rowcnt, const int32_t* __restrict__ rowoff,
                                                     int32_t* __restrict__ sets2, int32_t* __restrict__ sub2,
                                                     int32_t* __restrict__ counts2) {
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= F) return;
    int total = 0;
    const int32_t* row = P + (int64_t)i * F;
    for (int jb = i + 1; jb < F; jb += 32) {
        const int j = jb + lane;
        const int c = j < F ? row[j] : 0;
        const bool ok = j < F && c >= min_count;
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (WRITE && ok) {
            const int pos = rowoff[i] + total + __popc(m & ((1u << lane) - 1));
            sets2[2 * (int64_t)pos] = i;
            sets2[2 * (int64_t)pos + 1] = j;
            sub2[2 * (int64_t)pos] = j;      // {i,j} - {i} = {j}: level-1 index == rank
            sub2[2 * (int64_t)pos + 1] = i;  // {i,j} - {j} = {i}: the parent
            counts2[pos] = c;
        }
        total += __popc(m);
    }
    if (!WRITE && lane == 0) rowcnt[i] = total;
}

// Multi-GPU pair pass: only P[i][j], j > i, carries counts, so before the all-reduce the useful part of every row of the
// bottom half (row i, F-1-i entries) is folded into the unused head of row F-1-i of the top half (which has F-1-i free
// entries before its diagonal), the all-reduce covers the top ceil(F/2) rows only — half the bytes on the links — and
// the entries are folded back afterwards.  One CTA per bottom row.
template <bool UNFOLD>
__global__ void k_pair_fold(int32_t* __restrict__ P, int F) {
    const int h = (F + 1) / 2;
    const int i = h + blockIdx.x;      // bottom-half row
    if (i >= F) return;
    const int t = F - 1 - i;           // its partner in the top half; t < h, and t entries to move
    int32_t* src = P + (int64_t)i * F + i + 1;
    int32_t* dst = P + (int64_t)t * F;
    for (int c = threadIdx.x; c < t; c += blockDim.x) {
        if (UNFOLD) { src[c] = dst[c]; dst[c] = 0; }
        else dst[c] = src[c];
    }
}

// ------------------------------------------------------------------------------------------
// K5 — candidate join + subset prune (FastApriori.scala:167-193) on the itemset trie.
// Every level keeps, next to its sorted rows, `sub[s*k+p]` = index one level down of the set
// without its p-th item, and `child[z]..child[z+1]` = the rows one level up that extend row z
// (rows are sorted, so children are contiguous and ordered by their last item).
// For x = row i of level m = k-1 the only possible ys are the last items of the siblings that
// follow x (same parent).  x+y is a candidate iff for every p < m-1 the set (x - x[p]) + y is
// frequent, i.e. y is a child of row sub[i][p]: one short binary search inside that family.
// The searches also yield the candidate's own `sub` row: [q_0 .. q_{m-2}, sibling, i].
// LPX lanes per x (8, 16 or 32, picked from the average number of following siblings: deep
// levels have ~1-2 per x, level 3 has hundreds), one lane per sibling; every (x, sibling) pair
// owns a slot, so there is a single pass; valid slots are compacted afterwards.  Latency/L2-bound.
// ------------------------------------------------------------------------------------------
// (the level may have been compacted by a kernel that is still ahead in the stream: then its size is only on the device,
// `n_dev`, and the host sizes grid and arrays by a bound; cnt[n..n_bound] = 0, so the scan's total sits at [n_bound])
__global__ void k_sibling_counts(const int32_t* __restrict__ sub, const int32_t* __restrict__ pchild, int n_bound,
                                 const int32_t* __restrict__ n_dev, int m, int32_t* __restrict__ cnt) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_bound) return;
    const int n = n_dev ? *n_dev : n_bound;
    if (i >= n) { cnt[i] = 0; return; }
    const int par = sub[(int64_t)i * m + (m - 1)];
    cnt[i] = pchild[par + 1] - i - 1;
}

template <int LPX>
__global__ void __launch_bounds__(256) k_join_candidates(const int32_t* __restrict__ sets, const int32_t* __restrict__ sub,
                                                         const int32_t* __restrict__ pchild, int n, int m,
                                                         const int32_t* __restrict__ soff, int32_t* __restrict__ flag,
                                                         int32_t* __restrict__ slot_x, int32_t* __restrict__ slot_sub) {
    const int i = blockIdx.x * (256 / LPX) + threadIdx.x / LPX;
    const int lane = threadIdx.x % LPX;
    if (i >= n) return;
    const int32_t* xs = sub + (int64_t)i * m;
    const int fam_end = pchild[xs[m - 1] + 1];
    const int s0 = soff[i];
    const int k = m + 1;
    for (int j = i + 1 + lane; j < fam_end; j += LPX) {
        const int slot = s0 + (j - i - 1);
        const int y = sets[(int64_t)j * m + (m - 1)];
        int32_t* qs = slot_sub + (int64_t)slot * k;
        bool valid = true;
        for (int p = 0; valid && p < m - 1; ++p) {
            const int z = xs[p];
            int lo = pchild[z], hi = pchild[z + 1];
            int found = -1;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                const int v = sets[(int64_t)mid * m + (m - 1)];
                if (v == y) { found = mid; break; }
                if (v < y) lo = mid + 1; else hi = mid;
            }
            valid = found >= 0;
            qs[p] = found;
        }
        if (valid) { qs[m - 1] = j; qs[m] = i; }
        flag[slot] = valid;
        slot_x[slot] = i;
    }
}

// per-x candidate counts from the scanned slot flags (pos has n_slots + 1 entries)
__global__ void k_x_counts(const int32_t* __restrict__ soff, const int32_t* __restrict__ pos, int n, int32_t* __restrict__ cnt,
                           int32_t* __restrict__ gflag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = pos[soff[i + 1]] - pos[soff[i]];
    cnt[i] = c;
    gflag[i] = c > 0;
}
// (the thread of the last x also closes the offsets and leaves the level's group and candidate totals in lvl[0], lvl[1]:
// K4 and the kernels around it read them there, so the host need not wait for them before it queues the level)
__global__ void k_write_groups2(const int32_t* __restrict__ cnt, const int32_t* __restrict__ soff,
                                const int32_t* __restrict__ pos, const int32_t* __restrict__ gpos, int n, int64_t n_slots,
                                int32_t* __restrict__ gx, int32_t* __restrict__ goff, int32_t* __restrict__ lvl) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && cnt[i] > 0) { gx[gpos[i]] = i; goff[gpos[i]] = pos[soff[i]]; }
    if (i == n - 1) {
        const int ng = gpos[i] + (cnt[i] > 0), nc = pos[n_slots];
        goff[ng] = nc;
        lvl[0] = ng;
        lvl[1] = nc;
    }
}
// valid slots -> dense candidate arrays (ys, cx, csub[k])
__global__ void k_compact_slots(const int32_t* __restrict__ sets, int m, const int32_t* __restrict__ soff,
                                const int32_t* __restrict__ flag, const int32_t* __restrict__ pos,
                                const int32_t* __restrict__ slot_x, const int32_t* __restrict__ slot_sub, int64_t n_slots,
                                int32_t* __restrict__ ys, int32_t* __restrict__ cx, int32_t* __restrict__ csub) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots || !flag[s]) return;
    const int k = m + 1;
    const int i = slot_x[s];
    const int j = i + 1 + (int)(s - soff[i]);
    const int c = pos[s];
    ys[c] = sets[(int64_t)j * m + (m - 1)];
    cx[c] = i;
    for (int p = 0; p < k; ++p) csub[(int64_t)c * k + p] = slot_sub[s * k + p];
}

// fa_load_transactions_u16: 16-bit ids as they crossed PCIe -> the int32 ids every kernel reads (8 ids per thread and step)
__global__ void k_widen_u16(const uint16_t* __restrict__ src, int32_t* __restrict__ dst, int64_t n) {
    const int64_t n8 = n >> 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(src) + i);
        int4 a, b;
        a.x = v.x & 0xffff; a.y = v.x >> 16; a.z = v.y & 0xffff; a.w = v.y >> 16;
        b.x = v.z & 0xffff; b.y = v.z >> 16; b.z = v.w & 0xffff; b.w = v.w >> 16;
        reinterpret_cast<int4*>(dst)[2 * i] = a;
        reinterpret_cast<int4*>(dst)[2 * i + 1] = b;
    }
    for (int64_t i = (n8 << 3) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
}
__global__ void k_set3_i32(int32_t* p, int32_t a, int32_t b, int32_t c) { p[0] = a; p[1] = b; p[2] = c; }
__global__ void k_flag_eq(const int32_t* __restrict__ v, int64_t n, int32_t x, uint8_t* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = v[i] == x;
}
__global__ void k_flag_positive(const int32_t* __restrict__ v, int64_t n, int32_t* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = v[i] > 0;
}
// (n is the host's bound, *n_dev the number of entries that exist)
__global__ void k_flag_ge(const int32_t* __restrict__ v, int64_t n, const int32_t* __restrict__ n_dev, int32_t thr,
                          int32_t* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = i < *n_dev && v[i] >= thr;
}
// AssociationRules.scala:45 on the device: dense user item ids -> ranks (-1 = not a frequent item)
__global__ void k_map_ids(int32_t* __restrict__ items, int64_t n, const int32_t* __restrict__ map, int32_t n_map) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int32_t v = items[i];
        items[i] = (v >= 0 && v < n_map) ? __ldg(map + v) : -1;
    }
}
// gathers up to four short device arrays into (mapped host) dst, back to back: how sizes and counters reach the host
// without the copy engine, one launch per read-back
struct WordSrcs { const uint32_t* p[4]; int n[4]; };
__global__ void k_copy_words(WordSrcs s, uint32_t* __restrict__ dst) {
    int off = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        for (int i = threadIdx.x; i < s.n[q]; i += blockDim.x) dst[off + i] = s.p[q][i];
        off += s.n[q];
    }
}

// ------------------------------------------------------------------------------------------
// K4 — k-itemset support counting (FastApriori.scala:132-165).  A warp owns a run of G
// consecutive prefix groups on one row chunk (32*U uint4 = 2 KB of every item row).  Per group
// it ANDs the k-1 prefix bitmaps of that chunk into registers (the reference's `commonArray`)
// — groups are sorted, so when the first k-2 items repeat only the last prefix row is re-read —
// then streams every extension bitmap against it with AND + POPC, one REDUX + one RED.ADD per
// candidate.  Units are ordered chunk-major, so at any moment the resident warps touch the same
// 2 KB column of the F bitmaps: that column is read from HBM once per level and served from L2
// afterwards.  The OR of the prefix tidsets is accumulated per run and flushed into `alive`:
// rows outside it cannot support any deeper candidate (see the trimming kernels below).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}

template <int U, bool TAIL>
__device__ __forceinline__ void load_row(uint4 (&dst)[U], const uint4* __restrict__ base, uint32_t pitch_q32, int item,
                                         const bool (&ok)[U]) {
    // base = bits + this lane's first uint4 of the chunk; rows are < 2^31 uint4 long, so one IMAD.WIDE.U32 per row
    const uint4* p = base + (uint64_t)(uint32_t)item * pitch_q32;
    if constexpr (!TAIL) {  // the row pitch is a whole number of chunks: no bounds predicates, no zero fill
#pragma unroll
        for (int u = 0; u < U; ++u) dst[u] = __ldg(p + u * 32);
    } else {
#pragma unroll
        for (int u = 0; u < U; ++u) dst[u] = ok[u] ? __ldg(p + u * 32) : make_uint4(0u, 0u, 0u, 0u);
    }
    // (__ldg = read-only path that allocates in L1: the warps of an SM work on the same row chunk and the rows of one
    // prefix family are re-read within microseconds, so most of these loads hit in L1)
}
template <int U>
__device__ __forceinline__ int popc_rows(const uint4 (&a)[U], const uint4 (&b)[U]) {
    int s = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) s += popc4(a[u], b[u]);
    return s;
}

// (the stand-alone form of K4: fa_count_groups_noalloc runs this kernel; the level loop runs k_count_groups2 below,
// which reads per-group descriptors written by K5 instead of comparing prefix items per chunk.  DESIGN.md section 5 lists
// what was measured and rejected: more CTAs/SM, 1 KB chunks, prefetching, family tiles.)
#ifndef K4_MIN_BLOCKS
#define K4_MIN_BLOCKS 2
#endif
template <int U, int G, bool TAIL>
__global__ void __launch_bounds__(256, K4_MIN_BLOCKS) k_count_groups(const uint4* __restrict__ bits, int64_t pitch_q,
                                                         const int32_t* __restrict__ prev_sets, int m,
                                                         const int32_t* __restrict__ gx, const int32_t* __restrict__ goff,
                                                         const int32_t* __restrict__ ys, int64_t n_groups, int64_t n_chunks,
                                                         int32_t* __restrict__ counts, uint32_t* __restrict__ alive) {
    static_assert(G <= 30, "run metadata lives in one warp's lanes");
    const int META_X = m > 32 ? 64 : 32;  // prefix-item slots per group (shared memory comes out of L1: keep it small)
    const int META_W = META_X + 32;      // + 32 extension ids
    extern __shared__ int32_t meta_all[];
    int32_t* meta = meta_all + (threadIdx.x >> 5) * (G * META_W);
    int32_t* pm2_all = meta_all + 8 * G * META_W;  // after the 8 warps' metadata: 32*U uint4 per warp
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t n_runs = (n_groups + G - 1) / G;
    const int64_t total = n_runs * n_chunks;
    for (int64_t unit = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < total; unit += warps) {
        const int64_t chunk = unit / n_runs;
        const int64_t run = unit - chunk * n_runs;
        const int64_t q0 = chunk * (32 * U) + lane;
        const uint4* base = bits + q0;
        const uint32_t pq32 = (uint32_t)pitch_q;
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) ok[u] = !TAIL || q0 + u * 32 < pitch_q;
        const int64_t gbase = run * G;
        const int ngr = (int)min((int64_t)G, n_groups - gbase);
        // run metadata, one group per lane: prefix row index and extension range
        const int my_gx = lane < ngr ? gx[gbase + lane] : 0;
        const int my_goff = lane <= ngr ? goff[gbase + lane] : 0;
        uint4 pm1[U], al[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { pm1[u] = make_uint4(~0u, ~0u, ~0u, ~0u); al[u] = make_uint4(0u, 0u, 0u, 0u); }
        // Stage the whole run's metadata in this warp's shared memory with one burst of independent loads: the prefix
        // items of every group (positions 0..63) and its first 32 extension ids.  These are random rows of prev_sets /
        // ys (L2 latency); fetched one group ahead they were the top stall of the kernel (ncu r1e, 18 % of samples).
        __syncwarp();
        for (int gi = 0; gi < ngr; ++gi) {  // cp.async: no registers held, every copy of the run in flight at once
            const int32_t* xrow = prev_sets + (int64_t)__shfl_sync(FULL, my_gx, gi) * m;
            const int jb_ = __shfl_sync(FULL, my_goff, gi), je_ = __shfl_sync(FULL, my_goff, gi + 1);
            if (lane < m) cp_async4(meta + gi * META_W + lane, xrow + lane);
            if (lane + 32 < m) cp_async4(meta + gi * META_W + 32 + lane, xrow + lane + 32);
            if (jb_ + lane < je_) cp_async4(meta + gi * META_W + META_X + lane, ys + jb_ + lane);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        int pxa = -2, pxb = -2;
        for (int gi = 0; gi < ngr; ++gi) {
            const int xa = lane < m ? meta[gi * META_W + lane] : -1;
            const int xb = lane + 32 < m ? meta[gi * META_W + 32 + lane] : -1;
            const int yv = meta[gi * META_W + META_X + lane];  // lanes past the group's extensions hold stale ids that are never used
            const int jb = __shfl_sync(FULL, my_goff, gi), je = __shfl_sync(FULL, my_goff, gi + 1);
            const bool eq = (lane >= m - 1 || xa == pxa) && (lane + 32 >= m - 1 || xb == pxb);
            const bool same = __all_sync(FULL, eq);
            const int nc = je - jb;
            const int last = __shfl_sync(FULL, (m - 1) < 32 ? xa : xb, (m - 1) & 31);
            const int y0 = __shfl_sync(FULL, yv, 0), y1 = __shfl_sync(FULL, yv, 1);
            // the last prefix row and the first two extension rows do not depend on each other: issue them together
            uint4 rl[U], r0[U], r1[U];
            load_row<U, TAIL>(rl, base, pq32, last, ok);
            load_row<U, TAIL>(r0, base, pq32, nc > 0 ? y0 : last, ok);  // (callers of the C-ABI may pass empty groups)
            if (nc > 1) load_row<U, TAIL>(r1, base, pq32, y1, ok);  // about half of the groups have exactly one
            if (!same) {
                // pm1 = AND of the first m-1 prefix rows.  The AND of the first m-2 (pm2) lives in this warp's shared
                // memory: moving on to the next family usually changes only item m-2, and then one row load suffices.
                uint4* pm2 = reinterpret_cast<uint4*>(pm2_all) + (threadIdx.x >> 5) * (32 * U);
                const bool eq2 = (lane >= m - 2 || xa == pxa) && (lane + 32 >= m - 2 || xb == pxb);
                const bool keep2 = m >= 3 && __all_sync(FULL, eq2);
                if (!keep2) {
#pragma unroll
                    for (int u = 0; u < U; ++u) pm1[u] = make_uint4(~0u, ~0u, ~0u, ~0u);
                    int t = 0;
                    for (; t + 1 < m - 2; t += 2) {  // two rows in flight per step
                        const int ia = __shfl_sync(FULL, t < 32 ? xa : xb, t & 31);
                        const int ib = __shfl_sync(FULL, (t + 1) < 32 ? xa : xb, (t + 1) & 31);
                        uint4 ra[U], rb[U];
                        load_row<U, TAIL>(ra, base, pq32, ia, ok);
                        load_row<U, TAIL>(rb, base, pq32, ib, ok);
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            pm1[u].x &= ra[u].x & rb[u].x; pm1[u].y &= ra[u].y & rb[u].y;
                            pm1[u].z &= ra[u].z & rb[u].z; pm1[u].w &= ra[u].w & rb[u].w;
                        }
                    }
                    if (t < m - 2) {
                        const int ia = __shfl_sync(FULL, t < 32 ? xa : xb, t & 31);
                        uint4 ra[U];
                        load_row<U, TAIL>(ra, base, pq32, ia, ok);
#pragma unroll
                        for (int u = 0; u < U; ++u) { pm1[u].x &= ra[u].x; pm1[u].y &= ra[u].y; pm1[u].z &= ra[u].z; pm1[u].w &= ra[u].w; }
                    }
                    if (m >= 3) {
#pragma unroll
                        for (int u = 0; u < U; ++u) pm2[u * 32 + lane] = pm1[u];
                    }
                } else {
#pragma unroll
                    for (int u = 0; u < U; ++u) pm1[u] = pm2[u * 32 + lane];
                }
                if (m >= 2) {  // item m-2 closes pm1 (for m == 2 this is the only prefix row)
                    const int ia = __shfl_sync(FULL, (m - 2) < 32 ? xa : xb, (m - 2) & 31);
                    uint4 ra[U];
                    load_row<U, TAIL>(ra, base, pq32, ia, ok);
#pragma unroll
                    for (int u = 0; u < U; ++u) { pm1[u].x &= ra[u].x; pm1[u].y &= ra[u].y; pm1[u].z &= ra[u].z; pm1[u].w &= ra[u].w; }
                }
            }
            uint4 c[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                c[u] = make_uint4(pm1[u].x & rl[u].x, pm1[u].y & rl[u].y, pm1[u].z & rl[u].z, pm1[u].w & rl[u].w);
                al[u].x |= c[u].x; al[u].y |= c[u].y; al[u].z |= c[u].z; al[u].w |= c[u].w;
            }
            {
                const int s0 = __reduce_add_sync(FULL, popc_rows<U>(c, r0));
                if (lane == 0 && nc > 0 && s0) atomicAdd(counts + jb, s0);
                if (nc > 1) {
                    const int s1 = __reduce_add_sync(FULL, popc_rows<U>(c, r1));
                    if (lane == 0 && s1) atomicAdd(counts + jb + 1, s1);
                }
            }
            for (int j = 2; j < nc; j += 2) {  // further extensions, two rows in flight per step
                const bool two = j + 1 < nc;
                int ya, yb;
                if (j + 1 < 32) { ya = __shfl_sync(FULL, yv, j); yb = __shfl_sync(FULL, yv, j + 1); }
                else { ya = ys[jb + j]; yb = two ? ys[jb + j + 1] : ya; }
                load_row<U, TAIL>(r0, base, pq32, ya, ok);
                load_row<U, TAIL>(r1, base, pq32, two ? yb : ya, ok);
                const int s0 = __reduce_add_sync(FULL, popc_rows<U>(c, r0));
                const int s1 = __reduce_add_sync(FULL, popc_rows<U>(c, r1));
                if (lane == 0) {
                    if (s0) atomicAdd(counts + jb + j, s0);
                    if (two && s1) atomicAdd(counts + jb + j + 1, s1);
                }
            }
            pxa = xa; pxb = xb;
        }
        if (alive) {
            // Every resident warp works on the same row chunk (chunk-major order), so they all OR into the same words:
            // L2 serialises atomics on one address.  Read the word first (L2, may be slightly stale — bits are only ever
            // set during the launch, so a stale value can only cause a redundant atomic, never a missing one) and issue
            // the atomic only for bits that are not there yet; after the first few runs of a chunk almost none are.
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (ok[u]) {
                    uint32_t* d = alive + (q0 + u * 32) * 4;
                    const uint4 cur = __ldcg(reinterpret_cast<const uint4*>(d));
                    if (al[u].x & ~cur.x) atomicOr(d + 0, al[u].x);
                    if (al[u].y & ~cur.y) atomicOr(d + 1, al[u].y);
                    if (al[u].z & ~cur.z) atomicOr(d + 2, al[u].z);
                    if (al[u].w & ~cur.w) atomicOr(d + 3, al[u].w);
                }
        }
    }
}

// ------------------------------------------------------------------------------------------
// K4 as the level loop runs it.  Same work decomposition as k_count_groups (warp = run of G
// prefix groups x one 2 KB row chunk, chunk-major, prefix AND re-used along the sorted groups,
// alive rows accumulated), with the per-group bookkeeping moved out of the chunk loop: K5
// writes one descriptor per group (k_group_desc: prefix row, last item, item k-3, how many
// leading items it shares with the group before), so a warp no longer stages and compares every
// group's prefix items for every chunk; it stages prefix items only for the groups whose
// prefix AND has to be rebuilt, and the run's extension ids with one contiguous copy.  The
// counting itself is popc(pm1 & last & ext) — one LOP3 + POPC per word, no intermediate AND.
// (ncu r1h/r1j: the group kernel was instruction-bound with 27 % of its instructions in the
// AND+POPC itself; this version executes ~25 % fewer instructions for the same counts.)
// ------------------------------------------------------------------------------------------
// (also totals, in `loaded_rows`, the item rows K4 will request per row chunk for these groups with runs of `run_len`
// groups: the last prefix item, the extensions, and the prefix rows a group has to rebuild — the measured side of
// fa_stats.count_load_bytes)
__global__ void k_group_desc(const int32_t* __restrict__ sets, int m, const int32_t* __restrict__ gx, const int32_t* __restrict__ goff,
                             const int32_t* __restrict__ n_groups_dev, int run_len, int4* __restrict__ desc,
                             unsigned long long* __restrict__ loaded_rows) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n_groups = *n_groups_dev;  // (the grid covers the host's bound, the size of the level below)
    int loads = 0;
    if (g < n_groups) {
        const int32_t* row = sets + (int64_t)gx[g] * m;
        int d = 0;
        if (g > 0) {
            const int32_t* prow = sets + (int64_t)gx[g - 1] * m;
            while (d < m - 1 && row[d] == prow[d]) ++d;
        }
        desc[g] = make_int4(gx[g], row[m - 1], m >= 2 ? row[m - 2] : -1, d);
        const int dd = (g % run_len == 0) ? 0 : d;  // a run starts from scratch
        loads = 1 + (goff[g + 1] - goff[g]) + (dd >= m - 1 ? 0 : (m >= 3 && dd >= m - 2) ? 1 : m - 1);
    }
    loads = __reduce_add_sync(0xffffffffu, loads);
    if ((threadIdx.x & 31) == 0 && loads) atomicAdd(loaded_rows, (unsigned long long)loads);
}

// L1-hit bandwidth probe (fa_probe_l1_bandwidth): every warp re-reads its own 2 KB with K4's 16-byte read-only loads.
// Each load's address depends on the value the same slot returned in the iteration before (AND-ed with a mask that is
// zero at run time, which the compiler cannot know): nothing can be hoisted or merged, and the 4 independent chains per
// lane x 16 warps per SM keep far more bytes in flight than the L1 latency x bandwidth product needs.
__global__ void __launch_bounds__(256) k_probe_l1(const uint4* __restrict__ src, int iters, uint32_t zero_mask, uint32_t* __restrict__ sink) {
    const uint4* p = src + ((size_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 128 + (threadIdx.x & 31);
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = make_uint4(0u, 0u, 0u, 0u);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = __ldg(p + u * 32 + (v[u].x & zero_mask));
    }
    if ((v[0].x ^ v[1].y ^ v[2].z ^ v[3].w) == 0x9e3779b9u) sink[0] = v[0].x;  // keeps the loads alive; practically never taken
}

#define K4_YS_CAP 256  // extension ids of one run kept in shared memory (a run of 16 groups has ~35; the rest is read from global)
#define COUNT2_SMEM_WORDS(m) (COUNT_G_ * ((m) > 32 ? 64 : 32) + K4_YS_CAP + 32 * COUNT_U_ * 4)

// Bit-sliced adders: POPC runs on the 16-lane XU pipe (a quarter of the LOP3 rate), and 16 POPCs per lane and
// candidate made it the busiest pipe of K4 (ncu r1k: XU 42 %, ALU 41 %).  Eight carry-save adders (2 LOP3 each) turn
// the 16 masked words into 2 vectors of weight 1, 5 of weight 2 and 1 of weight 4: 8 POPCs instead of 16.
__device__ __forceinline__ void csa(uint32_t a, uint32_t b, uint32_t c, uint32_t& sum, uint32_t& carry) {
    sum = a ^ b ^ c;                    // LOP3 0x96
    carry = (a & b) | (c & (a | b));    // LOP3 0xe8 (majority)
}
#ifndef K4_CSA
#define K4_CSA 1
#endif
#ifndef K4_WARPS_SM
#define K4_WARPS_SM (8 * K4_MIN_BLOCKS)  // resident K4 warps per SM the register allocation is made for (16 -> 128 registers)
#endif
template <int U>
__device__ __forceinline__ int popc_rows3(const uint4 (&a)[U], const uint4 (&b)[U], const uint4 (&c)[U]) {
    if constexpr (U == 4 && K4_CSA) {
        uint32_t w[16];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            w[4 * u + 0] = a[u].x & b[u].x & c[u].x; w[4 * u + 1] = a[u].y & b[u].y & c[u].y;
            w[4 * u + 2] = a[u].z & b[u].z & c[u].z; w[4 * u + 3] = a[u].w & b[u].w & c[u].w;
        }
        uint32_t s0, s1, s2, s3, s4, s5, s6, c0, c1, c2, c3, c4, c5, c6, t0, f0;
        csa(w[0], w[1], w[2], s0, c0);
        csa(w[3], w[4], w[5], s1, c1);
        csa(w[6], w[7], w[8], s2, c2);
        csa(w[9], w[10], w[11], s3, c3);
        csa(w[12], w[13], w[14], s4, c4);
        csa(s0, s1, s2, s5, c5);
        csa(s3, s4, w[15], s6, c6);
        csa(c0, c1, c2, t0, f0);
        const int ones = __popc(s5) + __popc(s6);
        const int twos = __popc(t0) + __popc(c3) + __popc(c4) + __popc(c5) + __popc(c6);
        return ones + 2 * twos + 4 * __popc(f0);
    } else if constexpr (U == 2 && K4_CSA) {
        uint32_t w[8];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            w[4 * u + 0] = a[u].x & b[u].x & c[u].x; w[4 * u + 1] = a[u].y & b[u].y & c[u].y;
            w[4 * u + 2] = a[u].z & b[u].z & c[u].z; w[4 * u + 3] = a[u].w & b[u].w & c[u].w;
        }
        uint32_t s0, s1, s2, c0, c1, c2;
        csa(w[0], w[1], w[2], s0, c0);
        csa(w[3], w[4], w[5], s1, c1);
        csa(s0, s1, w[6], s2, c2);
        return __popc(s2) + __popc(w[7]) + 2 * (__popc(c0) + __popc(c1) + __popc(c2));
    } else {
        int s = 0;
#pragma unroll
        for (int u = 0; u < U; ++u)
            s += __popc(a[u].x & b[u].x & c[u].x) + __popc(a[u].y & b[u].y & c[u].y) + __popc(a[u].z & b[u].z & c[u].z) +
                 __popc(a[u].w & b[u].w & c[u].w);
        return s;
    }
}

// One unit per warp (the launch has ceil(units / 8) CTAs), the prefix-slot width a compile-time constant (WIDE: prefixes
// of more than 32 items) and row addresses as base + item * pitch in BYTES (one IMAD.WIDE per row): the per-group part of
// the loop was re-deriving shared-memory bases and 64-bit row addresses from kernel parameters at the 128-register cap
// (ncu r2d source page: ~35 of its ~105 bookkeeping instructions).
template <int U>
__device__ __forceinline__ void load_row_b(uint4 (&dst)[U], const char* __restrict__ base_b, uint32_t pitch_b, int item) {
    const uint4* p = reinterpret_cast<const uint4*>(base_b + (uint64_t)(uint32_t)item * pitch_b);
#pragma unroll
    for (int u = 0; u < U; ++u) dst[u] = __ldg(p + u * 32);
}

template <int U, int G, bool WIDE, int WARPS>
__global__ void __launch_bounds__(32 * WARPS, K4_WARPS_SM / WARPS) k_count_groups2(const uint4* __restrict__ bits, int64_t pitch_q,
                                                                      const int32_t* __restrict__ prev_sets, int m,
                                                                      const int4* __restrict__ gdesc, const int32_t* __restrict__ goff,
                                                                      const int32_t* __restrict__ ys, const int32_t* __restrict__ n_groups_dev,
                                                                      int64_t n_chunks, int32_t* __restrict__ counts,
                                                                      uint32_t* __restrict__ alive, const uint8_t* __restrict__ chunk_shift) {
    static_assert(G <= 30, "run metadata lives in one warp's lanes");
    constexpr int META_X = WIDE ? 64 : 32;
    const int64_t n_groups = *n_groups_dev;  // written by K5 (k_write_groups2); the grid covers the host's bound on it
    extern __shared__ int32_t smem_all[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int32_t* pfx = smem_all + wid * (G * META_X + K4_YS_CAP + 32 * U * 4);  // prefix items of the groups that rebuild
    int32_t* ysm = pfx + G * META_X;                                        // the run's extension ids
    uint4* pm2 = reinterpret_cast<uint4*>(ysm + K4_YS_CAP);                 // AND of the first m-2 prefix rows
    const unsigned FULL = 0xffffffffu;
    const uint32_t pitch_b = (uint32_t)pitch_q * 16u;   // bytes per item row (< 2^32: a row is at most 2^31 bits)
    const int64_t n_runs = (n_groups + G - 1) / G;
    const int64_t unit = (int64_t)blockIdx.x * (blockDim.x >> 5) + wid;
    if (unit < n_runs * n_chunks) {
        const int64_t chunk = unit / n_runs;
        const int64_t run = unit - chunk * n_runs;
        const int64_t q0 = chunk * (32 * U) + lane;
        const char* base = reinterpret_cast<const char*>(bits + q0);
        const int sh = chunk_shift ? chunk_shift[chunk] : 0;  // weighted rows: this chunk counts with weight 2^sh
        const int64_t gbase = run * G;
        const int ngr = (int)min((int64_t)G, n_groups - gbase);
        int4 d = lane < ngr ? __ldg(gdesc + gbase + lane) : make_int4(0, 0, 0, 0);
        if (lane == 0) d.w = 0;  // nothing is carried between units: the run's first group rebuilds its prefix AND
        const int my_goff = lane <= ngr ? goff[gbase + lane] : 0;
        const int jb_run = __shfl_sync(FULL, my_goff, 0);
        const int n_ys = min(__shfl_sync(FULL, my_goff, ngr) - jb_run, K4_YS_CAP);
        __syncwarp();
        for (int i = lane; i < n_ys; i += 32) cp_async4(ysm + i, ys + jb_run + i);
        for (unsigned w = __ballot_sync(FULL, lane < ngr && d.w < m - 2); w; w &= w - 1) {
            const int gi = __ffs(w) - 1;
            const int32_t* xrow = prev_sets + (int64_t)__shfl_sync(FULL, d.x, gi) * m;
            if (lane < m) cp_async4(pfx + gi * META_X + lane, xrow + lane);
            if (lane + 32 < m) cp_async4(pfx + gi * META_X + 32 + lane, xrow + lane + 32);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        uint4 pm1[U], al[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { pm1[u] = make_uint4(~0u, ~0u, ~0u, ~0u); al[u] = make_uint4(0u, 0u, 0u, 0u); }
        for (int gi = 0; gi < ngr; ++gi) {
            const int last = __shfl_sync(FULL, d.y, gi);
            const int depth = __shfl_sync(FULL, d.w, gi);
            const int jb = __shfl_sync(FULL, my_goff, gi);
            const int nc = __shfl_sync(FULL, my_goff, gi + 1) - jb;
            const int off = jb - jb_run;
            const int y0 = nc > 0 ? (off < K4_YS_CAP ? ysm[off] : ys[jb]) : last;
            const int y1 = nc > 1 ? (off + 1 < K4_YS_CAP ? ysm[off + 1] : ys[jb + 1]) : last;
            // (tried: prefetch.global.L1 of the next group's rows -> 34.4 ms instead of 27.6: the L1 pipe is already the busiest unit)
            // the last prefix row and the first two extension rows do not depend on each other: issue them together
            uint4 rl[U], r0[U], r1[U];
            load_row_b<U>(rl, base, pitch_b, last);
            load_row_b<U>(r0, base, pitch_b, y0);
            if (nc > 1) load_row_b<U>(r1, base, pitch_b, y1);  // about half of the groups have a single extension
            if (depth < m - 1) {  // a new sibling family: pm1 = AND of the first m-1 prefix rows
                if (m >= 3 && depth >= m - 2) {  // only item m-2 changed: the AND of the first m-2 is cached
#pragma unroll
                    for (int u = 0; u < U; ++u) pm1[u] = pm2[u * 32 + lane];
                } else {
#pragma unroll
                    for (int u = 0; u < U; ++u) pm1[u] = make_uint4(~0u, ~0u, ~0u, ~0u);
                    const int32_t* items = pfx + gi * META_X;
                    int t = 0;
                    for (; t + 1 < m - 2; t += 2) {  // two rows in flight per step
                        uint4 ra[U], rb[U];
                        load_row_b<U>(ra, base, pitch_b, items[t]);
                        load_row_b<U>(rb, base, pitch_b, items[t + 1]);
#pragma unroll
                        for (int u = 0; u < U; ++u) {
                            pm1[u].x &= ra[u].x & rb[u].x; pm1[u].y &= ra[u].y & rb[u].y;
                            pm1[u].z &= ra[u].z & rb[u].z; pm1[u].w &= ra[u].w & rb[u].w;
                        }
                    }
                    if (t < m - 2) {
                        uint4 ra[U];
                        load_row_b<U>(ra, base, pitch_b, items[t]);
#pragma unroll
                        for (int u = 0; u < U; ++u) { pm1[u].x &= ra[u].x; pm1[u].y &= ra[u].y; pm1[u].z &= ra[u].z; pm1[u].w &= ra[u].w; }
                    }
                    if (m >= 3) {
#pragma unroll
                        for (int u = 0; u < U; ++u) pm2[u * 32 + lane] = pm1[u];
                    }
                }
                {  // item m-2 closes pm1 (m >= 2: the level loop starts at k = 3)
                    uint4 ra[U];
                    load_row_b<U>(ra, base, pitch_b, __shfl_sync(FULL, d.z, gi));
#pragma unroll
                    for (int u = 0; u < U; ++u) { pm1[u].x &= ra[u].x; pm1[u].y &= ra[u].y; pm1[u].z &= ra[u].z; pm1[u].w &= ra[u].w; }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {  // rows in this group's prefix tidset stay alive
                al[u].x |= pm1[u].x & rl[u].x; al[u].y |= pm1[u].y & rl[u].y;
                al[u].z |= pm1[u].z & rl[u].z; al[u].w |= pm1[u].w & rl[u].w;
            }
            {
                const int s0 = __reduce_add_sync(FULL, popc_rows3<U>(pm1, rl, r0));
                if (lane == 0 && nc > 0 && s0) atomicAdd(counts + jb, s0 << sh);
                if (nc > 1) {
                    const int s1 = __reduce_add_sync(FULL, popc_rows3<U>(pm1, rl, r1));
                    if (lane == 0 && s1) atomicAdd(counts + jb + 1, s1 << sh);
                }
            }
            for (int j = 2; j < nc; j += 2) {  // further extensions, two rows in flight per step
                const bool two = j + 1 < nc;
                const int ya = off + j < K4_YS_CAP ? ysm[off + j] : ys[jb + j];
                const int yb = !two ? ya : (off + j + 1 < K4_YS_CAP ? ysm[off + j + 1] : ys[jb + j + 1]);
                load_row_b<U>(r0, base, pitch_b, ya);
                load_row_b<U>(r1, base, pitch_b, yb);
                const int s0 = __reduce_add_sync(FULL, popc_rows3<U>(pm1, rl, r0));
                const int s1 = __reduce_add_sync(FULL, popc_rows3<U>(pm1, rl, r1));
                if (lane == 0) {
                    if (s0) atomicAdd(counts + jb + j, s0 << sh);
                    if (two && s1) atomicAdd(counts + jb + j + 1, s1 << sh);
                }
            }
        }
        if (alive) {  // see k_count_groups: read first, atomics only for bits that are not there yet
#pragma unroll
            for (int u = 0; u < U; ++u) {
                uint32_t* dst = alive + (q0 + u * 32) * 4;
                const uint4 cur = __ldcg(reinterpret_cast<const uint4*>(dst));
                if (al[u].x & ~cur.x) atomicOr(dst + 0, al[u].x);
                if (al[u].y & ~cur.y) atomicOr(dst + 1, al[u].y);
                if (al[u].z & ~cur.z) atomicOr(dst + 2, al[u].z);
                if (al[u].w & ~cur.w) atomicOr(dst + 3, al[u].w);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// K4 on sibling tiles.  The groups of one sibling family P+{a_0}, P+{a_1}, ... share the parent prefix P (k-2 items)
// and draw their extensions from the same set: the last items of the family's later members.  A TILE is up to four
// groups of one family (those whose prefix rows fall into the same aligned block of four family rows).  Per row chunk a
// warp builds t_i = AND(P) & row[a_i] for the tile's groups once, keeps them in registers, and then walks the tile's
// EXTENSION STREAM: the union of the groups' extension lists in ascending item order, one 32-bit entry per distinct
// item = item | (4-bit mask of the groups that have it) << 24.  Each entry costs one row load however many of the four
// groups count against it, where the group-wise kernel loads the row once per candidate.  Candidate indices need no
// table: group i's candidates are consecutive from its offset, in the order the stream presents them.  The stream and
// three 16-byte descriptors per tile are compiled ONCE per level by k_tile_build; the chunk loop only interprets them.
//   descA = {prefix item k-3 (the last item of P), n_entries | n_groups << 28, depth, stream offset}
//   descB = {a_0, a_1, a_2, a_3}            descC = {candidate base of group 1, 2, 3, prefix row of group 0}
// depth = leading items of P shared with the tile before (>= k-2: same family, AND(P) is kept; k-3: only P's last item
// changed, the AND of the first k-3 items is cached in shared memory).  A group's stream region starts at its first
// candidate's index, so the stream needs no offsets of its own and is never longer than the candidate array.
// ------------------------------------------------------------------------------------------
__global__ void k_tile_flags(const int32_t* __restrict__ sub, const int32_t* __restrict__ pchild, int m,
                             const int32_t* __restrict__ gx, int64_t n_groups, int32_t* __restrict__ flag) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_groups) return;
    const int x = gx[g];
    const int par = sub[(int64_t)x * m + (m - 1)];
    int f = 1;
    if (g > 0) {
        const int px = gx[g - 1];
        const int ppar = sub[(int64_t)px * m + (m - 1)];
        f = par != ppar || ((x - pchild[par]) >> 2) != ((px - pchild[par]) >> 2);
    }
    flag[g] = f;
}

__global__ void __launch_bounds__(128) k_tile_build(const int32_t* __restrict__ sets, int m, const int32_t* __restrict__ gx,
                                                    const int32_t* __restrict__ goff, const int32_t* __restrict__ ys,
                                                    const int32_t* __restrict__ flag, const int32_t* __restrict__ tpos,
                                                    int64_t n_groups, int run_len, int4* __restrict__ descA, int4* __restrict__ descB,
                                                    int4* __restrict__ descC, int32_t* __restrict__ stream,
                                                    int32_t* __restrict__ n_tiles_out, unsigned long long* __restrict__ loaded_rows) {
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int loads = 0;
    if (g < n_groups && flag[g]) {
        const int t = tpos[g];
        int ngr = 1;
        while (ngr < 4 && g + ngr < n_groups && !flag[g + ngr]) ++ngr;
        const int32_t* row = sets + (int64_t)gx[g] * m;
        int d = 0;  // leading items of the parent prefix (the first m-1 items) shared with the tile before
        if (g > 0) {
            const int32_t* prow = sets + (int64_t)gx[g - 1] * m;
            while (d < m - 1 && row[d] == prow[d]) ++d;
        }
        int a[4] = {-1, -1, -1, -1}, p[4] = {0, 0, 0, 0}, e[4] = {0, 0, 0, 0};
        for (int i = 0; i < ngr; ++i) {
            a[i] = sets[(int64_t)gx[g + i] * m + (m - 1)];
            p[i] = goff[g + i];
            e[i] = goff[g + i + 1];
        }
        const int soff = p[0];
        const int cb1 = p[1], cb2 = p[2], cb3 = p[3];
        int n = 0;
        for (;;) {  // 4-way merge of the ascending extension lists
            int y = 0x7fffffff;
            for (int i = 0; i < ngr; ++i)
                if (p[i] < e[i]) y = min(y, ys[p[i]]);
            if (y == 0x7fffffff) break;
            int mask = 0;
            for (int i = 0; i < ngr; ++i)
                if (p[i] < e[i] && ys[p[i]] == y) { mask |= 1 << i; ++p[i]; }
            stream[soff + n++] = y | (mask << 24);
        }
        descA[t] = make_int4(m >= 2 ? row[m - 2] : -1, n | (ngr << 28), d, soff);
        descB[t] = make_int4(a[0], a[1], a[2], a[3]);
        descC[t] = make_int4(cb1, cb2, cb3, gx[g]);
        const int dd = (t % run_len == 0) ? 0 : d;  // a run starts from scratch
        loads = ngr + n + (dd >= m - 1 ? 0 : (m >= 3 && dd >= m - 2) ? 1 : m - 1);
    }
    if (g == n_groups - 1) *n_tiles_out = tpos[g] + flag[g];
    loads = __reduce_add_sync(0xffffffffu, loads);
    if ((threadIdx.x & 31) == 0 && loads) atomicAdd(loaded_rows, (unsigned long long)loads);
}

#ifndef COUNT_GT
#define COUNT_GT 12   // tiles per warp run
#endif
#define TILE_SMEM_WORDS(m) (3 * 4 * COUNT_GT + K4_YS_CAP + COUNT_GT * ((m) > 32 ? 64 : 32) + 2 * 32 * COUNT_U_ * 4)

template <int U>
__device__ __forceinline__ int popc_rows2(const uint4 (&a)[U], const uint4 (&b)[U]) {
    if constexpr (U == 4 && K4_CSA) {
        uint32_t w[16];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            w[4 * u + 0] = a[u].x & b[u].x; w[4 * u + 1] = a[u].y & b[u].y;
            w[4 * u + 2] = a[u].z & b[u].z; w[4 * u + 3] = a[u].w & b[u].w;
        }
        uint32_t s0, s1, s2, s3, s4, s5, s6, c0, c1, c2, c3, c4, c5, c6, t0, f0;
        csa(w[0], w[1], w[2], s0, c0);
        csa(w[3], w[4], w[5], s1, c1);
        csa(w[6], w[7], w[8], s2, c2);
        csa(w[9], w[10], w[11], s3, c3);
        csa(w[12], w[13], w[14], s4, c4);
        csa(s0, s1, s2, s5, c5);
        csa(s3, s4, w[15], s6, c6);
        csa(c0, c1, c2, t0, f0);
        const int ones = __popc(s5) + __popc(s6);
        const int twos = __popc(t0) + __popc(c3) + __popc(c4) + __popc(c5) + __popc(c6);
        return ones + 2 * twos + 4 * __popc(f0);
    } else {
        int s = 0;
#pragma unroll
        for (int u = 0; u < U; ++u) s += popc4(a[u], b[u]);
        return s;
    }
}

template <int U>
__device__ __forceinline__ void and_rows(uint4 (&d)[U], const uint4 (&a)[U]) {
#pragma unroll
    for (int u = 0; u < U; ++u) { d[u].x &= a[u].x; d[u].y &= a[u].y; d[u].z &= a[u].z; d[u].w &= a[u].w; }
}

template <int U, int GT>
__global__ void __launch_bounds__(256, K4_MIN_BLOCKS) k_count_tiles(const uint4* __restrict__ bits, int64_t pitch_q,
                                                                    const int32_t* __restrict__ prev_sets, int m,
                                                                    const int4* __restrict__ descA, const int4* __restrict__ descB,
                                                                    const int4* __restrict__ descC, const int32_t* __restrict__ stream,
                                                                    const int32_t* __restrict__ n_tiles_ptr, int64_t n_chunks,
                                                                    int32_t* __restrict__ counts, uint32_t* __restrict__ alive,
                                                                    const uint8_t* __restrict__ chunk_shift,
                                                                    unsigned long long* __restrict__ next_unit) {
    const int META_X = m > 32 ? 64 : 32;
    extern __shared__ int32_t smem_all[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int32_t* base_w = smem_all + wid * (3 * 4 * GT + K4_YS_CAP + GT * META_X + 2 * 32 * U * 4);
    int4* sA = reinterpret_cast<int4*>(base_w);          // the run's tile descriptors
    int4* sB = sA + GT;
    int4* sC = sB + GT;
    int32_t* ssm = base_w + 3 * 4 * GT;                  // the run's extension stream
    int32_t* pfx = ssm + K4_YS_CAP;                      // prefix items of the tiles that rebuild from scratch
    uint4* pm2 = reinterpret_cast<uint4*>(pfx + GT * META_X);  // AND of the first k-3 prefix rows
    uint4* pps = pm2 + 32 * U;                           // AND of the parent prefix (k-2 rows)
    const unsigned FULL = 0xffffffffu;
    const bool none[U] = {};
    const uint32_t pq32 = (uint32_t)pitch_q;
    const int64_t n_tiles = *n_tiles_ptr;
    const int64_t n_runs = (n_tiles + GT - 1) / GT;
    const int64_t total = n_runs * n_chunks;
    // Units (row chunk, run of GT tiles) are handed out in order from one counter: at any moment the resident warps work
    // on neighbouring units, i.e. on the same 2 KB column of the item rows (chunk-major sweep: that column is fetched from
    // HBM once per level and lives in L1/L2 while it is in use), whatever the grid size, and a long unit delays nobody.
    for (;;) {
        unsigned long long u64 = 0;
        if (lane == 0) u64 = atomicAdd(next_unit, 1ull);
        const int64_t unit = (int64_t)__shfl_sync(FULL, u64, 0);
        if (unit >= total) break;
        const int64_t chunk = unit / n_runs;
        const int64_t run = unit - chunk * n_runs;
        const int64_t q0 = chunk * (32 * U) + lane;
        const uint4* base = bits + q0;
        const int sh = chunk_shift ? chunk_shift[chunk] : 0;  // weighted rows: this chunk counts with weight 2^sh
        const int64_t tbase = run * GT;
        const int ntr = (int)min((int64_t)GT, n_tiles - tbase);
        int4 dA = make_int4(0, 0, 0, 0), dB = dA, dC = dA;
        if (lane < ntr) { dA = __ldg(descA + tbase + lane); dB = __ldg(descB + tbase + lane); dC = __ldg(descC + tbase + lane); }
        if (lane == 0) dA.z = 0;  // nothing is carried between units: the run's first tile rebuilds its prefix AND
        __syncwarp();
        if (lane < ntr) { sA[lane] = dA; sB[lane] = dB; sC[lane] = dC; }
        const int soff_run = __shfl_sync(FULL, dA.w, 0);
        const int n_str = min(__shfl_sync(FULL, dA.w + (dA.y & 0x0fffffff), ntr - 1) - soff_run, K4_YS_CAP);
        for (int i = lane; i < n_str; i += 32) cp_async4(ssm + i, stream + soff_run + i);
        for (unsigned w = __ballot_sync(FULL, lane < ntr && dA.z < m - 2); w; w &= w - 1) {
            const int ti = __ffs(w) - 1;
            const int32_t* xrow = prev_sets + (int64_t)__shfl_sync(FULL, dC.w, ti) * m;
            if (lane < m) cp_async4(pfx + ti * META_X + lane, xrow + lane);
            if (lane + 32 < m) cp_async4(pfx + ti * META_X + 32 + lane, xrow + lane + 32);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        uint4 al[U];
#pragma unroll
        for (int u = 0; u < U; ++u) al[u] = make_uint4(0u, 0u, 0u, 0u);
        for (int ti = 0; ti < ntr; ++ti) {
            const int4 A = sA[ti], B = sB[ti], C = sC[ti];
            // (the descriptors are the same in every lane; passing them through a warp reduction tells the compiler so, and
            // the branches on them become uniform branches without reconvergence bookkeeping)
            const unsigned ay = __reduce_or_sync(FULL, (unsigned)A.y);
            const int ngr = ay >> 28, n_ent = ay & 0x0fffffff, depth = __reduce_or_sync(FULL, (unsigned)A.z);
            const int so = A.w - soff_run;
            // AND of the parent prefix first (usually a cached AND and one row), then the tile's own rows against it: holding
            // the four rows while the prefix is rebuilt would not fit the 128-register budget of 2 CTAs per SM
            uint4 t0[U], t1[U], t2[U], t3[U];
            {
                uint4 pp[U];
                if (depth >= m - 1) {  // same family as the tile before: the parent prefix AND is kept
#pragma unroll
                    for (int u = 0; u < U; ++u) pp[u] = pps[u * 32 + lane];
                } else {
                    if (m >= 3 && depth >= m - 2) {  // only the prefix's last item changed
#pragma unroll
                        for (int u = 0; u < U; ++u) pp[u] = pm2[u * 32 + lane];
                    } else {
#pragma unroll
                        for (int u = 0; u < U; ++u) pp[u] = make_uint4(~0u, ~0u, ~0u, ~0u);
                        const int32_t* items = pfx + ti * META_X;
                        int t = 0;
                        for (; t + 1 < m - 2; t += 2) {  // two rows in flight per step
                            uint4 ra[U], rb[U];
                            load_row<U, false>(ra, base, pq32, items[t], none);
                            load_row<U, false>(rb, base, pq32, items[t + 1], none);
                            and_rows<U>(pp, ra);
                            and_rows<U>(pp, rb);
                        }
                        if (t < m - 2) {
                            uint4 ra[U];
                            load_row<U, false>(ra, base, pq32, items[t], none);
                            and_rows<U>(pp, ra);
                        }
                        if (m >= 3) {
#pragma unroll
                            for (int u = 0; u < U; ++u) pm2[u * 32 + lane] = pp[u];
                        }
                    }
                    if (m >= 2) {  // item k-3 closes the parent prefix (m == 2: it is the whole of it)
                        uint4 ra[U];
                        load_row<U, false>(ra, base, pq32, A.x, none);
                        and_rows<U>(pp, ra);
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) pps[u * 32 + lane] = pp[u];
                }
                load_row<U, false>(t0, base, pq32, B.x, none);
                if (ngr > 1) load_row<U, false>(t1, base, pq32, B.y, none);
                if (ngr > 2) load_row<U, false>(t2, base, pq32, B.z, none);
                if (ngr > 3) load_row<U, false>(t3, base, pq32, B.w, none);
                and_rows<U>(t0, pp);
                if (ngr > 1) and_rows<U>(t1, pp);
                if (ngr > 2) and_rows<U>(t2, pp);
                if (ngr > 3) and_rows<U>(t3, pp);
            }
            // first stream row in flight while the alive rows are accumulated
            uint4 rs[U];
            int e = so < K4_YS_CAP ? ssm[so] : stream[A.w];
            load_row<U, false>(rs, base, pq32, e & 0xffffff, none);
#pragma unroll
            for (int u = 0; u < U; ++u) {  // rows in a group prefix's tidset stay alive
                al[u].x |= t0[u].x; al[u].y |= t0[u].y; al[u].z |= t0[u].z; al[u].w |= t0[u].w;
            }
            if (ngr > 1) {
#pragma unroll
                for (int u = 0; u < U; ++u) { al[u].x |= t1[u].x; al[u].y |= t1[u].y; al[u].z |= t1[u].z; al[u].w |= t1[u].w; }
            }
            if (ngr > 2) {
#pragma unroll
                for (int u = 0; u < U; ++u) { al[u].x |= t2[u].x; al[u].y |= t2[u].y; al[u].z |= t2[u].z; al[u].w |= t2[u].w; }
            }
            if (ngr > 3) {
#pragma unroll
                for (int u = 0; u < U; ++u) { al[u].x |= t3[u].x; al[u].y |= t3[u].y; al[u].z |= t3[u].z; al[u].w |= t3[u].w; }
            }
            int c0 = A.w, c1 = C.x, c2 = C.y, c3 = C.z;
            auto count_entry = [&](const uint4 (&r)[U], int mask) {
                if (mask & 1) {
                    const int v = __reduce_add_sync(FULL, popc_rows2<U>(t0, r));
                    if (lane == 0 && v) atomicAdd(counts + c0, v << sh);
                    ++c0;
                }
                if (mask & 2) {
                    const int v = __reduce_add_sync(FULL, popc_rows2<U>(t1, r));
                    if (lane == 0 && v) atomicAdd(counts + c1, v << sh);
                    ++c1;
                }
                if (mask & 4) {
                    const int v = __reduce_add_sync(FULL, popc_rows2<U>(t2, r));
                    if (lane == 0 && v) atomicAdd(counts + c2, v << sh);
                    ++c2;
                }
                if (mask & 8) {
                    const int v = __reduce_add_sync(FULL, popc_rows2<U>(t3, r));
                    if (lane == 0 && v) atomicAdd(counts + c3, v << sh);
                    ++c3;
                }
            };
            auto entry = [&](int s) { return so + s < K4_YS_CAP ? ssm[so + s] : stream[A.w + s]; };
            for (int s = 0; s < n_ent; ++s) {
                count_entry(rs, __reduce_or_sync(FULL, (unsigned)e >> 24));
                if (s + 1 < n_ent) {  // (one row buffer: a second one in flight does not fit next to four prefix tidsets)
                    e = entry(s + 1);
                    load_row<U, false>(rs, base, pq32, e & 0xffffff, none);
                }
            }
        }
        if (alive) {  // see k_count_groups: read first, atomics only for bits that are not there yet
#pragma unroll
            for (int u = 0; u < U; ++u) {
                uint32_t* dst = alive + (q0 + u * 32) * 4;
                const uint4 cur = __ldcg(reinterpret_cast<const uint4*>(dst));
                if (al[u].x & ~cur.x) atomicOr(dst + 0, al[u].x);
                if (al[u].y & ~cur.y) atomicOr(dst + 1, al[u].y);
                if (al[u].z & ~cur.z) atomicOr(dst + 2, al[u].z);
                if (al[u].w & ~cur.w) atomicOr(dst + 3, al[u].w);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// Transaction trimming between levels (exact; the reference keeps every distinct row for the
// whole run).  While counting level k, K4 ORs the tidset of every group prefix into `alive`.
// Any deeper candidate contains one of those prefixes, so a row outside `alive` can never be
// counted again: such rows are squeezed out of the bitmaps of the items that still occur in
// frequent k-itemsets, and deeper levels stream shorter rows.  Supports are unchanged.
// ------------------------------------------------------------------------------------------
// (a few big CTAs, each marking into a bitmap of the F ranks in shared memory and storing one flag per rank it saw: the
// level's n * k entries name only F different ranks, and one blind global store per entry — 1.8 M stores to 1000 words at
// level 9 of configs[1] — took 31 us per level; testing the global flag first was slower still, one hot L2 line read by
// every SM)
__global__ void __launch_bounds__(1024) k_mark_active(const int32_t* __restrict__ sets, int64_t n_entries, int F,
                                                      int32_t* __restrict__ active) {
    extern __shared__ uint32_t seen[];
    const int words = (F + 31) >> 5;
    for (int w = threadIdx.x; w < words; w += blockDim.x) seen[w] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_entries; i += (int64_t)gridDim.x * blockDim.x) {
        const int x = __ldg(sets + i);
        const uint32_t bit = 1u << (x & 31);
        if (!(seen[x >> 5] & bit)) atomicOr(&seen[x >> 5], bit);
    }
    __syncthreads();
    for (int x = threadIdx.x; x < F; x += blockDim.x)
        if (seen[x >> 5] >> (x & 31) & 1) active[x] = 1;
}
__global__ void k_popc_words(const uint32_t* __restrict__ w, int64_t n, int32_t* __restrict__ cnt) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) cnt[i] = __popc(w[i]);
}

// One thread per 32-row word: the bit-compress network for this word's alive mask is built once
// (Hacker's Delight 7-4) and applied to the word of every active item; results are OR-ed into
// the shorter destination rows (zeroed beforehand) at the word's bit offset.
__global__ void __launch_bounds__(256) k_compact_rows(const uint32_t* __restrict__ src, int64_t src_pitch,
                                                      const uint32_t* __restrict__ alive, const int32_t* __restrict__ alive_off,
                                                      const int32_t* __restrict__ active, int F, uint32_t* __restrict__ dst,
                                                      int64_t dst_pitch) {
    extern __shared__ int32_t act_list[];
    __shared__ int n_act;
    if (threadIdx.x == 0) n_act = 0;
    __syncthreads();
    // blockIdx.y takes a slice of the item RANKS (each CTA lists the active items of its own slice; the list order is
    // arbitrary, the slice is not): with a thread per word alone, the short bitmaps of the deep levels gave a grid of a
    // few dozen CTAs, each walking all ~1000 items serially (0.11 ms per trim at configs[1])
    const int per = (F + gridDim.y - 1) / gridDim.y;
    const int i_lo = blockIdx.y * per, i_hi = min(F, i_lo + per);
    for (int i = i_lo + threadIdx.x; i < i_hi; i += blockDim.x)
        if (active[i]) act_list[atomicAdd(&n_act, 1)] = i;
    __syncthreads();
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= src_pitch) return;
    const uint32_t a = alive[w];
    if (!a) return;
    uint32_t mv[5];
    {
        uint32_t m = a, mk = ~m << 1;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            uint32_t mp = mk ^ (mk << 1);
            mp ^= mp << 2; mp ^= mp << 4; mp ^= mp << 8; mp ^= mp << 16;
            mv[i] = mp & m;
            m = (m ^ mv[i]) | (mv[i] >> (1 << i));
            mk &= ~mp;
        }
    }
    const int off = alive_off[w];
    const int sh = off & 31;
    const int64_t dw = off >> 5;
    const int na = n_act;
    for (int t = 0; t < na; ++t) {
        const int item = act_list[t];
        uint32_t x = __ldg(src + (int64_t)item * src_pitch + w) & a;
        if (!x) continue;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const uint32_t tt = x & mv[i];
            x = (x ^ tt) | (tt >> (1 << i));
        }
        uint32_t* d = dst + (int64_t)item * dst_pitch + dw;
        atomicOr(d, x << sh);
        if (sh && (x >> (32 - sh))) atomicOr(d + 1, x >> (32 - sh));
    }
}

// frequent candidates -> rows of the next level (order preserved, so the level stays sorted)
__global__ void k_compact_level(const int32_t* __restrict__ prev_sets, int m, const int32_t* __restrict__ cx,
                                const int32_t* __restrict__ ys, const int32_t* __restrict__ csub,
                                const int32_t* __restrict__ counts, const int32_t* __restrict__ flag,
                                const int32_t* __restrict__ pos, int64_t n_cands, int32_t* __restrict__ sets,
                                int32_t* __restrict__ sub, int32_t* __restrict__ out_counts, int32_t* __restrict__ n_out) {
    // one thread per (candidate, position): neighbouring threads read and write neighbouring words of the row arrays
    const int k = m + 1;
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_cands * k) return;
    const int64_t j = e / k;
    const int t = (int)(e - j * k);
    if (e == n_cands * k - 1) *n_out = pos[j] + flag[j];  // size of the new level, for the kernels queued behind this one
    if (!flag[j]) return;
    const int64_t d = (int64_t)pos[j] * k + t;
    sets[d] = t < m ? prev_sets[(int64_t)cx[j] * m + t] : ys[j];
    sub[d] = csub[e];
    if (t == 0) out_counts[pos[j]] = counts[j];
}

// ------------------------------------------------------------------------------------------
// Resume at the rule stage: levels installed from the host get the same
// `sub` index the mining kernels produce.  One thread per (row s, dropped position j): binary
// search of S - {S[j]} in the sorted rows of the level below.  Also checks what the rest of
// the pipeline assumes: items ascending inside a row, rows strictly ascending, ranks in range.
// err bit 0: malformed level, bit 1: a subset is missing one level down (not downward closed).
// ------------------------------------------------------------------------------------------
__global__ void k_find_subsets(const int32_t* __restrict__ sets, int64_t n, int k, const int32_t* __restrict__ prev,
                               int64_t n_prev, int n_ranks, int32_t* __restrict__ sub, int32_t* __restrict__ err) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n * k) return;
    const int64_t s = r / k;
    const int j = (int)(r - s * k);
    const int32_t* row = sets + s * k;
    if (j == 0) {  // one thread per row validates it
        bool bad = row[0] < 0 || row[k - 1] >= n_ranks;
        for (int t = 1; t < k; ++t) bad |= row[t] <= row[t - 1];
        if (s > 0) {
            const int32_t* before = row - k;
            int t = 0;
            while (t < k && before[t] == row[t]) ++t;
            bad |= t == k || before[t] > row[t];
        }
        if (bad) atomicOr(err, 1);
    }
    const int m = k - 1;
    if (m == 1) {  // level 1 is the identity: row z is {z}
        const int x = row[j == 0 ? 1 : 0];
        sub[r] = x;
        if (x < 0 || x >= n_prev) atomicOr(err, 2);
        return;
    }
    int64_t lo = 0, hi = n_prev;  // first row of prev that is >= the subset
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        const int32_t* p = prev + mid * m;
        int cmp = 0;
        for (int t = 0; t < m && cmp == 0; ++t) {
            const int x = row[t + (t >= j)];
            cmp = (p[t] < x) ? -1 : (p[t] > x);
        }
        if (cmp < 0) lo = mid + 1; else hi = mid;
    }
    bool found = lo < n_prev;
    if (found) {
        const int32_t* p = prev + lo * m;
        for (int t = 0; t < m; ++t) found &= p[t] == row[t + (t >= j)];
    }
    sub[r] = found ? (int32_t)lo : -1;
    if (!found) atomicOr(err, 2);
}

// child[z] = first row of the level above whose parent (sub[.][k-1]) is >= z, z in 0..n_parents
__global__ void k_child_index(const int32_t* __restrict__ sub, int64_t n, const int32_t* __restrict__ n_dev, int k, int n_parents,
                              int32_t* __restrict__ child) {
    const int z = blockIdx.x * blockDim.x + threadIdx.x;
    if (z > n_parents) return;
    int64_t lo = 0, hi = n_dev ? (int64_t)*n_dev : n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (sub[mid * k + (k - 1)] < z) lo = mid + 1; else hi = mid;
    }
    child[z] = (int32_t)lo;
}
__global__ void k_iota(int32_t* p, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}

// ------------------------------------------------------------------------------------------
// K6 — rule generation (AssociationRules.scala:126-145).  Rule id = base_k + s*k + j is the
// rule (S - {S[j]} -> S[j]) of the s-th frequent k-itemset.  One IEEE double division each.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gen_rules(const int32_t* __restrict__ counts, const int32_t* __restrict__ sub,
                                                   int64_t n, int k, const int32_t* __restrict__ pcounts,
                                                   double* __restrict__ conf) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n * k) return;
    // sub[r] indexes S - {S[j]} one level down (downward closure guarantees it exists)
    conf[r] = (double)counts[r / k] / (double)pcounts[sub[r]];
}

// ------------------------------------------------------------------------------------------
// K7 — level-wise prune (AssociationRules.scala:152-179): keep (A -> c) iff every
// (A - {a} -> c) was kept one level down with strictly smaller confidence.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_prune_rules(int64_t n, int k, const int32_t* __restrict__ sub,
                                                     const double* __restrict__ conf, const double* __restrict__ pconf,
                                                     const uint8_t* __restrict__ pkept, uint8_t* __restrict__ kept,
                                                     unsigned long long* __restrict__ kept_count) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool ok = false;
    if (r < n * k) {
        const int64_t s = r / k;
        const int j = (int)(r - s * k);
        const double c = conf[r];
        ok = true;
        for (int i = 0; ok && i < k; ++i) {
            if (i == j) continue;
            const int64_t pr = (int64_t)sub[s * k + i] * (k - 1) + (j - (j > i ? 1 : 0));
            ok = pkept[pr] && pconf[pr] < c;
        }
        kept[r] = ok;
    }
    // survivors of this level, counted on the device (the reference's "After cut level" printout, :177)
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(kept_count, (unsigned long long)__popc(m));
}

__device__ __forceinline__ void decode_rule(const LevelTab& T, int64_t rid, int& k, int64_t& s, int& j) {
    int kk = 2;
    while (kk < T.max_k && rid >= T.base[kk + 1]) ++kk;
    const int64_t local = rid - T.base[kk];
    k = kk;
    s = local / kk;
    j = (int)(local - s * kk);
}

// K8 keys (AssociationRules.scala:116-120): secondary = consequent's item id as Int, ascending
__global__ void k_rule_key_item(LevelTab T, const int32_t* __restrict__ list, int64_t n,
                                const int32_t* __restrict__ item_int, uint32_t* __restrict__ key) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int k, j; int64_t s;
    decode_rule(T, list[p], k, s, j);
    const int cons = T.sets[k][s * k + j];
    key[p] = (uint32_t)item_int[cons] ^ 0x80000000u;  // signed -> unsigned order
}
// primary = confidence descending: positive doubles order like their bit patterns
__global__ void k_rule_key_conf(const int32_t* __restrict__ list, int64_t n, const double* __restrict__ conf,
                                uint64_t* __restrict__ key) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) key[p] = ~(uint64_t)__double_as_longlong(conf[list[p]]);
}

// K9 index: kept rules grouped by their antecedent's trie node, priority order inside each group
__global__ void k_rule_key_ant(LevelTab T, const int32_t* __restrict__ order, int64_t n, uint32_t* __restrict__ key,
                               int32_t* __restrict__ prio) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int k, j; int64_t s;
    decode_rule(T, order[p], k, s, j);
    key[p] = (uint32_t)(T.nodebase[k - 1] + T.sub[k][s * k + j]);  // node of S - {S[j]} one level down
    prio[p] = (int32_t)p;
}
// records in node order: (priority, consequent)
__global__ void k_rule_records(LevelTab T, const int32_t* __restrict__ order, const int32_t* __restrict__ lp, int64_t n,
                               int2* __restrict__ rec) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = lp[i];
    int k, j; int64_t s;
    decode_rule(T, order[p], k, s, j);
    rec[i] = make_int2(p, T.sets[k][s * k + j]);
}
__global__ void k_lower_bounds(const uint32_t* __restrict__ keys, int64_t n, int64_t n_keys, int32_t* __restrict__ lo) {
    const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (a > n_keys) return;
    int64_t l = 0, h = n;
    while (l < h) {
        const int64_t mid = (l + h) >> 1;
        if (keys[mid] < (uint32_t)a) l = mid + 1; else h = mid;
    }
    lo[a] = (int32_t)l;
}

// Digest of the kept, sorted rules (fa_rules_digest); the CPU checker under tests/ restates this arithmetic.
// d[0] += h(rule), d[1] ^= h(rule) over the rule multiset; d[2] += g(position, confidence bits, consequent item id).
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
    return x;
}
__global__ void __launch_bounds__(256) k_rule_digest(LevelTab T, const int32_t* __restrict__ order, int64_t n,
                                                     const double* __restrict__ conf, const int32_t* __restrict__ item_int,
                                                     unsigned long long* __restrict__ d) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t hsh = 0, ord = 0;
    if (p < n) {
        const int rid = order[p];
        int k, j; int64_t s;
        decode_rule(T, rid, k, s, j);
        const int32_t* S = T.sets[k] + s * k;
        const uint64_t bits = (uint64_t)__double_as_longlong(conf[rid]);
        hsh = mix64((uint64_t)(uint32_t)S[j] + 0x9E3779B97F4A7C15ull);
        for (int t = 0; t < k; ++t)
            if (t != j) hsh = mix64(hsh ^ ((uint64_t)(uint32_t)(S[t] + 1) * 0xD6E8FEB86659FD93ull));
        hsh = mix64(hsh ^ bits);
        ord = mix64(mix64((uint64_t)p + 1) ^ bits) ^ mix64(((uint64_t)p << 32) ^ (uint64_t)(uint32_t)item_int[S[j]]);
    }
    uint64_t sum = hsh, x = hsh;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        x ^= __shfl_xor_sync(0xffffffffu, x, o);
        ord += __shfl_xor_sync(0xffffffffu, ord, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(d, (unsigned long long)sum);
        atomicXor(d + 1, (unsigned long long)x);
        atomicAdd(d + 2, (unsigned long long)ord);
    }
}

// ------------------------------------------------------------------------------------------
// K9 — per-user first match (AssociationRules.scala:80-106).  A rule A -> c can fire for basket
// u only if A is a frequent itemset inside u, so one warp walks the itemset trie restricted to
// u: level-1 nodes are u's items, and a node's children (contiguous rows one level up, see K5)
// are tested 32 at a time against u's F-bit mask in shared memory.  Every visited node looks up
// the rules that have it as antecedent (grouped by node, best priority first) and takes the
// first whose consequent is outside u; the warp keeps the minimum priority.  Work per basket is
// the number of frequent subsets of the basket, not the number of rules.  Pending child ranges
// live on a per-warp stack (shared memory, spilling to a global scratch region).
// L2/latency-bound.
// ------------------------------------------------------------------------------------------
#define MATCH_STACK_SMEM 160   // ranges per warp kept in shared memory
#define MATCH_WARPS 8

__global__ void __launch_bounds__(32 * MATCH_WARPS) k_recommend(LevelTab T, const int32_t* __restrict__ aoff,
                                                                const int2* __restrict__ arec, int F,
                                                                const int64_t* __restrict__ uoff,
                                                                const int32_t* __restrict__ uitems, int64_t n_users,
                                                                int32_t* __restrict__ out, int3* __restrict__ gstack,
                                                                int gcap, int32_t* __restrict__ errflag,
                                                                unsigned long long* __restrict__ dbg,
                                                                const int32_t* __restrict__ user_list, int scap) {
    extern __shared__ uint32_t smem_match[];
    const int words = (F + 31) >> 5;
    const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1;
    uint32_t* mask = smem_match + (size_t)wib * (words + 3 * MATCH_STACK_SMEM);
    int3* sstack = reinterpret_cast<int3*>(mask + words);
    const int64_t gw = (int64_t)blockIdx.x * MATCH_WARPS + wib;
    const int64_t n_warps = (int64_t)gridDim.x * MATCH_WARPS;
    int3* mystack = gstack + gw * gcap;
    auto has = [&](int it) -> bool { return (mask[it >> 5] >> (it & 31)) & 1u; };
    for (int64_t ui = gw; ui < n_users; ui += n_warps) {
        const int64_t u = user_list ? user_list[ui] : ui;   // second pass: only the baskets whose stack overflowed
        bool ovf = false;
        for (int w = lane; w < words; w += 32) mask[w] = 0;
        __syncwarp();
        const int64_t ub = uoff[u], ue = uoff[u + 1];
        for (int64_t e = ub + lane; e < ue; e += 32) {
            const int it = uitems[e];
            if ((unsigned)it < (unsigned)F) atomicOr(&mask[it >> 5], 1u << (it & 31));
        }
        __syncwarp();
        int best = 0x7fffffff, best_cons = -1, sp = 0;
        unsigned long long nodes = 0;
        // visit node (lvl, idx) on the lanes where `in` holds: its rules, then push its child range
        auto visit = [&](int lvl, int idx, bool in) {
            int c0 = 0, c1 = 0;
            if (in) {
                const int64_t gid = T.nodebase[lvl] + idx;
                const int e1 = aoff[gid + 1];
                for (int e = aoff[gid]; e < e1; ++e) {
                    const int2 rc = __ldg(arec + e);
                    if (rc.x >= best) break;
                    if (!has(rc.y)) { best = rc.x; best_cons = rc.y; break; }
                }
                if (T.child[lvl]) { c0 = T.child[lvl][idx]; c1 = T.child[lvl][idx + 1]; }
                ++nodes;
            }
            const bool push = c0 < c1;
            const unsigned pm = __ballot_sync(0xffffffffu, push);
            if (pm) {
                if (push) {
                    const int pos = sp + __popc(pm & lt);
                    const int3 ent = make_int3(lvl + 1, c0, c1);
                    if (pos < scap) sstack[pos] = ent;
                    else if (pos - scap < gcap) mystack[pos - scap] = ent;
                    else ovf = true;   // dropped: this basket is matched again with a stack sized for the worst case
                }
                sp += __popc(pm);
                __syncwarp();
            }
        };
        for (int w = 0; w < words; ++w) {
            const uint32_t bits = mask[w];
            if (bits) visit(1, (w << 5) + lane, (bits >> lane) & 1u);
        }
        while (sp > 0) {
            --sp;
            int3 ent;
            if (sp < scap) ent = sstack[sp];
            else if (sp - scap < gcap) ent = mystack[sp - scap];
            else ent = make_int3(2, 0, 0);  // dropped on overflow (errflag is set)
            __syncwarp();
            const int lvl = ent.x;
            const int32_t* rows = T.sets[lvl];
            for (int b = ent.y; b < ent.z; b += 32) {
                const int r = b + lane;
                bool in = r < ent.z;
                if (in) in = has(rows[(int64_t)r * lvl + (lvl - 1)]);
                visit(lvl, r, in);
            }
            // share the best priority so far: it prunes every later rule-list walk
            int mv = best, mc = best_cons;
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                const int ov = __shfl_xor_sync(0xffffffffu, mv, d), oc = __shfl_xor_sync(0xffffffffu, mc, d);
                if (ov < mv) { mv = ov; mc = oc; }
            }
            best = mv; best_cons = mc;
        }
        int mv = best, mc = best_cons;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            const int ov = __shfl_xor_sync(0xffffffffu, mv, d), oc = __shfl_xor_sync(0xffffffffu, mc, d);
            if (ov < mv) { mv = ov; mc = oc; }
        }
        if (__any_sync(0xffffffffu, ovf)) { mc = -2; if (lane == 0) *errflag = 2; }
        if (lane == 0) out[u] = mc;
        if (dbg) {
            unsigned long long tot = nodes;
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, d);
            if (lane == 0) { atomicAdd(dbg, tot); atomicMax(dbg + 1, tot); }
        }
        __syncwarp();
    }
}
