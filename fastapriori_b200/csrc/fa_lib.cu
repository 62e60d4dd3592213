// libfastapriori_b200 — C-ABI entry points (include/fastapriori_b200.h) over the sm_100a kernels.
// Host logic here is launch plumbing only: sizes, buffers, scans/sorts (CUB utilities) and the
// exchange hook.  There is no CPU compute path: every entry point fails if CUDA fails.

#include <cub/cub.cuh>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <memory>
#include <mutex>
#include <dlfcn.h>

#define COUNT_U_ COUNT_U
#define COUNT_G_ COUNT_G
#include "fa_kernels.cuh"
#include "fa_pair_tc.cuh"

#ifndef COUNT_U
#define COUNT_U 4   // uint4 per lane per chunk in K4: chunk = 32 * 4 * 16 B = 2 KB of every item row
#endif
#ifndef COUNT_G
#define COUNT_G 30  // consecutive prefix groups per warp run (each run starts by rebuilding a prefix AND: 16 -> 30 took K4 from 30.3 to 28.4 ms)
#endif
#define PITCH_ALIGN (32 * COUNT_U * 4)  // words: item rows are a whole number of K4 chunks (2 KB), so K4 needs no tail predicates
#define NOALLOC_G 16  // run length of the stand-alone entry point's kernel (k_count_groups), unchanged since it was added
#define COUNT_META_BYTES(m) (8 * NOALLOC_G * ((m) > 32 ? 96 : 64) * 4 + 8 * 32 * COUNT_U * 16)  // per warp: a run's group metadata + the cached prefix AND (48-64 KB per CTA)

namespace {

template <typename T>
int scan_exclusive(fa_ctx* ctx, const T* in, T* out, int64_t n) {
    if (n <= 0) return FA_OK;
    if (n > 0x7fffffff) return ctx->fail(FA_ELIMIT, "scan longer than 2^31-1");
    size_t bytes = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceScan::ExclusiveSum(ctx->scratch.p, bytes, in, out, (int)n, ctx->stream));
    return FA_OK;
}

struct ScalarSrc { const void* src; int n_words; };

// Brings a few device scalars to the host with ONE stream sync and no copy-engine traffic (see fa_ctx::h_scalars).
int read_scalars(fa_ctx* ctx, std::initializer_list<ScalarSrc> srcs, uint32_t* out) {
    WordSrcs w = {};
    int q = 0, off = 0;
    for (const ScalarSrc& s : srcs) {
        if (q == 4 || off + s.n_words > 240) return ctx->fail(FA_ELIMIT, "read_scalars: too many sources or words");
        w.p[q] = reinterpret_cast<const uint32_t*>(s.src);
        w.n[q] = s.n_words;
        off += s.n_words;
        ++q;
    }
    k_copy_words<<<1, 64, 0, ctx->stream>>>(w, ctx->d_scalars);
    LAUNCH_CHECK();
    CK(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < off; ++i) out[i] = ctx->h_scalars[i];
    return FA_OK;
}

constexpr int PRE_SLOT = 256;  // word of h_scalars that receives the next level's slot count (read_scalars uses < 240)

// First step of K5 for level P.k + 1 (one slot per (x, following sibling) pair), queued without a host sync at the end
// of the call that built P: by the time the host comes back with fa_gen_candidates the total is already in h_scalars.
// `n_bound` rows are covered; when the level's size is still on its way (`n_dev`, written by a kernel queued before),
// the rows past it count zero slots.  The total stays on the device at c_off[n_bound]; `to_host` also sends it to
// h_scalars[PRE_SLOT] (the caller that has its own read-back coming passes false and reads c_off[n_bound] with it).
int enqueue_sibling_slots(fa_ctx* ctx, Level& P, Level& Q, int64_t n_bound, const int32_t* n_dev, bool to_host) {
    ctx->pre_level = 0;
    ctx->pre_read = false;
    if (n_bound == 0 || n_bound > 0x7fffffff / 2) return FA_OK;
    CK(ctx->c_cnt.reserve(4 * (size_t)(n_bound + 1), ctx->stream));
    CK(ctx->c_off.reserve(4 * (size_t)(n_bound + 1), ctx->stream));
    int32_t* cnt = ctx->c_cnt.as<int32_t>();
    int32_t* soff = ctx->c_off.as<int32_t>();
    k_sibling_counts<<<(unsigned)cdiv(n_bound + 1, 256), 256, 0, ctx->stream>>>(P.sub.as<int32_t>(), Q.child.as<int32_t>(), (int)n_bound, n_dev,
                                                                                P.k, cnt);
    LAUNCH_CHECK();
    if (int rc = scan_exclusive(ctx, cnt, soff, n_bound + 1)) return rc;
    if (to_host) {
        WordSrcs w = {};
        w.p[0] = reinterpret_cast<const uint32_t*>(soff + n_bound);
        w.n[0] = 1;
        k_copy_words<<<1, 32, 0, ctx->stream>>>(w, ctx->d_scalars + PRE_SLOT);
        LAUNCH_CHECK();
    }
    ctx->pre_level = P.k + 1;
    return FA_OK;
}

// total of an exclusive scan = last scanned + last input; reads two ints back (one sync)
int scan_total(fa_ctx* ctx, const int32_t* in, const int32_t* scanned, int64_t n, int64_t* total) {
    if (n <= 0) { *total = 0; return FA_OK; }
    uint32_t h[2];
    if (int rc = read_scalars(ctx, {{in + n - 1, 1}, {scanned + n - 1, 1}}, h)) return rc;
    *total = (int64_t)(int32_t)h[0] + (int64_t)(int32_t)h[1];
    return FA_OK;
}

// ---- built-in NCCL exchange: libnccl.so.2 is opened at run time (in a torch process that is the copy torch already
// loaded), so the library has no link-time dependency on it and single-GPU users never touch it.
struct FaNcclId { char internal[FA_COMM_ID_BYTES]; };  // same layout as ncclUniqueId (nccl.h: char internal[128])
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, FaNcclId /* ncclUniqueId, by value */, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int NCCL_INT32 = 2, NCCL_SUM = 0;            // ncclInt32 / ncclSum in nccl.h

const NcclApi* nccl_api(std::string* why) {
    static NcclApi api;
    static std::string err;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { err = std::string("cannot open libnccl.so.2: ") + dlerror(); return; }
        api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
        api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
        api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
        api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
        if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy || !api.GetErrorString) {
            err = "libnccl.so.2 lacks an expected symbol";
            return;
        }
        api.lib = h;
    });
    if (!api.lib) { if (why) *why = err; return nullptr; }
    return &api;
}

int exchange(fa_ctx* ctx, int32_t* dev, int64_t count) {
    if (count <= 0 || (!ctx->nccl_comm && !ctx->xchg)) return FA_OK;
    // timed as its own class: the events bracket the collective on the ctx stream, so the span includes the wait for
    // the slowest rank's counting kernel (rank skew), which is what the 1 -> 8 GPU curve loses
    ctx->tbegin(T_EXCHANGE);
    int rc;
    if (ctx->nccl_comm) {
        const NcclApi* api = nccl_api(nullptr);
        rc = api->AllReduce(dev, dev, (size_t)count, NCCL_INT32, NCCL_SUM, ctx->nccl_comm, ctx->stream);
        if (rc != 0) ctx->fail(FA_EXCHANGE, std::string("ncclAllReduce: ") + api->GetErrorString(rc));
    } else {
        rc = ctx->xchg(ctx->xchg_user, dev, count, (void*)ctx->stream);
        if (rc != 0) ctx->fail(FA_EXCHANGE, "exchange callback returned " + std::to_string(rc));
    }
    ctx->tend();
    return rc != 0 ? FA_EXCHANGE : FA_OK;
}

// child ranges of level `P` (k-1) over level `L` (k): rows of L sorted by parent index sub[.][k-1]
// (`n_dev`: L's size when only the device knows it yet, else null and L.n counts)
int make_child_index(fa_ctx* ctx, Level& P, Level& L, const int32_t* n_dev = nullptr) {
    CK(P.child.reserve(sizeof(int32_t) * (size_t)(P.n + 2), ctx->stream));
    k_child_index<<<(unsigned)cdiv(P.n + 1, 256), 256, 0, ctx->stream>>>(L.sub.as<int32_t>(), L.n, n_dev, L.k, (int)P.n,
                                                                         P.child.as<int32_t>());
    LAUNCH_CHECK();
    return FA_OK;
}

void drop_levels(fa_ctx* ctx, size_t from) {
    if (ctx->copy_stream && ctx->levels.size() > from) cudaStreamSynchronize(ctx->copy_stream);  // copies of these levels may be in flight
    for (size_t k = from; k < ctx->levels.size(); ++k) {
        ctx->levels[k].sets.release(ctx->stream);
        ctx->levels[k].counts.release(ctx->stream);
        ctx->levels[k].sub.release(ctx->stream);
        ctx->levels[k].child.release(ctx->stream);
    }
    if (ctx->levels.size() > from) ctx->levels.resize(from);
    ctx->have_rules = ctx->have_sorted = false;
    ctx->pre_level = 0;
    ctx->pre_read = false;
}

int build_level_tab(fa_ctx* ctx) {
    LevelTab& T = ctx->tab;
    memset(&T, 0, sizeof T);
    int max_k = (int)ctx->levels.size() - 1;
    while (max_k >= 2 && ctx->levels[max_k].n == 0) --max_k;
    if (max_k >= FA_MAX_LEVELS) return ctx->fail(FA_ELIMIT, "more than 63 itemset levels");
    T.max_k = max_k;
    int64_t base = 0, nodes = 0;
    for (int k = 1; k <= max_k; ++k) {
        T.sets[k] = ctx->levels[k].sets.as<int32_t>();
        T.sub[k] = ctx->levels[k].sub.as<int32_t>();
        T.child[k] = (k < max_k) ? ctx->levels[k].child.as<int32_t>() : nullptr;
        T.nodebase[k] = nodes;
        nodes += ctx->levels[k].n;
        if (k >= 2) { T.base[k] = base; base += ctx->levels[k].n * k; }
    }
    for (int k = std::max(max_k, 1) + 1; k <= FA_MAX_LEVELS; ++k) { T.base[k] = base; T.nodebase[k] = nodes; }
    ctx->n_nodes = nodes;
    if (nodes > 0x7fffffff) return ctx->fail(FA_ELIMIT, "more than 2^31-1 frequent itemsets");
    if (base > 0x7fffffff) return ctx->fail(FA_ELIMIT, "more than 2^31-1 raw rules");
    ctx->n_raw = base;
    return FA_OK;
}

// After counting level k: K4 left in t_alive the rows that lie in some group prefix's tidset.  trim_prepare counts them
// on the device (the caller reads the total together with the frequent-candidate total: one host sync for both);
// trim_apply squeezes the dead rows out when that is much less than what the bitmaps hold (see the trimming kernels).
int trim_prepare(fa_ctx* ctx) {
    const int64_t pitch = ctx->pitch_words;
    CK(ctx->t_cnt.reserve(4 * (size_t)(pitch + 1), ctx->stream));
    CK(ctx->t_off.reserve(4 * (size_t)(pitch + 1), ctx->stream));
    k_popc_words<<<(unsigned)cdiv(pitch, 256), 256, 0, ctx->stream>>>(ctx->t_alive.as<uint32_t>(), pitch, ctx->t_cnt.as<int32_t>());
    LAUNCH_CHECK();
    CK(cudaMemsetAsync(ctx->t_cnt.as<int32_t>() + pitch, 0, 4, ctx->stream));
    return scan_exclusive(ctx, ctx->t_cnt.as<int32_t>(), ctx->t_off.as<int32_t>(), pitch + 1);
}

int trim_apply(fa_ctx* ctx, Level& L, int32_t alive_rows) {
    const int k = L.k;
    const int F = ctx->F;
    if (L.n < k + 1 || ctx->no_trim || ctx->weighted) return FA_OK;  // the level loop ends here anyway; weight planes: squeezing rows would break the chunk -> plane map
    const int64_t pitch = ctx->pitch_words;
    const int64_t new_pitch = std::max<int64_t>(PITCH_ALIGN, cdiv(cdiv((int64_t)alive_rows, 32), PITCH_ALIGN) * PITCH_ALIGN);
    if (new_pitch * 20 > pitch * 17) return FA_OK;  // less than 15 % shorter: not worth a pass over the bitmaps
    const size_t smem = 4 * (size_t)F;
    if (smem > 200 * 1024) return FA_OK;  // too many items for the in-kernel active list: skip trimming
    CK(ctx->t_active.reserve(4 * (size_t)F, ctx->stream));
    int32_t* active = ctx->t_active.as<int32_t>();
    CK(cudaMemsetAsync(active, 0, 4 * (size_t)F, ctx->stream));
    k_mark_active<<<(unsigned)std::min<int64_t>(cdiv(L.n * k, 4096), ctx->sm_count), 1024, 4 * (size_t)cdiv(F, 32), ctx->stream>>>(L.sets.as<int32_t>(), L.n * k, F, active);
    LAUNCH_CHECK();
    const size_t bytes = 4 * (size_t)new_pitch * (size_t)F;
    // (reserved at the size of the untrimmed bitmaps: the two buffers swap roles at every trim, and the next run's
    // fa_build_bitmaps must find a full-size buffer whichever of them is current, or it re-allocates tens of GB per run)
    CK(ctx->bits2.reserve(std::max(bytes, ctx->bits.cap - ctx->bits.cap / 5 - 256), ctx->stream));
    CK(cudaMemsetAsync(ctx->bits2.p, 0, bytes, ctx->stream));
    if (smem > 48 * 1024) CK(cudaFuncSetAttribute(k_compact_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // enough CTAs for a few waves: the word axis alone is short once the rows have been trimmed a few times
    const unsigned wblocks = (unsigned)cdiv(pitch, 256);
    const unsigned islices = (unsigned)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(cdiv(F, 32), 64), cdiv((int64_t)ctx->sm_count * 8, wblocks)));
    k_compact_rows<<<dim3(wblocks, islices), 256, smem, ctx->stream>>>(ctx->bits.as<uint32_t>(), pitch, ctx->t_alive.as<uint32_t>(),
                                                                          ctx->t_off.as<int32_t>(), active, F,
                                                                          ctx->bits2.as<uint32_t>(), new_pitch);
    LAUNCH_CHECK();
    std::swap(ctx->bits, ctx->bits2);
    ctx->pitch_words = new_pitch;
    ctx->live_rows = alive_rows;
    ctx->bits_trimmed = true;
    ctx->trims++;
    if (ctx->debug_trim) fprintf(stderr, "[trim] after level %d: %d rows alive, pitch %lld -> %lld words\n", k, alive_rows, (long long)pitch, (long long)new_pitch);
    return FA_OK;
}

// SURVEY.md 8 f-4 (FastApriori.scala:66-79): merge identical filtered baskets into weighted rows.  Called with F <= 64, so a
// basket is a 64-bit mask and the merge is exact.  On success with ctx->weighted set, ctx->bits holds the weight planes
// (see the kernels' comment), ctx->pitch_words their length and ctx->chunk_shift the plane of every 2 KB chunk; otherwise
// (the merge would not at least halve the row axis, and it was not forced) nothing but ctx->total_count has changed.
int build_weighted(fa_ctx* ctx) {
    const int F = ctx->F;
    const int64_t n = ctx->n_rows;
    if (n > 0x7fffffff) return FA_OK;
    CK(ctx->w_masks.reserve(8 * (size_t)n, ctx->stream));
    CK(ctx->w_sorted.reserve(8 * (size_t)n, ctx->stream));
    CK(ctx->w_umask.reserve(8 * (size_t)n, ctx->stream));
    CK(ctx->w_weights.reserve(4 * (size_t)n, ctx->stream));
    CK(ctx->w_misc.reserve(4 * 80, ctx->stream));
    auto* masks = ctx->w_masks.as<unsigned long long>();
    auto* sorted = ctx->w_sorted.as<unsigned long long>();
    auto* umask = ctx->w_umask.as<unsigned long long>();
    int32_t* weights = ctx->w_weights.as<int32_t>();
    int32_t* d_runs = ctx->w_misc.as<int32_t>();                   // [0] number of runs
    unsigned* d_plane = ctx->w_misc.as<unsigned>() + 4;            // [4..36] plane counts, [36] distinct baskets
    unsigned* d_cursor = ctx->w_misc.as<unsigned>() + 40;          // [40..71] per-plane cursors of the scatter
    CK(cudaMemsetAsync(ctx->w_misc.p, 0, 4 * 80, ctx->stream));
    k_row_masks<<<(unsigned)cdiv(n, 256), 256, 0, ctx->stream>>>(ctx->d_off, ctx->d_items, ctx->id2rank.as<int32_t>(), n, 2, masks);
    LAUNCH_CHECK();
    size_t bytes = 0;
    CK(cub::DeviceRadixSort::SortKeys(nullptr, bytes, masks, sorted, (int)n, 0, F, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceRadixSort::SortKeys(ctx->scratch.p, bytes, masks, sorted, (int)n, 0, F, ctx->stream));
    CK(cub::DeviceRunLengthEncode::Encode(nullptr, bytes, sorted, umask, weights, d_runs, (int)n, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceRunLengthEncode::Encode(ctx->scratch.p, bytes, sorted, umask, weights, d_runs, (int)n, ctx->stream));
    k_plane_counts<<<(unsigned)std::min<int64_t>(cdiv(n, 256), 1024), 256, 0, ctx->stream>>>(umask, weights, d_runs, d_plane);
    LAUNCH_CHECK();
    uint32_t h[34];
    if (int rc = read_scalars(ctx, {{d_runs, 1}, {d_plane, 33}}, h)) return rc;
    const int64_t runs = (int32_t)h[0];
    ctx->total_count = (int64_t)h[33];
    PlaneOffsets po;
    int64_t v = 0;
    for (int b = 0; b < 32; ++b) {
        po.row0[b] = v;
        v += cdiv((int64_t)h[1 + b], 32 * PITCH_ALIGN) * 32 * PITCH_ALIGN;  // planes start on chunk boundaries
    }
    po.row0[32] = v;
    if (v == 0) v = 32 * PITCH_ALIGN;
    // Unless forced: only when the merge removes three quarters of the row axis AND the pair pass is big enough to repay the
    // sort (measured on B200 at 10 M rows, 35 frequent items: merging costs 0.9 ms more than K2 and saves 0.4 ms of K3/K4)
    if (ctx->weighted_mode != 1 && v * 4 > n) return FA_OK;
    ctx->pitch_words = v / 32;
    const int64_t n_chunks = ctx->pitch_words / PITCH_ALIGN;
    const size_t bbytes = 4 * (size_t)ctx->pitch_words * (size_t)F;
    CK(ctx->bits.reserve(bbytes, ctx->stream));
    CK(cudaMemsetAsync(ctx->bits.p, 0, bbytes, ctx->stream));
    CK(ctx->chunk_shift.reserve((size_t)n_chunks, ctx->stream));
    k_chunk_shifts<<<(unsigned)cdiv(n_chunks, 256), 256, 0, ctx->stream>>>(po, n_chunks, ctx->chunk_shift.as<uint8_t>());
    LAUNCH_CHECK();
    if (runs > 0) {
        k_weighted_scatter<<<(unsigned)cdiv(runs, 256), 256, 0, ctx->stream>>>(umask, weights, (int)runs, po, d_cursor, ctx->bits.as<uint32_t>(), ctx->pitch_words);
        LAUNCH_CHECK();
    }
    ctx->weighted = true;
    ctx->live_rows = 0;
    for (int b = 0; b < 32; ++b) ctx->live_rows += h[1 + b];  // virtual rows in use
    return FA_OK;
}

}  // namespace

extern "C" {

int fa_create(int device, void* cuda_stream, fa_ctx** ctx_out) {
    if (!ctx_out) return FA_EINVAL;
    *ctx_out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return FA_ECUDA;
    if (cudaSetDevice(device) != cudaSuccess) return FA_ECUDA;
    fa_ctx* ctx = new fa_ctx();
    ctx->device = device;
    if (cuda_stream) ctx->stream = (cudaStream_t)cuda_stream;
    else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return FA_ECUDA; }
        ctx->own_stream = true;
    }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    ctx->no_trim = getenv("FA_NO_TRIM") != nullptr;
    { const char* w = getenv("FA_WEIGHTED"); ctx->weighted_mode = w ? atoi(w) : -1; }
    { const char* g = getenv("FA_MATCH_GCAP"); if (g && atoi(g) >= 0) ctx->match_gcap = atoi(g); }
    { const char* g = getenv("FA_MATCH_SCAP"); if (g && atoi(g) >= 1 && atoi(g) <= MATCH_STACK_SMEM) ctx->match_scap = atoi(g); }
    { const char* u = getenv("FA_USER_DEDUP"); ctx->user_dedup = !(u && atoi(u) == 0); }
    { const char* k4 = getenv("FA_K4"); ctx->k4_groups = !(k4 && strcmp(k4, "tiles") == 0); }
    { const char* w = getenv("FA_K5_WAIT"); ctx->k5_wait = w && strcmp(w, "1") == 0; }
    { const char* w = getenv("FA_K2_SMEM_KB"); const int v = w ? atoi(w) : 0; if (v >= 16 && v <= 200) ctx->k2_smem_kb = v; }
    { const char* w = getenv("FA_K4_WARPS"); const int v = w ? atoi(w) : 2; ctx->k4_warps = (v == 1 || v == 4 || v == 8) ? v : 2; }
    ctx->debug_trim = getenv("FA_DEBUG_TRIM") != nullptr;
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->copy_event, cudaEventDisableTiming) != cudaSuccess) { delete ctx; return FA_ECUDA; }
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t thr = ~0ull;  // keep freed blocks cached: the level loop re-allocates the same sizes
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    if (cudaHostAlloc((void**)&ctx->h_scalars, 2048, cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer((void**)&ctx->d_scalars, ctx->h_scalars, 0) != cudaSuccess) { delete ctx; return FA_ENOMEM; }
    ctx->levels.resize(1);
    if (ctx->scalars.reserve(512, ctx->stream) != cudaSuccess || ctx->errflag.reserve(16, ctx->stream) != cudaSuccess) {
        delete ctx;
        return FA_ENOMEM;
    }
    cudaMemsetAsync(ctx->errflag.p, 0, 16, ctx->stream);
    *ctx_out = ctx;
    return FA_OK;
}

void fa_destroy(fa_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
    ctx->resolve_timers();
    for (auto e : ctx->ev_pool) cudaEventDestroy(e);
    drop_levels(ctx, 0);
    DevBuf* bufs[] = {&ctx->own_off, &ctx->own_items, &ctx->own_items16, &ctx->id2rank, &ctx->bits, &ctx->bits2, &ctx->t_active, &ctx->t_alive, &ctx->t_cnt, &ctx->t_off, &ctx->c_cnt, &ctx->c_off, &ctx->c_gflag,
                      &ctx->c_gpos, &ctx->c_gx, &ctx->c_gdesc, &ctx->c_goff, &ctx->c_ys, &ctx->c_cx, &ctx->c_counts, &ctx->c_flag,
                      &ctx->c_pos, &ctx->w_masks, &ctx->w_sorted, &ctx->w_umask, &ctx->w_weights, &ctx->w_misc, &ctx->chunk_shift, &ctx->c_tflag, &ctx->c_tpos, &ctx->c_tdesc, &ctx->c_tstream, &ctx->c_sflag, &ctx->c_spos, &ctx->c_slotx, &ctx->c_slotsub, &ctx->c_csub, &ctx->pair_counts, &ctx->pair_tiles, &ctx->pair_rowcnt, &ctx->pair_rowoff,
                      &ctx->scratch, &ctx->scalars, &ctx->errflag, &ctx->r_conf, &ctx->r_kept, &ctx->r_list,
                      &ctx->r_key32, &ctx->r_key64, &ctx->r_tmp32, &ctx->r_tmp64, &ctx->r_val, &ctx->r_val2,
                      &ctx->r_order, &ctx->m_aoff, &ctx->m_lp, &ctx->m_rec, &ctx->m_stack, &ctx->u_off, &ctx->u_items, &ctx->u_out, &ctx->id2rank_u, &ctx->r_itemint,
                      &ctx->ud_idx, &ctx->ud_user, &ctx->ud_flag, &ctx->ud_run, &ctx->ud_len, &ctx->ud_off, &ctx->ud_ranks, &ctx->ud_rec};
    for (DevBuf* b : bufs) b->release(ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->nccl_comm) { cudaStreamSynchronize(ctx->stream); nccl_api(nullptr)->CommDestroy(ctx->nccl_comm); }
    if (ctx->copy_event) cudaEventDestroy(ctx->copy_event);
    if (ctx->h_scalars) cudaFreeHost(ctx->h_scalars);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* fa_last_error(const fa_ctx* ctx) { return ctx ? ctx->err.c_str() : "null ctx"; }

int fa_set_exchange(fa_ctx* ctx, fa_allreduce_fn fn, void* user) {
    if (!ctx) return FA_EINVAL;
    ctx->xchg = fn;
    ctx->xchg_user = user;
    return FA_OK;
}

int fa_comm_unique_id(uint8_t* id_out, char* err_out, int64_t err_cap) {
    auto fail = [&](int code, const std::string& msg) {
        if (err_out && err_cap > 0) snprintf(err_out, (size_t)err_cap, "%s", msg.c_str());
        return code;
    };
    if (!id_out) return fail(FA_EINVAL, "id_out is null");
    std::string why;
    const NcclApi* api = nccl_api(&why);
    if (!api) return fail(FA_EXCHANGE, why);
    FaNcclId id;
    int rc = api->GetUniqueId(&id);
    if (rc != 0) return fail(FA_EXCHANGE, std::string("ncclGetUniqueId: ") + api->GetErrorString(rc));
    memcpy(id_out, id.internal, FA_COMM_ID_BYTES);
    return FA_OK;
}

int fa_comm_init(fa_ctx* ctx, const uint8_t* id, int32_t rank, int32_t world_size) {
    if (!ctx) return FA_EINVAL;
    if (!id || world_size < 1 || rank < 0 || rank >= world_size) return ctx->fail(FA_EINVAL, "fa_comm_init: bad id/rank/world_size");
    if (ctx->nccl_comm) return ctx->fail(FA_EINVAL, "fa_comm_init: this context already has a communicator");
    std::string why;
    const NcclApi* api = nccl_api(&why);
    if (!api) return ctx->fail(FA_EXCHANGE, why);
    CK(cudaSetDevice(ctx->device));
    FaNcclId nid;
    memcpy(nid.internal, id, FA_COMM_ID_BYTES);
    void* comm = nullptr;
    int rc = api->CommInitRank(&comm, world_size, nid, rank);
    if (rc != 0) return ctx->fail(FA_EXCHANGE, std::string("ncclCommInitRank: ") + api->GetErrorString(rc));
    ctx->nccl_comm = comm;
    return FA_OK;
}

static int reset_tx(fa_ctx* ctx, int64_t n_rows, int32_t n_items) {
    if (n_rows < 0 || n_items < 0) return ctx->fail(FA_EINVAL, "negative size");
    if (n_rows > 0x7fffffff) return ctx->fail(FA_ELIMIT, "more than 2^31-1 rows per context (counts are Int32, FastApriori.scala:74)");
    CK(cudaSetDevice(ctx->device));
    drop_levels(ctx, 1);
    ctx->have_bits = false;
    ctx->F = 0;
    ctx->n_rows = n_rows;
    ctx->n_items = n_items;
    return FA_OK;
}

int fa_load_transactions(fa_ctx* ctx, const int64_t* row_offsets, const int32_t* item_ids, int64_t n_rows,
                         int32_t n_items) {
    if (!ctx) return FA_EINVAL;
    if (!row_offsets) return ctx->fail(FA_EINVAL, "row_offsets is null");
    if (int rc = reset_tx(ctx, n_rows, n_items)) return rc;
    const int64_t nnz = row_offsets[n_rows] - row_offsets[0];
    if (row_offsets[0] != 0 || nnz < 0) return ctx->fail(FA_EINVAL, "row_offsets must start at 0 and be non-decreasing");
    if (nnz > 0 && !item_ids) return ctx->fail(FA_EINVAL, "item_ids is null");
    CK(ctx->own_off.reserve(sizeof(int64_t) * (size_t)(n_rows + 1), ctx->stream));
    CK(ctx->own_items.reserve(sizeof(int32_t) * (size_t)std::max<int64_t>(nnz, 1), ctx->stream));
    CK(cudaMemcpyAsync(ctx->own_off.p, row_offsets, sizeof(int64_t) * (size_t)(n_rows + 1), cudaMemcpyHostToDevice, ctx->stream));
    if (nnz) CK(cudaMemcpyAsync(ctx->own_items.p, item_ids, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, ctx->stream));
    ctx->d_off = ctx->own_off.as<int64_t>();
    ctx->d_items = ctx->own_items.as<int32_t>();
    ctx->nnz = nnz;
    ctx->have_tx = true;
    return FA_OK;
}

int fa_load_transactions_u16(fa_ctx* ctx, const int64_t* row_offsets, const uint16_t* item_ids, int64_t n_rows,
                             int32_t n_items) {
    if (!ctx) return FA_EINVAL;
    if (!row_offsets) return ctx->fail(FA_EINVAL, "row_offsets is null");
    if (n_items > 65536) return ctx->fail(FA_EINVAL, "fa_load_transactions_u16: more than 65536 item ids do not fit 16 bits");
    if (int rc = reset_tx(ctx, n_rows, n_items)) return rc;
    const int64_t nnz = row_offsets[n_rows] - row_offsets[0];
    if (row_offsets[0] != 0 || nnz < 0) return ctx->fail(FA_EINVAL, "row_offsets must start at 0 and be non-decreasing");
    if (nnz > 0 && !item_ids) return ctx->fail(FA_EINVAL, "item_ids is null");
    CK(ctx->own_off.reserve(sizeof(int64_t) * (size_t)(n_rows + 1), ctx->stream));
    CK(ctx->own_items.reserve(sizeof(int32_t) * (size_t)std::max<int64_t>(nnz, 1), ctx->stream));
    CK(ctx->own_items16.reserve(sizeof(uint16_t) * (size_t)std::max<int64_t>(nnz, 1), ctx->stream));
    CK(cudaMemcpyAsync(ctx->own_off.p, row_offsets, sizeof(int64_t) * (size_t)(n_rows + 1), cudaMemcpyHostToDevice, ctx->stream));
    if (nnz) {
        CK(cudaMemcpyAsync(ctx->own_items16.p, item_ids, sizeof(uint16_t) * (size_t)nnz, cudaMemcpyHostToDevice, ctx->stream));
        const unsigned blocks = (unsigned)std::max<int64_t>(1, std::min<int64_t>(cdiv(cdiv(nnz, 8), 256), (int64_t)ctx->sm_count * 8));
        k_widen_u16<<<blocks, 256, 0, ctx->stream>>>(ctx->own_items16.as<uint16_t>(), ctx->own_items.as<int32_t>(), nnz);
        LAUNCH_CHECK();
    }
    ctx->d_off = ctx->own_off.as<int64_t>();
    ctx->d_items = ctx->own_items.as<int32_t>();
    ctx->nnz = nnz;
    ctx->have_tx = true;
    return FA_OK;
}

int fa_attach_transactions(fa_ctx* ctx, const int64_t* dev_row_offsets, const int32_t* dev_item_ids, int64_t n_rows,
                           int32_t n_items) {
    if (!ctx) return FA_EINVAL;
    if (!dev_row_offsets) return ctx->fail(FA_EINVAL, "dev_row_offsets is null");
    if (int rc = reset_tx(ctx, n_rows, n_items)) return rc;
    int64_t ends[2] = {0, 0};
    CK(cudaMemcpyAsync(&ends[0], dev_row_offsets, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(&ends[1], dev_row_offsets + n_rows, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ends[0] != 0 || ends[1] < 0) return ctx->fail(FA_EINVAL, "row_offsets must start at 0 and be non-decreasing");
    ctx->d_off = dev_row_offsets;
    ctx->d_items = dev_item_ids;
    ctx->nnz = ends[1];
    ctx->have_tx = true;
    return FA_OK;
}

// ---- K1 ------------------------------------------------------------------------------------
int fa_count_items(fa_ctx* ctx, int32_t* counts_out) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_tx) return ctx->fail(FA_EINVAL, "fa_count_items before fa_load_transactions");
    if (!counts_out && ctx->n_items) return ctx->fail(FA_EINVAL, "counts_out is null");
    CK(cudaSetDevice(ctx->device));
    const int32_t n_items = ctx->n_items;
    const bool sharded = ctx->nccl_comm || ctx->xchg;
    ctx->global_rows = sharded ? -1 : ctx->n_rows;
    if (n_items == 0) return FA_OK;
    CK(ctx->c_counts.reserve(sizeof(int32_t) * (size_t)(n_items + 3), ctx->stream));
    int32_t* d_counts = ctx->c_counts.as<int32_t>();
    CK(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)n_items, ctx->stream));
    CK(cudaMemsetAsync(ctx->errflag.p, 0, 4, ctx->stream));
    ctx->tbegin(T_HIST);
    if (ctx->nnz > 0) {
        const int threads = 512;
        const int64_t want = cdiv(cdiv(ctx->nnz, 4), threads);
        const unsigned blocks = (unsigned)std::max<int64_t>(1, std::min<int64_t>(want, (int64_t)ctx->sm_count * 4));
        const size_t smem = sizeof(int32_t) * (size_t)n_items;
        if (smem <= 96 * 1024) {
            CK(cudaFuncSetAttribute(k_item_hist<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
            k_item_hist<true><<<blocks, threads, smem, ctx->stream>>>(ctx->d_items, ctx->nnz, n_items, d_counts, ctx->errflag.as<int32_t>());
        } else {
            k_item_hist<false><<<blocks, threads, 0, ctx->stream>>>(ctx->d_items, ctx->nnz, n_items, d_counts, ctx->errflag.as<int32_t>());
        }
        LAUNCH_CHECK();
    }
    ctx->tend();
    // the shards' row counts ride along as three 20-bit limbs behind the item counts: data.count() (FastApriori.scala:38)
    // then needs no collective of its own
    int32_t limbs[3] = {0, 0, 0};
    if (sharded) {
        k_set3_i32<<<1, 1, 0, ctx->stream>>>(d_counts + n_items, (int32_t)(ctx->n_rows & 0xFFFFF), (int32_t)((ctx->n_rows >> 20) & 0xFFFFF),
                                             (int32_t)(ctx->n_rows >> 40));
        LAUNCH_CHECK();
    }
    if (int rc = exchange(ctx, d_counts, sharded ? n_items + 3 : n_items)) return rc;
    int32_t bad = 0;
    CK(cudaMemcpyAsync(counts_out, d_counts, sizeof(int32_t) * (size_t)n_items, cudaMemcpyDeviceToHost, ctx->stream));
    if (sharded) CK(cudaMemcpyAsync(limbs, d_counts + n_items, 12, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(&bad, ctx->errflag.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (sharded) ctx->global_rows = (int64_t)limbs[0] + ((int64_t)limbs[1] << 20) + ((int64_t)limbs[2] << 40);
    if (bad) return ctx->fail(FA_EINVAL, "item id outside [0, n_items)");
    return FA_OK;
}

int fa_global_rows(fa_ctx* ctx, int64_t* n_rows_out) {
    if (!ctx) return FA_EINVAL;
    if (!n_rows_out) return ctx->fail(FA_EINVAL, "n_rows_out is null");
    if (ctx->global_rows < 0) return ctx->fail(FA_EINVAL, "fa_global_rows before fa_count_items (or no items)");
    *n_rows_out = ctx->global_rows;
    return FA_OK;
}

// ---- K2 ------------------------------------------------------------------------------------
int fa_build_bitmaps(fa_ctx* ctx, const int32_t* id_to_rank, int32_t n_frequent, const int32_t* rank_counts) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_tx) return ctx->fail(FA_EINVAL, "fa_build_bitmaps before fa_load_transactions");
    if (n_frequent < 0 || n_frequent > ctx->n_items) return ctx->fail(FA_EINVAL, "n_frequent out of range");
    if (ctx->n_items && !id_to_rank) return ctx->fail(FA_EINVAL, "id_to_rank is null");
    if (n_frequent && !rank_counts) return ctx->fail(FA_EINVAL, "rank_counts is null");
    for (int32_t i = 0; i < ctx->n_items; ++i)
        if (id_to_rank[i] < -1 || id_to_rank[i] >= n_frequent) return ctx->fail(FA_EINVAL, "id_to_rank entry out of range");
    CK(cudaSetDevice(ctx->device));
    drop_levels(ctx, 1);
    const int32_t F = n_frequent;
    ctx->F = F;
    const int64_t words = cdiv(ctx->n_rows, 32);
    ctx->pitch_words = std::max<int64_t>(PITCH_ALIGN, cdiv(words, PITCH_ALIGN) * PITCH_ALIGN);
    CK(ctx->id2rank.reserve(sizeof(int32_t) * (size_t)std::max(ctx->n_items, 1), ctx->stream));
    if (ctx->n_items) CK(cudaMemcpyAsync(ctx->id2rank.p, id_to_rank, sizeof(int32_t) * (size_t)ctx->n_items, cudaMemcpyHostToDevice, ctx->stream));
    // level 1: identity rank rows + level-1 counts (FastApriori.scala:83)
    ctx->levels.resize(2);
    Level& L1 = ctx->levels[1];
    L1.k = 1;
    L1.n = F;
    CK(L1.sets.reserve(sizeof(int32_t) * (size_t)std::max(F, 1), ctx->stream));
    CK(L1.counts.reserve(sizeof(int32_t) * (size_t)std::max(F, 1), ctx->stream));
    if (F) {
        k_iota<<<(unsigned)cdiv(F, 256), 256, 0, ctx->stream>>>(L1.sets.as<int32_t>(), F);
        LAUNCH_CHECK();
        CK(cudaMemcpyAsync(L1.counts.p, rank_counts, sizeof(int32_t) * (size_t)F, cudaMemcpyHostToDevice, ctx->stream));
    }
    ctx->weighted = false;
    ctx->total_count = -1;
    ctx->sup_max = 0;
    ctx->sup_min = 0;
    for (int32_t r = 0; r < F; ++r) ctx->sup_max = std::max<int64_t>(ctx->sup_max, rank_counts[r]);
    ctx->tbegin(T_BITMAP);
    // FA_WEIGHTED=1: whenever a basket fits a 64-bit mask; unset: only from 48 frequent items up, where the pair pass is
    // big enough to repay the sort of the row masks (build_weighted then still requires a 4x shorter row axis)
    if (F >= 2 && F <= 64 && ctx->n_rows > 0 && (ctx->weighted_mode == 1 || (ctx->weighted_mode < 0 && F >= 48))) {
        if (int rc = build_weighted(ctx)) return rc;
    }
    const size_t bytes = sizeof(uint32_t) * (size_t)ctx->pitch_words * (size_t)std::max(F, 1);
    if (!ctx->weighted) CK(ctx->bits.reserve(bytes, ctx->stream));
    if (F > 0 && !ctx->weighted) {
        if (ctx->n_rows == 0) {
            CK(cudaMemsetAsync(ctx->bits.p, 0, bytes, ctx->stream));
        } else {
            // largest row tile whose [F x tile_rows] bit tile fits in the CTA's shared-memory budget; else chunk the ranks
            size_t budget = (size_t)ctx->k2_smem_kb * 1024;
            int tile_rows = 1024, rank_chunk = F;
            for (;;) {
                tile_rows = 1024;
                while (tile_rows > 256 && (size_t)F * (tile_rows / 8) + 4 * (size_t)(tile_rows + 1) > budget) tile_rows >>= 1;
                rank_chunk = F;
                if ((size_t)F * (tile_rows / 8) + 4 * (size_t)(tile_rows + 1) > budget)
                    rank_chunk = (int)((budget - 4 * (size_t)(tile_rows + 1)) / (tile_rows / 8));
                if (rank_chunk == F || budget >= 200 * 1024) break;
                budget = 200 * 1024;  // rather one CTA per SM than a second pass over the transactions
            }
            const int passes = (int)cdiv(F, rank_chunk);
            const int64_t n_tiles = cdiv(ctx->n_rows, tile_rows);
            const size_t smem = (size_t)rank_chunk * (tile_rows / 8) + 4 * (size_t)(tile_rows + 1);
            CK(cudaFuncSetAttribute(k_build_bitmaps, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            dim3 grid((unsigned)n_tiles, (unsigned)passes);
            k_build_bitmaps<<<grid, 512, smem, ctx->stream>>>(ctx->d_off, ctx->d_items, ctx->id2rank.as<int32_t>(), ctx->n_rows,
                                                             tile_rows, rank_chunk, F, ctx->bits.as<uint32_t>(),
                                                             ctx->pitch_words, n_tiles);
            LAUNCH_CHECK();
        }
    }
    ctx->tend();
    ctx->have_bits = true;
    ctx->bits_trimmed = false;
    if (!ctx->weighted) ctx->live_rows = ctx->n_rows;
    return FA_OK;
}

// ---- K3 ------------------------------------------------------------------------------------
int fa_count_pairs(fa_ctx* ctx, int32_t min_count, int64_t* n_frequent_out) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_bits) return ctx->fail(FA_EINVAL, "fa_count_pairs before fa_build_bitmaps");
    if (ctx->bits_trimmed) return ctx->fail(FA_EINVAL, "the level loop trimmed the bitmaps; call fa_build_bitmaps again before a new pair pass");
    if (!n_frequent_out) return ctx->fail(FA_EINVAL, "n_frequent_out is null");
    CK(cudaSetDevice(ctx->device));
    drop_levels(ctx, 2);
    const int F = ctx->F;
    ctx->sup_min = min_count > 0 ? min_count : -1;  // -1: a threshold below 1 bounds nothing
    ctx->levels.resize(3);
    Level& L2 = ctx->levels[2];
    L2.k = 2;
    L2.n = 0;
    *n_frequent_out = 0;
    if (F < 2) {
        CK(L2.sets.reserve(8, ctx->stream));
        CK(L2.sub.reserve(8, ctx->stream));
        CK(L2.counts.reserve(4, ctx->stream));
        return make_child_index(ctx, ctx->levels[1], L2);
    }
    const size_t pbytes = sizeof(int32_t) * (size_t)F * (size_t)F;
    CK(ctx->pair_counts.reserve(pbytes, ctx->stream));
    int32_t* P = ctx->pair_counts.as<int32_t>();
    CK(cudaMemsetAsync(P, 0, pbytes, ctx->stream));
    // tensor-core path (tcgen05 int8 GEMM) from 256 frequent items up; below that a 128x128 tile is mostly padding
    // and the POPC kernel wins.  FA_PAIR_IMPL=tc|popc forces one of them (tests compare the two).
    const char* impl_env = getenv("FA_PAIR_IMPL");
    const bool use_tc = !ctx->weighted && (impl_env ? strcmp(impl_env, "tc") == 0 : F >= 256);  // (weighted rows exist only for F <= 64)
    // tensor-core kernel: above 1024 items the CTA-pair kernel (k_pair_tc4: cta_group::2, 256 x 256 items per pair, tensor
    // pipe 80 % active), else 128 x 128 per CTA (k_pair_tc: more, smaller CTAs for the small pair matrices; measured at 1000
    // items x 1 M rows: 0.56 ms).  FA_PAIR_TILE = 128 | 256 (k_pair_tc2) | 512 (k_pair_tc3) | p256 (k_pair_tc4) forces one.
    const char* tile_env = getenv("FA_PAIR_TILE");
    const bool tcp256 = use_tc && (tile_env ? strcmp(tile_env, "p256") == 0 : F > 1024);
    const bool tc256 = use_tc && tile_env && atoi(tile_env) == 256;
    const bool tc512 = use_tc && tile_env && atoi(tile_env) == 512;
    const int tile = tcp256 ? 256 : use_tc ? TC_TILE : PAIR_T;
    const int nt = (int)cdiv(F, tile);           // column (B) blocks
    const int row_step = tcp256 ? 1 : tc512 ? 4 : tc256 ? 2 : 1;   // a 256- (512-) item row block spans two (four) column blocks
    const int nta = (int)cdiv(nt, row_step);     // row (A) blocks
    const int list_key = nt * 16 + row_step * 2 + (tcp256 ? 1 : 0);
    if (ctx->pair_tiles_nt != list_key) {  // the tile list depends on (nt, tile shape) only: built once, reused by every later run
        std::vector<int2> tiles;
        for (int a = 0; a < nta; ++a)
            for (int b = a * row_step; b < nt; ++b) tiles.push_back(make_int2(a, b));
        CK(ctx->pair_tiles.reserve(sizeof(int2) * tiles.size(), ctx->stream));
        CK(cudaMemcpyAsync(ctx->pair_tiles.p, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));  // `tiles` is pageable host memory about to go out of scope
        ctx->pair_tiles_nt = list_key;
        ctx->pair_tiles_count = (int64_t)tiles.size();
    }
    const int64_t n_tiles = ctx->pair_tiles_count;
    ctx->tbegin(T_PAIRS);
    if (use_tc) {
        const int64_t n_kb = ctx->pitch_words / 4;
        const int ctas_per_sm = (tc256 || tc512 || tcp256) ? 1 : 2;
        // enough CTAs for ~8 waves, so that the last, partly filled wave costs a few per cent (at 4687 items the 361 tiles
        // of 256 x 128 are 2.4 waves of 148 one-CTA SMs: a fifth of the machine idles in the third), but slices of at least
        // 512 k-blocks, so that a CTA's epilogue and pipeline fill stay small next to its MMAs
        const int64_t n_ctas = (tcp256 || tc512) ? 2 * n_tiles : n_tiles;   // the pair kernels run two CTAs per tile
        int64_t splits = std::max<int64_t>(1, std::min<int64_t>(n_kb / 512 + 1, cdiv((int64_t)ctx->sm_count * ctas_per_sm * 8, n_ctas)));
        // Blocks are dispatched tile-fastest, so all tiles of one row slice run together.  When a slice of ALL item rows
        // (F x kb x 16 B) of at least 2048 k-blocks fits in ~48 MB of L2, slice that way: every row then comes from HBM about
        // once instead of once per tile pair (cfg-4/5 shapes).  Shorter slices are not worth it — every slice costs each
        // tile an epilogue of atomic adds and a pipeline fill (measured at 16384 items: 183-k-block slices, 133 ms vs 114 ms).
        const int64_t l2_kb = (48ll << 20) / (16ll * std::max(F, 1));
        if (l2_kb >= 2048) splits = std::max(splits, cdiv(n_kb, l2_kb));
        const int64_t kb_per_split = cdiv(n_kb, splits);
        splits = cdiv(n_kb, kb_per_split);
        dim3 grid((unsigned)n_tiles, (unsigned)splits);
        if (tcp256) {
            dim3 grid2((unsigned)(2 * n_tiles), (unsigned)splits);
            CK(cudaFuncSetAttribute(k_pair_tc4, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
            k_pair_tc4<<<grid2, TC4_THREADS, TC_SMEM_BYTES, ctx->stream>>>(ctx->bits.as<uint32_t>(), ctx->pitch_words, F, kb_per_split,
                                                                         ctx->pair_tiles.as<int2>(), P);
        } else if (tc512) {
            dim3 grid2((unsigned)(2 * n_tiles), (unsigned)splits);   // one CTA pair (2-cluster) per tile
            k_pair_tc3<<<grid2, TC3_THREADS, TC3_SMEM_BYTES, ctx->stream>>>(ctx->bits.as<uint32_t>(), ctx->pitch_words, F, kb_per_split,
                                                                          ctx->pair_tiles.as<int2>(), P);
        } else if (tc256) {
            CK(cudaFuncSetAttribute(k_pair_tc2, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
            k_pair_tc2<<<grid, TC2_THREADS, TC_SMEM_BYTES, ctx->stream>>>(ctx->bits.as<uint32_t>(), ctx->pitch_words, F, kb_per_split,
                                                                         ctx->pair_tiles.as<int2>(), P);
        } else {
            CK(cudaFuncSetAttribute(k_pair_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
            k_pair_tc<<<grid, TC_THREADS, TC_SMEM_BYTES, ctx->stream>>>(ctx->bits.as<uint32_t>(), ctx->pitch_words, F, kb_per_split,
                                                                       ctx->pair_tiles.as<int2>(), P);
        }
        LAUNCH_CHECK();
    } else {
        // split the row axis so that tiles x splits covers the machine a few times over
        const int64_t chunks = ctx->pitch_words / PAIR_KW;
        int64_t splits = std::max<int64_t>(1, std::min<int64_t>(chunks, cdiv((int64_t)ctx->sm_count * 8, n_tiles)));
        const int64_t words_per_split = cdiv(chunks, splits) * PAIR_KW;
        splits = cdiv(ctx->pitch_words, words_per_split);
        dim3 grid((unsigned)n_tiles, (unsigned)splits);
        k_pair_popc<<<grid, 256, 0, ctx->stream>>>(ctx->bits.as<uint32_t>(), ctx->pitch_words, F, words_per_split,
                                                   ctx->pair_tiles.as<int2>(), P, ctx->weighted ? ctx->chunk_shift.as<uint8_t>() : nullptr);
        LAUNCH_CHECK();
    }
    ctx->tend();
    ctx->pair_ops += 2.0 * (double)ctx->n_rows * (double)F * (double)F;
    if (ctx->nccl_comm || ctx->xchg) {
        // the lower triangle carries nothing: fold the bottom rows' counts into it and exchange the top half only
        const int h = (F + 1) / 2;
        if (F - h > 0) { k_pair_fold<false><<<F - h, 256, 0, ctx->stream>>>(P, F); LAUNCH_CHECK(); }
        if (int rc = exchange(ctx, P, (int64_t)h * F)) return rc;
        if (F - h > 0) { k_pair_fold<true><<<F - h, 256, 0, ctx->stream>>>(P, F); LAUNCH_CHECK(); }
    }
    ctx->tbegin(T_THRESH);
    CK(ctx->pair_rowcnt.reserve(sizeof(int32_t) * (size_t)F, ctx->stream));
    CK(ctx->pair_rowoff.reserve(sizeof(int32_t) * (size_t)F, ctx->stream));
    const unsigned rblocks = (unsigned)cdiv(F, 8);
    k_pair_select<false><<<rblocks, 256, 0, ctx->stream>>>(P, F, min_count, ctx->pair_rowcnt.as<int32_t>(), nullptr, nullptr, nullptr, nullptr);
    LAUNCH_CHECK();
    if (int rc = scan_exclusive(ctx, ctx->pair_rowcnt.as<int32_t>(), ctx->pair_rowoff.as<int32_t>(), F)) return rc;
    int64_t n2 = 0;
    if (int rc = scan_total(ctx, ctx->pair_rowcnt.as<int32_t>(), ctx->pair_rowoff.as<int32_t>(), F, &n2)) return rc;
    L2.n = n2;
    CK(L2.sets.reserve(sizeof(int32_t) * 2 * (size_t)std::max<int64_t>(n2, 1), ctx->stream));
    CK(L2.sub.reserve(sizeof(int32_t) * 2 * (size_t)std::max<int64_t>(n2, 1), ctx->stream));
    CK(L2.counts.reserve(sizeof(int32_t) * (size_t)std::max<int64_t>(n2, 1), ctx->stream));
    if (n2) {
        k_pair_select<true><<<rblocks, 256, 0, ctx->stream>>>(P, F, min_count, nullptr, ctx->pair_rowoff.as<int32_t>(),
                                                              L2.sets.as<int32_t>(), L2.sub.as<int32_t>(), L2.counts.as<int32_t>());
        LAUNCH_CHECK();
    }
    if (int rc = make_child_index(ctx, ctx->levels[1], L2)) return rc;
    ctx->tend();
    if (n2 >= 3) {
        if (int rc = enqueue_sibling_slots(ctx, L2, ctx->levels[1], n2, nullptr, true)) return rc;
    }
    *n_frequent_out = n2;
    return FA_OK;
}

// ---- K5 ------------------------------------------------------------------------------------
int fa_gen_candidates(fa_ctx* ctx, int32_t k, int64_t* n_groups_out, int64_t* n_candidates_out) {
    if (!ctx) return FA_EINVAL;
    if (k < 3 || (size_t)k != ctx->levels.size()) return ctx->fail(FA_EINVAL, "fa_gen_candidates(k) needs levels 2..k-1 and no level k yet");
    if ((n_groups_out == nullptr) != (n_candidates_out == nullptr)) return ctx->fail(FA_EINVAL, "fa_gen_candidates: pass both outputs or neither");
    if (k >= FA_MAX_LEVELS) return ctx->fail(FA_ELIMIT, "more than 63 itemset levels");
    CK(cudaSetDevice(ctx->device));
    Level& P = ctx->levels[k - 1];
    const int64_t n = P.n;
    const int m = k - 1;
    ctx->cand_k = k;
    ctx->n_groups = ctx->n_cands = ctx->ng_bound = ctx->nc_bound = ctx->n_slots = 0;
    if (n_groups_out) *n_groups_out = *n_candidates_out = 0;
    if (n == 0) return FA_OK;
    if (n > 0x7fffffff / 2) return ctx->fail(FA_ELIMIT, "level too large");
    Level& Q = ctx->levels[k - 2];  // parents of level k-1: their child ranges delimit the sibling families
    const int32_t* sets = P.sets.as<int32_t>();
    const int32_t* sub = P.sub.as<int32_t>();
    const int32_t* pchild = Q.child.as<int32_t>();
    // one slot per (x, following sibling) pair: counted and scanned right after level k-1 was built, and in the level
    // loop already read by the host together with that level's size (no sync here); otherwise one read-back now
    int64_t n_slots = 0;
    if (ctx->pre_level == k && ctx->pre_read) {
        n_slots = ctx->pre_slots;
    } else {
        if (ctx->pre_level != k) {
            if (int rc = enqueue_sibling_slots(ctx, P, Q, n, nullptr, true)) return rc;
        }
        CK(cudaStreamSynchronize(ctx->stream));
        n_slots = (int32_t)ctx->h_scalars[PRE_SLOT];
    }
    // (the slot offsets in c_off were scanned over a bound >= n: the entries past row n all equal the total)
    ctx->pre_level = 0;
    ctx->pre_read = false;
    ctx->tbegin(T_CANDGEN);
    CK(ctx->c_gflag.reserve(4 * (size_t)n, ctx->stream));
    CK(ctx->c_gpos.reserve(4 * (size_t)n, ctx->stream));
    int32_t* cnt = ctx->c_cnt.as<int32_t>();
    int32_t* soff = ctx->c_off.as<int32_t>();
    if (n_slots < 0 || n_slots * k > 0x7fffffffll * 4 || n_slots > 0x7ffffff0) return ctx->fail(FA_ELIMIT, "too many sibling pairs in one level");
    int32_t* lvl = ctx->scalars.as<int32_t>() + 52;  // [0] groups, [1] candidates of this level, [2] its frequent itemsets (K4's call)
    ctx->n_slots = n_slots;
    if (n_slots > 0) {
        CK(ctx->c_sflag.reserve(4 * (size_t)(n_slots + 1), ctx->stream));
        CK(ctx->c_spos.reserve(4 * (size_t)(n_slots + 1), ctx->stream));
        CK(ctx->c_slotx.reserve(4 * (size_t)n_slots, ctx->stream));
        CK(ctx->c_slotsub.reserve(4 * (size_t)n_slots * (size_t)k, ctx->stream));
        int32_t* sflag = ctx->c_sflag.as<int32_t>();
        int32_t* spos = ctx->c_spos.as<int32_t>();
        // lanes per x from the average number of following siblings (deep levels: 1-2, level 3: hundreds)
        const int64_t avg_sib = n_slots / n;
        if (avg_sib >= 12)
            k_join_candidates<32><<<(unsigned)cdiv(n, 8), 256, 0, ctx->stream>>>(sets, sub, pchild, (int)n, m, soff, sflag,
                                                                                ctx->c_slotx.as<int32_t>(), ctx->c_slotsub.as<int32_t>());
        else if (avg_sib >= 4)
            k_join_candidates<16><<<(unsigned)cdiv(n, 16), 256, 0, ctx->stream>>>(sets, sub, pchild, (int)n, m, soff, sflag,
                                                                                 ctx->c_slotx.as<int32_t>(), ctx->c_slotsub.as<int32_t>());
        else  // (4 lanes instead of 8 at the sparse levels: K5 1.70 -> 1.55 ms per configs[1] pass)
            k_join_candidates<4><<<(unsigned)cdiv(n, 64), 256, 0, ctx->stream>>>(sets, sub, pchild, (int)n, m, soff, sflag,
                                                                                ctx->c_slotx.as<int32_t>(), ctx->c_slotsub.as<int32_t>());
        LAUNCH_CHECK();
        CK(cudaMemsetAsync(sflag + n_slots, 0, 4, ctx->stream));
        if (int rc = scan_exclusive(ctx, sflag, spos, n_slots + 1)) return rc;
        k_x_counts<<<(unsigned)cdiv(n, 256), 256, 0, ctx->stream>>>(soff, spos, (int)n, cnt, ctx->c_gflag.as<int32_t>());
        LAUNCH_CHECK();
        if (int rc = scan_exclusive(ctx, ctx->c_gflag.as<int32_t>(), ctx->c_gpos.as<int32_t>(), n)) return rc;
        // Everything below is sized by what the host knows without asking the device: candidates <= slots, groups <= n.
        // The exact totals are left in lvl[] by k_write_groups2 for the kernels that need them.
        CK(ctx->c_gx.reserve(4 * (size_t)n, ctx->stream));
        CK(ctx->c_goff.reserve(4 * (size_t)(n + 1), ctx->stream));
        CK(ctx->c_ys.reserve(4 * (size_t)n_slots, ctx->stream));
        CK(ctx->c_cx.reserve(4 * (size_t)n_slots, ctx->stream));
        CK(ctx->c_csub.reserve(4 * (size_t)n_slots * (size_t)k, ctx->stream));
        k_compact_slots<<<(unsigned)cdiv(n_slots, 256), 256, 0, ctx->stream>>>(sets, m, soff, ctx->c_sflag.as<int32_t>(),
                                                                              ctx->c_spos.as<int32_t>(), ctx->c_slotx.as<int32_t>(),
                                                                              ctx->c_slotsub.as<int32_t>(), n_slots,
                                                                              ctx->c_ys.as<int32_t>(), ctx->c_cx.as<int32_t>(),
                                                                              ctx->c_csub.as<int32_t>());
        LAUNCH_CHECK();
        k_write_groups2<<<(unsigned)cdiv(n, 256), 256, 0, ctx->stream>>>(cnt, soff, ctx->c_spos.as<int32_t>(),
                                                                        ctx->c_gpos.as<int32_t>(), (int)n, n_slots,
                                                                        ctx->c_gx.as<int32_t>(), ctx->c_goff.as<int32_t>(), lvl);
        LAUNCH_CHECK();
        ctx->ng_bound = n;
        ctx->nc_bound = n_slots;
        if (n_groups_out || !ctx->k4_groups || ctx->k5_wait) {  // the caller wants the totals now (or the tile schedule is sized by them)
            uint32_t h[2];
            if (int rc = read_scalars(ctx, {{lvl, 2}}, h)) return rc;
            ctx->n_groups = ctx->ng_bound = (int32_t)h[0];
            ctx->n_cands = ctx->nc_bound = (int32_t)h[1];
                }
        const int64_t ng = ctx->ng_bound, nc = ctx->nc_bound;
        unsigned long long* d_loaded = ctx->scalars.as<unsigned long long>() + 16;  // rows K4 will request per chunk (read back with the level's totals)
        if (nc > 0 && !ctx->k4_groups) {
            // K4's per-level schedule: sibling tiles (<= 4 groups of one family) and their merged extension streams.  The
            // tile count stays on the device (scalars word 48); the host sizes everything by its bound, the group count.
            CK(ctx->c_tflag.reserve(4 * (size_t)(ng + 1), ctx->stream));
            CK(ctx->c_tpos.reserve(4 * (size_t)(ng + 1), ctx->stream));
            CK(ctx->c_tdesc.reserve(48 * (size_t)ng, ctx->stream));
            CK(ctx->c_tstream.reserve(4 * (size_t)nc, ctx->stream));
            CK(cudaMemsetAsync(d_loaded, 0, 8, ctx->stream));
            k_tile_flags<<<(unsigned)cdiv(ng, 256), 256, 0, ctx->stream>>>(sub, pchild, m, ctx->c_gx.as<int32_t>(), ng, ctx->c_tflag.as<int32_t>());
            LAUNCH_CHECK();
            if (int rc = scan_exclusive(ctx, ctx->c_tflag.as<int32_t>(), ctx->c_tpos.as<int32_t>(), ng)) return rc;
            int4* td = ctx->c_tdesc.as<int4>();
            k_tile_build<<<(unsigned)cdiv(ng, 128), 128, 0, ctx->stream>>>(sets, m, ctx->c_gx.as<int32_t>(), ctx->c_goff.as<int32_t>(),
                                                                         ctx->c_ys.as<int32_t>(), ctx->c_tflag.as<int32_t>(),
                                                                         ctx->c_tpos.as<int32_t>(), ng, COUNT_GT, td, td + ng, td + 2 * ng,
                                                                         ctx->c_tstream.as<int32_t>(), ctx->scalars.as<int32_t>() + 48, d_loaded);
            LAUNCH_CHECK();
        } else if (nc > 0) {
            CK(ctx->c_gdesc.reserve(16 * (size_t)ng, ctx->stream));
            CK(cudaMemsetAsync(d_loaded, 0, 8, ctx->stream));
            k_group_desc<<<(unsigned)cdiv(ng, 256), 256, 0, ctx->stream>>>(sets, m, ctx->c_gx.as<int32_t>(), ctx->c_goff.as<int32_t>(), lvl, COUNT_G,
                                                                          ctx->c_gdesc.as<int4>(), d_loaded);
            LAUNCH_CHECK();
        }
    }
    ctx->tend();
    if (n_groups_out) {
        *n_groups_out = ctx->n_groups;
        *n_candidates_out = ctx->n_cands;
    }
    return FA_OK;
}

// ---- resume at the rule stage ---------------------------------------------------------------
int fa_put_level(fa_ctx* ctx, int32_t k, const int32_t* sets, const int32_t* counts, int64_t n) {
    if (!ctx) return FA_EINVAL;
    if (k < 1 || k >= FA_MAX_LEVELS) return ctx->fail(FA_EINVAL, "fa_put_level: k out of range");
    if (n < 0 || n > 0x7fffffff / k) return ctx->fail(FA_ELIMIT, "fa_put_level: level too large");
    if (n > 0 && (!counts || (k >= 2 && !sets))) return ctx->fail(FA_EINVAL, "fa_put_level: null sets/counts");
    CK(cudaSetDevice(ctx->device));
    ctx->sup_min = -1;  // supports installed by the host are not known to be anti-monotone: fa_sort_rules then sorts on all 64 key bits
    if (k == 1) {
        drop_levels(ctx, 1);
        ctx->have_bits = false;  // itemsets without transactions: the mining entry points need a new load + fa_build_bitmaps
        ctx->F = (int32_t)n;
        ctx->levels.resize(2);
        Level& L1 = ctx->levels[1];
        L1.k = 1;
        L1.n = n;
        CK(L1.sets.reserve(sizeof(int32_t) * (size_t)std::max<int64_t>(n, 1), ctx->stream));
        CK(L1.counts.reserve(sizeof(int32_t) * (size_t)std::max<int64_t>(n, 1), ctx->stream));
        if (n) {
            k_iota<<<(unsigned)cdiv(n, 256), 256, 0, ctx->stream>>>(L1.sets.as<int32_t>(), (int)n);
            LAUNCH_CHECK();
            CK(cudaMemcpyAsync(L1.counts.p, counts, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));  // the caller's buffer may be pageable and short-lived
        }
        return FA_OK;
    }
    if ((size_t)k != ctx->levels.size()) return ctx->fail(FA_EINVAL, "fa_put_level: levels must be installed in order k = 1, 2, ...");
    drop_levels(ctx, (size_t)k);
    ctx->levels.resize((size_t)k + 1);
    Level& P = ctx->levels[k - 1];
    Level& L = ctx->levels[k];
    L.k = k;
    L.n = n;
    const size_t cells = (size_t)std::max<int64_t>(n, 1) * (size_t)k;
    CK(L.sets.reserve(4 * cells, ctx->stream));
    CK(L.sub.reserve(4 * cells, ctx->stream));
    CK(L.counts.reserve(4 * (size_t)std::max<int64_t>(n, 1), ctx->stream));
    if (n) {
        CK(cudaMemcpyAsync(L.sets.p, sets, 4 * (size_t)n * (size_t)k, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(L.counts.p, counts, 4 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemsetAsync(ctx->errflag.p, 0, 4, ctx->stream));
        k_find_subsets<<<(unsigned)cdiv(n * k, 256), 256, 0, ctx->stream>>>(L.sets.as<int32_t>(), n, k, P.sets.as<int32_t>(), P.n, ctx->F,
                                                                            L.sub.as<int32_t>(), ctx->errflag.as<int32_t>());
        LAUNCH_CHECK();
    }
    if (int rc = make_child_index(ctx, P, L)) return rc;
    uint32_t bad = 0;
    if (n) { if (int rc = read_scalars(ctx, {{ctx->errflag.p, 1}}, &bad)) return rc; }
    else CK(cudaStreamSynchronize(ctx->stream));
    if (bad) {
        drop_levels(ctx, (size_t)k);
        return ctx->fail(FA_EINVAL, (bad & 1) ? "fa_put_level: rows must hold ascending ranks in [0, F) and be strictly ascending in lexicographic order"
                                              : "fa_put_level: an itemset has a subset that is not in the level below (not downward closed)");
    }
    return FA_OK;
}

// ---- K4 ------------------------------------------------------------------------------------
int fa_count_candidates(fa_ctx* ctx, int32_t k, int32_t min_count, int64_t* n_frequent_out) {
    if (!ctx) return FA_EINVAL;
    if (k != ctx->cand_k || (size_t)k != ctx->levels.size()) return ctx->fail(FA_EINVAL, "fa_count_candidates(k) must follow fa_gen_candidates(k)");
    if (!n_frequent_out) return ctx->fail(FA_EINVAL, "n_frequent_out is null");
    CK(cudaSetDevice(ctx->device));
    const int m = k - 1;
    // exact totals when fa_gen_candidates was asked for them, else the host's bounds (the device holds the totals)
    const int64_t ncb = ctx->nc_bound, ngb = ctx->ng_bound;
    if (min_count <= 0) ctx->sup_min = -1;
    else if (ctx->sup_min > 0) ctx->sup_min = std::min<int64_t>(ctx->sup_min, min_count);
    ctx->levels.resize(k + 1);
    Level& P = ctx->levels[k - 1];
    Level& L = ctx->levels[k];
    L.k = k;
    L.n = 0;
    *n_frequent_out = 0;
    ctx->cand_k = 0;
    if (ncb == 0) {
        CK(L.sets.reserve(4 * (size_t)k, ctx->stream));
        CK(L.sub.reserve(4 * (size_t)k, ctx->stream));
        CK(L.counts.reserve(4, ctx->stream));
        return make_child_index(ctx, P, L);
    }
    int32_t* lvl = ctx->scalars.as<int32_t>() + 52;
    CK(ctx->c_counts.reserve(4 * (size_t)ncb, ctx->stream));
    int32_t* counts = ctx->c_counts.as<int32_t>();
    CK(cudaMemsetAsync(counts, 0, 4 * (size_t)ncb, ctx->stream));
    const int64_t pitch_q = ctx->pitch_words / 4;
    const int64_t n_chunks = cdiv(pitch_q, 32 * COUNT_U);
    CK(ctx->t_alive.reserve(4 * (size_t)ctx->pitch_words, ctx->stream));
    CK(cudaMemsetAsync(ctx->t_alive.p, 0, 4 * (size_t)ctx->pitch_words, ctx->stream));
    ctx->tbegin(T_COUNT);
    if (!ctx->k4_groups) {
        // persistent grid (the tile count is known to the device only; its bound, the group count, sizes the grid):
        // warps take units in order, unit = (row chunk, run of COUNT_GT tiles), chunk-major
        const int64_t ng = ngb;  // exact on this path (fa_gen_candidates read it back)
        const int64_t units_bound = cdiv(ng, COUNT_GT) * n_chunks;
        const unsigned blocks = (unsigned)std::max<int64_t>(1, std::min<int64_t>(cdiv(units_bound, 8), (int64_t)ctx->sm_count * K4_MIN_BLOCKS));
        unsigned long long* d_next = ctx->scalars.as<unsigned long long>() + 20;  // the kernel's unit counter
        CK(cudaMemsetAsync(d_next, 0, 8, ctx->stream));
        const int smem = 8 * 4 * TILE_SMEM_WORDS(m);
        CK(cudaFuncSetAttribute(k_count_tiles<COUNT_U, COUNT_GT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 4 * TILE_SMEM_WORDS(64)));
        const int4* td = ctx->c_tdesc.as<int4>();
        k_count_tiles<COUNT_U, COUNT_GT><<<blocks, 256, smem, ctx->stream>>>(
            ctx->bits.as<uint4>(), pitch_q, P.sets.as<int32_t>(), m, td, td + ng, td + 2 * ng, ctx->c_tstream.as<int32_t>(),
            ctx->scalars.as<int32_t>() + 48, n_chunks, counts, ctx->t_alive.as<uint32_t>(),
            ctx->weighted ? ctx->chunk_shift.as<uint8_t>() : nullptr, d_next);
        LAUNCH_CHECK();
    } else {
        // one warp per unit, blocks dispatched in unit order => chunk-major sweep.  The grid covers the bound on the
        // groups; the kernel places its units by the exact count (lvl[0]) and the warps past the last unit leave at once.
        const int64_t units = cdiv(ngb, COUNT_G) * n_chunks;
        const int W = ctx->k4_warps;
        if (cdiv(units, W) > 0x7fffffff) return ctx->fail(FA_ELIMIT, "fa_count_candidates: too many (group run, row chunk) units in one level");
        const unsigned blocks = (unsigned)cdiv(units, W);
        static_assert(PITCH_ALIGN % (32 * COUNT_U * 4) == 0, "K4 reads whole chunks without bounds checks");
        const int smem = W * 4 * COUNT2_SMEM_WORDS(m);
        const uint8_t* shifts = ctx->weighted ? ctx->chunk_shift.as<uint8_t>() : nullptr;
#define FA_LAUNCH_K4(WIDE_, W_)                                                                                                          \
    do {                                                                                                                                 \
        CK(cudaFuncSetAttribute(k_count_groups2<COUNT_U, COUNT_G, WIDE_, W_>, cudaFuncAttributeMaxDynamicSharedMemorySize,               \
                                W_ * 4 * COUNT2_SMEM_WORDS(64)));                                                                        \
        k_count_groups2<COUNT_U, COUNT_G, WIDE_, W_><<<blocks, 32 * W_, smem, ctx->stream>>>(                                            \
            ctx->bits.as<uint4>(), pitch_q, P.sets.as<int32_t>(), m, ctx->c_gdesc.as<int4>(), ctx->c_goff.as<int32_t>(),                 \
            ctx->c_ys.as<int32_t>(), lvl, n_chunks, counts, ctx->t_alive.as<uint32_t>(), shifts);                                        \
    } while (0)
        if (m > 32) {
            if (W == 1) FA_LAUNCH_K4(true, 1); else if (W == 2) FA_LAUNCH_K4(true, 2); else if (W == 4) FA_LAUNCH_K4(true, 4); else FA_LAUNCH_K4(true, 8);
        } else {
            if (W == 1) FA_LAUNCH_K4(false, 1); else if (W == 2) FA_LAUNCH_K4(false, 2); else if (W == 4) FA_LAUNCH_K4(false, 4); else FA_LAUNCH_K4(false, 8);
        }
#undef FA_LAUNCH_K4
        LAUNCH_CHECK();
    }
    ctx->tend();
    const size_t k4_timer = ctx->pending.size();
    ctx->count_launches++;
    if (int rc = exchange(ctx, counts, ncb)) return rc;
    ctx->tbegin(T_THRESH);
    // Threshold, the new level's rows, its parents' child ranges and the first step of the NEXT level's K5 are all queued
    // behind K4 with sizes taken from the device; the host then waits ONCE per level for: groups, candidates, frequent
    // candidates, alive rows, rows K4 requested, and the next level's slot count.
    CK(ctx->c_flag.reserve(4 * (size_t)ncb, ctx->stream));
    CK(ctx->c_pos.reserve(4 * (size_t)ncb, ctx->stream));
    k_flag_ge<<<(unsigned)cdiv(ncb, 256), 256, 0, ctx->stream>>>(counts, ncb, lvl + 1, min_count, ctx->c_flag.as<int32_t>());
    LAUNCH_CHECK();
    if (int rc = scan_exclusive(ctx, ctx->c_flag.as<int32_t>(), ctx->c_pos.as<int32_t>(), ncb)) return rc;
    if (int rc = trim_prepare(ctx)) return rc;
    CK(L.sets.reserve(4 * (size_t)k * (size_t)ncb, ctx->stream));
    CK(L.sub.reserve(4 * (size_t)k * (size_t)ncb, ctx->stream));
    CK(L.counts.reserve(4 * (size_t)ncb, ctx->stream));
    k_compact_level<<<(unsigned)cdiv(ncb * k, 256), 256, 0, ctx->stream>>>(P.sets.as<int32_t>(), m, ctx->c_cx.as<int32_t>(),
                                                                      ctx->c_ys.as<int32_t>(), ctx->c_csub.as<int32_t>(), counts,
                                                                      ctx->c_flag.as<int32_t>(), ctx->c_pos.as<int32_t>(), ncb,
                                                                      L.sets.as<int32_t>(), L.sub.as<int32_t>(), L.counts.as<int32_t>(), lvl + 2);
    LAUNCH_CHECK();
    if (int rc = make_child_index(ctx, P, L, lvl + 2)) return rc;
    const bool pre = ncb >= k + 1 && ncb <= 0x7fffffff / 2;
    if (pre) {
        if (int rc = enqueue_sibling_slots(ctx, L, P, ncb, lvl + 2, false)) return rc;
    }
    int64_t nk = 0, ng = 0, nc = 0;
    int32_t alive_rows = 0;
    {
        uint32_t h[7];
        if (int rc = read_scalars(ctx, {{lvl, 3}, {ctx->t_off.as<int32_t>() + ctx->pitch_words, 1},
                                       {ctx->scalars.as<unsigned long long>() + 16, 2},
                                       {pre ? ctx->c_off.as<int32_t>() + ncb : lvl, 1}}, h)) return rc;
        ng = (int32_t)h[0];
        nc = (int32_t)h[1];
        nk = (int32_t)h[2];
        alive_rows = (int32_t)h[3];
        ctx->count_load_bytes += (double)((uint64_t)h[4] | ((uint64_t)h[5] << 32)) * 4.0 * (double)ctx->pitch_words;
        ctx->pre_slots = (int32_t)h[6];
        ctx->pre_read = pre;
    }
    ctx->n_groups = L.groups = ng;
    ctx->n_cands = L.cands = nc;
    ctx->count_bytes += (double)ctx->pitch_words * 4.0 * ((double)ng * m + (double)nc) + 4.0 * (double)nc;
    if (getenv("FA_DEBUG_LEVELS") && k4_timer >= 1 && k4_timer <= ctx->pending.size()) {
        float t = 0;
        cudaEventElapsedTime(&t, ctx->pending[k4_timer - 1].a, ctx->pending[k4_timer - 1].b);
        fprintf(stderr, "[level %d] groups=%lld cands=%lld slots=%lld pitch_words=%lld K4=%.3f ms  (%.1f GB algorithmic, %.2f TB/s)\n", k, (long long)ng,
                (long long)nc, (long long)ctx->n_slots, (long long)ctx->pitch_words, t, (double)ctx->pitch_words * 4.0 * ((double)ng * m + (double)nc) / 1e9,
                (double)ctx->pitch_words * 4.0 * ((double)ng * m + (double)nc) / 1e9 / t);
    }
    L.n = nk;
    if (int rc = trim_apply(ctx, L, alive_rows)) return rc;
    ctx->tend();
    if (nk < k + 1) { ctx->pre_level = 0; ctx->pre_read = false; }  // the level loop ends here (FastApriori.scala:110)
    *n_frequent_out = nk;
    return FA_OK;
}

// K5's totals for level k (the reference prints the group count, FastApriori.scala:114): valid once fa_count_candidates(k)
// has returned, whether or not fa_gen_candidates(k) was asked to wait for them.
int fa_level_counters(fa_ctx* ctx, int32_t k, int64_t* n_groups_out, int64_t* n_candidates_out) {
    if (!ctx) return FA_EINVAL;
    if (!n_groups_out || !n_candidates_out) return ctx->fail(FA_EINVAL, "fa_level_counters: null output");
    const bool ok = k >= 3 && (size_t)k < ctx->levels.size();
    *n_groups_out = ok ? ctx->levels[k].groups : 0;
    *n_candidates_out = ok ? ctx->levels[k].cands : 0;
    return FA_OK;
}

int fa_level_size(fa_ctx* ctx, int32_t k, int64_t* n_out) {
    if (!ctx || !n_out) return FA_EINVAL;
    *n_out = (k >= 1 && (size_t)k < ctx->levels.size()) ? ctx->levels[k].n : 0;
    return FA_OK;
}

int fa_get_level(fa_ctx* ctx, int32_t k, int32_t* sets_out, int32_t* counts_out) {
    if (!ctx) return FA_EINVAL;
    if (k < 1 || (size_t)k >= ctx->levels.size()) return ctx->fail(FA_EINVAL, "no such level");
    CK(cudaSetDevice(ctx->device));
    Level& L = ctx->levels[k];
    if (L.n == 0) return FA_OK;
    if (sets_out) CK(cudaMemcpyAsync(sets_out, L.sets.p, 4 * (size_t)k * (size_t)L.n, cudaMemcpyDeviceToHost, ctx->stream));
    if (counts_out) CK(cudaMemcpyAsync(counts_out, L.counts.p, 4 * (size_t)L.n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return FA_OK;
}

int fa_get_level_async(fa_ctx* ctx, int32_t k, int32_t* sets_out, int32_t* counts_out) {
    if (!ctx) return FA_EINVAL;
    if (k < 1 || (size_t)k >= ctx->levels.size()) return ctx->fail(FA_EINVAL, "no such level");
    CK(cudaSetDevice(ctx->device));
    Level& L = ctx->levels[k];
    if (L.n == 0) return FA_OK;
    // the copy stream waits for everything issued so far on the main stream, then copies behind its own queue
    CK(cudaEventRecord(ctx->copy_event, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->copy_event, 0));
    if (sets_out) CK(cudaMemcpyAsync(sets_out, L.sets.p, 4 * (size_t)k * (size_t)L.n, cudaMemcpyDeviceToHost, ctx->copy_stream));
    if (counts_out) CK(cudaMemcpyAsync(counts_out, L.counts.p, 4 * (size_t)L.n, cudaMemcpyDeviceToHost, ctx->copy_stream));
    return FA_OK;
}

int fa_wait_copies(fa_ctx* ctx) {
    if (!ctx) return FA_EINVAL;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->copy_stream));
    return FA_OK;
}

// ---- K6 ------------------------------------------------------------------------------------
int fa_gen_rules(fa_ctx* ctx, int64_t* n_raw_out) {
    if (!ctx) return FA_EINVAL;
    if (ctx->levels.size() < 2) return ctx->fail(FA_EINVAL, "fa_gen_rules before mining");
    if (!n_raw_out) return ctx->fail(FA_EINVAL, "n_raw_out is null");
    CK(cudaSetDevice(ctx->device));
    ctx->have_rules = ctx->have_sorted = false;
    if (int rc = build_level_tab(ctx)) return rc;
    const LevelTab& T = ctx->tab;
    const int64_t nr = ctx->n_raw;
    *n_raw_out = nr;
    ctx->tbegin(T_RULES);
    CK(ctx->r_conf.reserve(8 * (size_t)std::max<int64_t>(nr, 1), ctx->stream));
    for (int k = 2; k <= T.max_k; ++k) {
        Level& L = ctx->levels[k];
        Level& P = ctx->levels[k - 1];
        if (L.n == 0) continue;
        const int64_t n = L.n * k;
        k_gen_rules<<<(unsigned)cdiv(n, 256), 256, 0, ctx->stream>>>(L.counts.as<int32_t>(), L.sub.as<int32_t>(), L.n, k,
                                                                    P.counts.as<int32_t>(), ctx->r_conf.as<double>() + T.base[k]);
        LAUNCH_CHECK();
    }
    ctx->tend();
    ctx->have_rules = true;
    return FA_OK;
}

// ---- K7 ------------------------------------------------------------------------------------
int fa_prune_rules(fa_ctx* ctx, int64_t* n_kept_out) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_rules) return ctx->fail(FA_EINVAL, "fa_prune_rules before fa_gen_rules");
    if (!n_kept_out) return ctx->fail(FA_EINVAL, "n_kept_out is null");
    CK(cudaSetDevice(ctx->device));
    const LevelTab& T = ctx->tab;
    const int64_t nr = ctx->n_raw;
    ctx->have_sorted = false;
    ctx->cut_before.assign(FA_MAX_LEVELS, 0);
    ctx->cut_after.assign(FA_MAX_LEVELS, 0);
    ctx->n_kept = 0;
    *n_kept_out = 0;
    if (nr == 0) return FA_OK;
    ctx->tbegin(T_RULES);
    CK(ctx->r_kept.reserve((size_t)nr, ctx->stream));
    CK(ctx->r_list.reserve(4 * (size_t)nr, ctx->stream));
    uint8_t* kept = ctx->r_kept.as<uint8_t>();
    const double* conf = ctx->r_conf.as<double>();
    int32_t* list = ctx->r_list.as<int32_t>();
    // device scalars: [0] number of selected rule ids, [8..] kept rules per level (unsigned long long each)
    int32_t* d_nsel = ctx->scalars.as<int32_t>();
    CK(ctx->r_tmp64.reserve(8 * FA_MAX_LEVELS, ctx->stream));
    unsigned long long* d_kept = ctx->r_tmp64.as<unsigned long long>();
    CK(cudaMemsetAsync(d_kept, 0, 8 * FA_MAX_LEVELS, ctx->stream));
    for (int k = 2; k <= T.max_k; ++k) {
        const int64_t n = ctx->levels[k].n * k;
        if (n == 0) continue;
        if (k == 2) CK(cudaMemsetAsync(kept + T.base[2], 1, (size_t)n, ctx->stream));  // AssociationRules.scala:149-151
        else {
            k_prune_rules<<<(unsigned)cdiv(n, 256), 256, 0, ctx->stream>>>(ctx->levels[k].n, k, ctx->levels[k].sub.as<int32_t>(), conf + T.base[k],
                                                                          conf + T.base[k - 1], kept + T.base[k - 1],
                                                                          kept + T.base[k], d_kept + k);
            LAUNCH_CHECK();
        }
    }
    // kept rule ids of all levels, ascending (one select over the whole id space, one host sync)
    cub::CountingInputIterator<int32_t> ids(0);
    size_t bytes = 0;
    CK(cub::DeviceSelect::Flagged(nullptr, bytes, ids, kept, list, d_nsel, (int)nr, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceSelect::Flagged(ctx->scratch.p, bytes, ids, kept, list, d_nsel, (int)nr, ctx->stream));
    uint32_t hw[1 + 2 * FA_MAX_LEVELS];
    if (int rc = read_scalars(ctx, {{d_nsel, 1}, {d_kept, 2 * FA_MAX_LEVELS}}, hw)) return rc;
    unsigned long long h_kept[FA_MAX_LEVELS];
    memcpy(h_kept, hw + 1, sizeof h_kept);
    int64_t total = (int32_t)hw[0];
    for (int k = 2; k <= T.max_k; ++k) {
        const int64_t n = ctx->levels[k].n * k;
        ctx->cut_before[k - 1] = n;   // indexed by antecedent size, like the reference's printout
        ctx->cut_after[k - 1] = k == 2 ? n : (int64_t)h_kept[k];
    }
    ctx->tend();
    ctx->n_kept = total;
    *n_kept_out = total;
    return FA_OK;
}

// Host-only helper of fa_sort_rules (exported so that the argument below can be tested exhaustively without a GPU).
int fa_conf_key_bits(int64_t sup_min, int64_t sup_max, int32_t* lo_bit_out, int32_t* hi_bit_out) {
    if (!lo_bit_out || !hi_bit_out) return FA_EINVAL;
    *lo_bit_out = 0;
    *hi_bit_out = 64;
    if (sup_min <= 0 || sup_max < sup_min) return FA_OK;  // no usable bounds: all 64 bits
    const double lo = (double)sup_min / (double)sup_max, one = 1.0;
    uint64_t blo, bone;
    memcpy(&blo, &lo, 8);
    memcpy(&bone, &one, 8);
    const uint64_t x = blo ^ bone;
    int hi = 1;
    while (hi < 64 && (x >> hi)) ++hi;
    int len = 1;
    while (len < 62 && (1ll << len) <= sup_max) ++len;
    const int lo_bit = std::max(0, 52 - (2 * len + 2));
    *lo_bit_out = lo_bit;
    *hi_bit_out = std::max(hi, lo_bit + 1);
    return FA_OK;
}

int fa_rule_cut(fa_ctx* ctx, int32_t level, int64_t* before_out, int64_t* after_out) {
    if (!ctx || !before_out || !after_out) return FA_EINVAL;
    const bool ok = level >= 1 && (size_t)level < ctx->cut_before.size();
    *before_out = ok ? ctx->cut_before[level] : 0;
    *after_out = ok ? ctx->cut_after[level] : 0;
    return FA_OK;
}

// ---- K8 (+ the K9 bucket index) ------------------------------------------------------------
int fa_sort_rules(fa_ctx* ctx, const int32_t* rank_to_item_int) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_rules || ctx->cut_before.empty()) return ctx->fail(FA_EINVAL, "fa_sort_rules before fa_prune_rules");
    if (ctx->F && !rank_to_item_int) return ctx->fail(FA_EINVAL, "rank_to_item_int is null");
    CK(cudaSetDevice(ctx->device));
    const int64_t R = ctx->n_kept;
    const int F = ctx->F;
    const int64_t n_nodes = ctx->n_nodes;
    CK(ctx->m_aoff.reserve(4 * (size_t)(n_nodes + 2), ctx->stream));
    if (R == 0) {
        CK(cudaMemsetAsync(ctx->m_aoff.p, 0, 4 * (size_t)(n_nodes + 2), ctx->stream));
        ctx->have_sorted = true;
        return FA_OK;
    }
    if (R > 0x7fffffff) return ctx->fail(FA_ELIMIT, "too many rules");
    ctx->tbegin(T_SORT);
    const unsigned rb = (unsigned)cdiv(R, 256);
    CK(ctx->r_itemint.reserve(4 * (size_t)F, ctx->stream));
    CK(cudaMemcpyAsync(ctx->r_itemint.p, rank_to_item_int, 4 * (size_t)F, cudaMemcpyHostToDevice, ctx->stream));
    CK(ctx->r_key32.reserve(4 * (size_t)R * 2, ctx->stream));
    CK(ctx->r_key64.reserve(8 * (size_t)R * 2, ctx->stream));
    CK(ctx->r_val.reserve(4 * (size_t)R, ctx->stream));
    CK(ctx->r_val2.reserve(4 * (size_t)R, ctx->stream));
    CK(ctx->r_order.reserve(4 * (size_t)R, ctx->stream));
    uint32_t* k32a = ctx->r_key32.as<uint32_t>();
    uint32_t* k32b = k32a + R;
    uint64_t* k64a = ctx->r_key64.as<uint64_t>();
    uint64_t* k64b = k64a + R;
    int32_t* list = ctx->r_list.as<int32_t>();
    int32_t* v1 = ctx->r_val.as<int32_t>();
    int32_t* order = ctx->r_order.as<int32_t>();
    size_t bytes = 0;
    // pass 1 (secondary key): consequent item id ascending; rule ids are already ascending, LSD radix is stable
    k_rule_key_item<<<rb, 256, 0, ctx->stream>>>(ctx->tab, list, R, ctx->r_itemint.as<int32_t>(), k32a);
    LAUNCH_CHECK();
    // (keys are item ^ 0x80000000: when no item id is negative the top bit is constant and only bit_length(max id) low
    // bits differ, so the radix sort needs 1-2 digit passes instead of 4)
    int item_bits = 32;
    {
        int32_t lo = 0, hi = 0;
        for (int i = 0; i < F; ++i) { lo = std::min(lo, rank_to_item_int[i]); hi = std::max(hi, rank_to_item_int[i]); }
        if (lo >= 0) { item_bits = 1; while (item_bits < 31 && (1ll << item_bits) <= hi) ++item_bits; }
    }
    CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, k32a, k32b, list, v1, (int)R, 0, item_bits, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, bytes, k32a, k32b, list, v1, (int)R, 0, item_bits, ctx->stream));
    // pass 2 (primary key): confidence descending
    k_rule_key_conf<<<rb, 256, 0, ctx->stream>>>(v1, R, ctx->r_conf.as<double>(), k64a);
    LAUNCH_CHECK();
    // Which of the 64 key bits can differ between two rules?  Every confidence is fl(a / b) with integer supports
    // sup_min <= a <= b <= sup_max, so (1) it lies in [fl(sup_min / sup_max), 1] — rounding is monotone — and the key
    // bits above the first bit where those two bounds differ are the same in all keys; (2) two different confidences
    // are at least 1 / sup_max^2 apart, so they already differ in the top 2 * bit_length(sup_max) + 2 mantissa bits
    // (a bucket of the dropped low bits is at most 2^-(m+1) wide).  Sorting on that bit range alone gives the same
    // order — equal truncated keys are equal confidences, which the stable sort leaves in item order — in 6 digit
    // passes instead of 8 at configs[1].
    int conf_lo_bit = 0, conf_hi_bit = 64;
    fa_conf_key_bits(ctx->sup_min, ctx->sup_max, &conf_lo_bit, &conf_hi_bit);
    CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, k64a, k64b, v1, order, (int)R, conf_lo_bit, conf_hi_bit, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, bytes, k64a, k64b, v1, order, (int)R, conf_lo_bit, conf_hi_bit, ctx->stream));
    // K9 index: group by antecedent node (stable LSD radix keeps priority order inside each group)
    CK(ctx->m_lp.reserve(4 * (size_t)R, ctx->stream));
    CK(ctx->m_rec.reserve(8 * (size_t)R, ctx->stream));
    int32_t* prio = ctx->r_val2.as<int32_t>();
    k_rule_key_ant<<<rb, 256, 0, ctx->stream>>>(ctx->tab, order, R, k32a, prio);
    LAUNCH_CHECK();
    int bits = 1;
    while (bits < 32 && (1ll << bits) < n_nodes + 1) ++bits;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, k32a, k32b, prio, ctx->m_lp.as<int32_t>(), (int)R, 0, bits, ctx->stream));
    CK(ctx->scratch.reserve(bytes, ctx->stream));
    CK(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, bytes, k32a, k32b, prio, ctx->m_lp.as<int32_t>(), (int)R, 0, bits, ctx->stream));
    k_rule_records<<<rb, 256, 0, ctx->stream>>>(ctx->tab, order, ctx->m_lp.as<int32_t>(), R, ctx->m_rec.as<int2>());
    LAUNCH_CHECK();
    k_lower_bounds<<<(unsigned)cdiv(n_nodes + 1, 256), 256, 0, ctx->stream>>>(k32b, R, n_nodes, ctx->m_aoff.as<int32_t>());
    LAUNCH_CHECK();
    ctx->tend();
    ctx->have_sorted = true;
    return FA_OK;
}

int fa_rules_digest(fa_ctx* ctx, uint64_t* out) {
    if (!ctx) return FA_EINVAL;
    if (!out) return ctx->fail(FA_EINVAL, "out is null");
    if (!ctx->have_rules || !ctx->have_sorted) return ctx->fail(FA_EINVAL, "fa_rules_digest before fa_sort_rules");
    CK(cudaSetDevice(ctx->device));
    const int64_t R = ctx->n_kept;
    out[0] = (uint64_t)R;
    out[1] = out[2] = out[3] = 0;
    if (R == 0) return FA_OK;
    unsigned long long* d = ctx->scalars.as<unsigned long long>() + 8;  // 3 cells, clear of the words other entry points use
    CK(cudaMemsetAsync(d, 0, 24, ctx->stream));
    k_rule_digest<<<(unsigned)cdiv(R, 256), 256, 0, ctx->stream>>>(ctx->tab, ctx->r_order.as<int32_t>(), R, ctx->r_conf.as<double>(),
                                                                  ctx->r_itemint.as<int32_t>(), d);
    LAUNCH_CHECK();
    uint32_t h[6];
    if (int rc = read_scalars(ctx, {{d, 6}}, h)) return rc;
    for (int i = 0; i < 3; ++i) out[1 + i] = (uint64_t)h[2 * i] | ((uint64_t)h[2 * i + 1] << 32);
    return FA_OK;
}

int fa_rules_max_len(fa_ctx* ctx, int32_t* max_antecedent_out) {
    if (!ctx || !max_antecedent_out) return FA_EINVAL;
    *max_antecedent_out = ctx->have_rules ? std::max(ctx->tab.max_k - 1, 0) : 0;
    return FA_OK;
}

int fa_get_rules(fa_ctx* ctx, int32_t which, int32_t stride, int32_t* ant_out, int32_t* cons_out, double* conf_out) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_rules) return ctx->fail(FA_EINVAL, "fa_get_rules before fa_gen_rules");
    if (which == 1 && !ctx->have_sorted) return ctx->fail(FA_EINVAL, "fa_get_rules(1) before fa_sort_rules");
    CK(cudaSetDevice(ctx->device));
    const LevelTab& T = ctx->tab;
    const int64_t n = which ? ctx->n_kept : ctx->n_raw;
    if (n == 0) return FA_OK;
    if (stride < T.max_k - 1) return ctx->fail(FA_EINVAL, "stride smaller than the longest antecedent");
    std::vector<double> conf((size_t)ctx->n_raw);
    std::vector<int32_t> order;
    std::vector<std::vector<int32_t>> sets(T.max_k + 1);
    CK(cudaMemcpyAsync(conf.data(), ctx->r_conf.p, 8 * (size_t)ctx->n_raw, cudaMemcpyDeviceToHost, ctx->stream));
    if (which) {
        order.resize((size_t)n);
        CK(cudaMemcpyAsync(order.data(), ctx->r_order.p, 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    }
    for (int k = 2; k <= T.max_k; ++k) {
        sets[k].resize((size_t)ctx->levels[k].n * k);
        if (!sets[k].empty()) CK(cudaMemcpyAsync(sets[k].data(), ctx->levels[k].sets.p, 4 * sets[k].size(), cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    for (int64_t p = 0; p < n; ++p) {
        const int64_t rid = which ? order[p] : p;
        int k = 2;
        while (k < T.max_k && rid >= T.base[k + 1]) ++k;
        const int64_t local = rid - T.base[k], s = local / k;
        const int j = (int)(local - s * k);
        const int32_t* S = sets[k].data() + s * k;
        int w = 0;
        for (int t = 0; t < k; ++t)
            if (t != j) ant_out[p * stride + w++] = S[t];
        for (; w < stride; ++w) ant_out[p * stride + w] = -1;
        cons_out[p] = S[j];
        conf_out[p] = conf[rid];
    }
    return FA_OK;
}

// ---- K9 ------------------------------------------------------------------------------------
static int recommend_impl(fa_ctx* ctx, const int64_t* user_offsets, const int32_t* user_ranks, int64_t n_users, bool mapped,
                          const int32_t* id_to_rank, int32_t n_ids, int32_t* consequent_rank_out);

int fa_recommend(fa_ctx* ctx, const int64_t* user_offsets, const int32_t* user_ranks, int64_t n_users,
                 int32_t* consequent_rank_out) {
    return recommend_impl(ctx, user_offsets, user_ranks, n_users, false, nullptr, 0, consequent_rank_out);
}

int fa_recommend_ids(fa_ctx* ctx, const int64_t* user_offsets, const int32_t* user_item_ids, int64_t n_users,
                     const int32_t* id_to_rank, int32_t n_ids, int32_t* consequent_rank_out) {
    if (!ctx) return FA_EINVAL;
    if (n_ids < 0 || (n_ids && !id_to_rank)) return ctx->fail(FA_EINVAL, "fa_recommend_ids: bad id_to_rank");
    return recommend_impl(ctx, user_offsets, user_item_ids, n_users, true, id_to_rank, n_ids, consequent_rank_out);
}

static int recommend_impl(fa_ctx* ctx, const int64_t* user_offsets, const int32_t* user_ranks, int64_t n_users, bool mapped,
                          const int32_t* id_to_rank, int32_t n_ids, int32_t* consequent_rank_out) {
    if (!ctx) return FA_EINVAL;
    if (!ctx->have_sorted) return ctx->fail(FA_EINVAL, "fa_recommend before fa_sort_rules");
    if (n_users < 0 || (n_users && (!user_offsets || !consequent_rank_out))) return ctx->fail(FA_EINVAL, "bad user arguments");
    if (n_users == 0) return FA_OK;
    CK(cudaSetDevice(ctx->device));
    const int64_t nnz = user_offsets[n_users];
    if (user_offsets[0] != 0 || nnz < 0 || (nnz && !user_ranks)) return ctx->fail(FA_EINVAL, "bad user CSR");
    const int F = ctx->F;
    const size_t smem = MATCH_WARPS * sizeof(uint32_t) * (size_t)((F + 31) / 32 + 3 * MATCH_STACK_SMEM);
    if (smem > 200 * 1024) return ctx->fail(FA_ELIMIT, "too many frequent items for the basket mask in fa_recommend");
    CK(ctx->u_off.reserve(8 * (size_t)(n_users + 1), ctx->stream));
    CK(ctx->u_items.reserve(4 * (size_t)std::max<int64_t>(nnz, 1), ctx->stream));
    CK(ctx->u_out.reserve(4 * (size_t)n_users, ctx->stream));
    CK(cudaMemcpyAsync(ctx->u_off.p, user_offsets, 8 * (size_t)(n_users + 1), cudaMemcpyHostToDevice, ctx->stream));
    if (nnz) CK(cudaMemcpyAsync(ctx->u_items.p, user_ranks, 4 * (size_t)nnz, cudaMemcpyHostToDevice, ctx->stream));
    if (mapped && nnz) {  // dense ids -> ranks on the device (ids outside the table or mapped to -1 are not frequent)
        CK(ctx->id2rank_u.reserve(4 * (size_t)std::max(n_ids, 1), ctx->stream));
        if (n_ids) CK(cudaMemcpyAsync(ctx->id2rank_u.p, id_to_rank, 4 * (size_t)n_ids, cudaMemcpyHostToDevice, ctx->stream));
        k_map_ids<<<(unsigned)std::min<int64_t>(cdiv(nnz, 256), 1 << 20), 256, 0, ctx->stream>>>(ctx->u_items.as<int32_t>(), nnz,
                                                                                          ctx->id2rank_u.as<int32_t>(), n_ids);
        LAUNCH_CHECK();
    }
    ctx->tbegin(T_MATCH);
    ctx->distinct_baskets = -1;
    // what K9 matches: every user line, or — with at most 64 frequent items, where a basket is exactly a 64-bit mask — the
    // distinct baskets, each once, like the reference's grouping of identical baskets (AssociationRules.scala:51-58)
    const int64_t* k9_off = ctx->u_off.as<int64_t>();
    const int32_t* k9_items = ctx->u_items.as<int32_t>();
    int64_t k9_n = n_users;
    int32_t* k9_out = ctx->u_out.as<int32_t>();
    const bool merge = ctx->user_dedup && ctx->n_kept > 0 && F <= 64 && n_users >= 1024 && n_users <= 0x7fffffff;
    if (merge) {
        const int64_t n = n_users;
        CK(ctx->w_masks.reserve(8 * (size_t)n, ctx->stream));
        CK(ctx->w_sorted.reserve(8 * (size_t)n, ctx->stream));
        CK(ctx->w_umask.reserve(8 * (size_t)n, ctx->stream));
        CK(ctx->ud_idx.reserve(4 * (size_t)n, ctx->stream));
        CK(ctx->ud_user.reserve(4 * (size_t)n, ctx->stream));
        CK(ctx->ud_flag.reserve(4 * (size_t)n, ctx->stream));
        CK(ctx->ud_run.reserve(4 * (size_t)n, ctx->stream));
        CK(ctx->ud_len.reserve(8 * (size_t)(n + 1), ctx->stream));
        CK(ctx->ud_off.reserve(8 * (size_t)(n + 1), ctx->stream));
        CK(ctx->ud_ranks.reserve(4 * (size_t)std::max<int64_t>(nnz, 1), ctx->stream));
        CK(ctx->ud_rec.reserve(4 * (size_t)n, ctx->stream));
        auto* masks = ctx->w_masks.as<unsigned long long>();
        auto* sorted = ctx->w_sorted.as<unsigned long long>();
        auto* umask = ctx->w_umask.as<unsigned long long>();
        const unsigned nb = (unsigned)cdiv(n, 256);
        k_row_masks<<<nb, 256, 0, ctx->stream>>>(k9_off, k9_items, nullptr, n, 0, masks);  // (the items are ranks by now)
        LAUNCH_CHECK();
        k_iota<<<nb, 256, 0, ctx->stream>>>(ctx->ud_idx.as<int32_t>(), (int)n);
        LAUNCH_CHECK();
        size_t bytes = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, bytes, masks, sorted, ctx->ud_idx.as<int32_t>(), ctx->ud_user.as<int32_t>(), (int)n, 0, F, ctx->stream));
        CK(ctx->scratch.reserve(bytes, ctx->stream));
        CK(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, bytes, masks, sorted, ctx->ud_idx.as<int32_t>(), ctx->ud_user.as<int32_t>(), (int)n, 0, F, ctx->stream));
        k_run_starts<<<nb, 256, 0, ctx->stream>>>(sorted, n, ctx->ud_flag.as<int32_t>());
        LAUNCH_CHECK();
        CK(cub::DeviceScan::InclusiveSum(nullptr, bytes, ctx->ud_flag.as<int32_t>(), ctx->ud_run.as<int32_t>(), (int)n, ctx->stream));
        CK(ctx->scratch.reserve(bytes, ctx->stream));
        CK(cub::DeviceScan::InclusiveSum(ctx->scratch.p, bytes, ctx->ud_flag.as<int32_t>(), ctx->ud_run.as<int32_t>(), (int)n, ctx->stream));
        uint32_t hd = 0;
        if (int rc = read_scalars(ctx, {{ctx->ud_run.as<int32_t>() + n - 1, 1}}, &hd)) return rc;
        const int64_t D = (int64_t)(int32_t)hd + 1;
        ctx->distinct_baskets = D;
        k_run_heads<<<nb, 256, 0, ctx->stream>>>(sorted, ctx->ud_flag.as<int32_t>(), ctx->ud_run.as<int32_t>(), n, umask, ctx->ud_len.as<int64_t>());
        LAUNCH_CHECK();
        CK(cudaMemsetAsync(ctx->ud_len.as<int64_t>() + D, 0, 8, ctx->stream));
        if (int rc = scan_exclusive(ctx, ctx->ud_len.as<int64_t>(), ctx->ud_off.as<int64_t>(), D + 1)) return rc;
        k_mask_expand<<<(unsigned)cdiv(D, 256), 256, 0, ctx->stream>>>(umask, ctx->ud_off.as<int64_t>(), (int)D, ctx->ud_ranks.as<int32_t>());
        LAUNCH_CHECK();
        k9_off = ctx->ud_off.as<int64_t>();
        k9_items = ctx->ud_ranks.as<int32_t>();
        k9_n = D;
        k9_out = ctx->ud_rec.as<int32_t>();
    }
    if (ctx->n_kept == 0) {
        CK(cudaMemsetAsync(ctx->u_out.p, 0xff, 4 * (size_t)n_users, ctx->stream));
    } else {
        unsigned long long* dbg = nullptr;
        if (getenv("FA_DEBUG_MATCH")) { dbg = (unsigned long long*)ctx->scalars.p; CK(cudaMemsetAsync(dbg, 0, 32, ctx->stream)); }
        CK(cudaFuncSetAttribute(k_recommend, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem, 48 * 1024)));
        int per_sm = 1;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_recommend, 32 * MATCH_WARPS, smem));
        const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>(cdiv(k9_n, MATCH_WARPS), (int64_t)ctx->sm_count * std::max(per_sm, 1)));
        const int gcap = ctx->match_gcap;  // spill ranges per warp in the first pass; baskets that need more are matched again below
        CK(ctx->m_stack.reserve(sizeof(int3) * (size_t)gcap * (size_t)blocks * MATCH_WARPS, ctx->stream));
        CK(cudaMemsetAsync(ctx->errflag.p, 0, 4, ctx->stream));
        k_recommend<<<(unsigned)blocks, 32 * MATCH_WARPS, smem, ctx->stream>>>(ctx->tab, ctx->m_aoff.as<int32_t>(), ctx->m_rec.as<int2>(), F,
                                                                            k9_off, k9_items, k9_n, k9_out, ctx->m_stack.as<int3>(), gcap,
                                                                            ctx->errflag.as<int32_t>(), dbg, nullptr, ctx->match_scap);
        LAUNCH_CHECK();
        if (merge) {
            k_scatter_runs<<<(unsigned)cdiv(n_users, 256), 256, 0, ctx->stream>>>(ctx->ud_user.as<int32_t>(), ctx->ud_run.as<int32_t>(), k9_out,
                                                                                 n_users, ctx->u_out.as<int32_t>());
            LAUNCH_CHECK();
        }
    }
    ctx->tend();
    int32_t overflow = 0;
    CK(cudaMemcpyAsync(consequent_rank_out, ctx->u_out.p, 4 * (size_t)n_users, cudaMemcpyDeviceToHost, ctx->stream));
    if (ctx->n_kept) CK(cudaMemcpyAsync(&overflow, ctx->errflag.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (overflow) {
        // Rare path: some basket had more pending trie ranges than the first pass keeps per warp.  Those baskets (marked -2)
        // are matched again by a small grid whose per-warp stack holds the worst case: a range is pushed per visited node,
        // and at most (basket size <= F) ranges are pending per level of the depth-first walk.
        const int64_t cap2 = (int64_t)F * (ctx->tab.max_k + 1) + 64;
        CK(ctx->ud_flag.reserve((size_t)k9_n, ctx->stream));
        CK(ctx->ud_idx.reserve(4 * (size_t)k9_n, ctx->stream));
        uint8_t* oflag = ctx->ud_flag.as<uint8_t>();
        int32_t* olist = ctx->ud_idx.as<int32_t>();
        int32_t* d_nsel = ctx->scalars.as<int32_t>();
        k_flag_eq<<<(unsigned)cdiv(k9_n, 256), 256, 0, ctx->stream>>>(k9_out, k9_n, -2, oflag);
        LAUNCH_CHECK();
        cub::CountingInputIterator<int32_t> ids(0);
        size_t bytes = 0;
        CK(cub::DeviceSelect::Flagged(nullptr, bytes, ids, oflag, olist, d_nsel, (int)k9_n, ctx->stream));
        CK(ctx->scratch.reserve(bytes, ctx->stream));
        CK(cub::DeviceSelect::Flagged(ctx->scratch.p, bytes, ids, oflag, olist, d_nsel, (int)k9_n, ctx->stream));
        uint32_t hn = 0;
        if (int rc = read_scalars(ctx, {{d_nsel, 1}}, &hn)) return rc;
        const int64_t n2 = (int32_t)hn;
        if (n2 > 0) {
            const int64_t blocks2 = std::max<int64_t>(1, std::min<int64_t>(cdiv(n2, MATCH_WARPS), 32));
            CK(ctx->m_stack.reserve(sizeof(int3) * (size_t)cap2 * (size_t)blocks2 * MATCH_WARPS, ctx->stream));
            CK(cudaMemsetAsync(ctx->errflag.p, 0, 4, ctx->stream));
            k_recommend<<<(unsigned)blocks2, 32 * MATCH_WARPS, smem, ctx->stream>>>(ctx->tab, ctx->m_aoff.as<int32_t>(), ctx->m_rec.as<int2>(), F,
                                                                                 k9_off, k9_items, n2, k9_out, ctx->m_stack.as<int3>(), (int)cap2,
                                                                                 ctx->errflag.as<int32_t>(), nullptr, olist, MATCH_STACK_SMEM);
            LAUNCH_CHECK();
            if (merge) {
                k_scatter_runs<<<(unsigned)cdiv(n_users, 256), 256, 0, ctx->stream>>>(ctx->ud_user.as<int32_t>(), ctx->ud_run.as<int32_t>(), k9_out,
                                                                                     n_users, ctx->u_out.as<int32_t>());
                LAUNCH_CHECK();
            }
            overflow = 0;
            CK(cudaMemcpyAsync(consequent_rank_out, ctx->u_out.p, 4 * (size_t)n_users, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaMemcpyAsync(&overflow, ctx->errflag.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            ctx->match_retries += n2;
        }
        if (overflow) return ctx->fail(FA_ELIMIT, "fa_recommend: trie-walk stack overflow even with the worst-case stack (corrupt lattice?)");
    }
    if (getenv("FA_DEBUG_MATCH") && ctx->n_kept) {
        unsigned long long h[4];
        cudaMemcpy(h, ctx->scalars.p, 32, cudaMemcpyDeviceToHost);
        fprintf(stderr, "[match] users=%lld rules=%lld nodes_total=%llu nodes_max=%llu\n", (long long)n_users, (long long)ctx->n_kept, h[0], h[1]);
    }
    return FA_OK;
}

int fa_get_stats(fa_ctx* ctx, fa_stats* out) {
    if (!ctx || !out) return FA_EINVAL;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->resolve_timers();
    out->ms_hist = ctx->ms[T_HIST];
    out->ms_bitmap = ctx->ms[T_BITMAP];
    out->ms_pairs = ctx->ms[T_PAIRS];
    out->ms_candgen = ctx->ms[T_CANDGEN];
    out->ms_count = ctx->ms[T_COUNT];
    out->ms_threshold = ctx->ms[T_THRESH];
    out->ms_rules = ctx->ms[T_RULES];
    out->ms_sort = ctx->ms[T_SORT];
    out->ms_match = ctx->ms[T_MATCH];
    out->ms_exchange = ctx->ms[T_EXCHANGE];
    out->count_bytes = ctx->count_bytes;
    out->count_load_bytes = ctx->count_load_bytes;
    out->pair_ops = ctx->pair_ops;
    out->count_launches = ctx->count_launches;
    out->kernel_launches = ctx->kernel_launches;
    out->bitmap_bytes = ctx->have_bits ? (int64_t)ctx->pitch_words * 4 * ctx->F : 0;
    out->row_trims = ctx->trims;
    out->live_rows = ctx->live_rows;
    out->total_count = ctx->total_count;
    out->weighted = ctx->weighted ? 1 : 0;
    out->distinct_baskets = ctx->distinct_baskets;
    out->match_retries = ctx->match_retries;
    return FA_OK;
}

int fa_probe_l1_bandwidth(fa_ctx* ctx, double* gb_per_s_out) {
    if (!ctx || !gb_per_s_out) return FA_EINVAL;
    CK(cudaSetDevice(ctx->device));
    // 2 resident CTAs of 8 warps per SM like K4, 2 KB per warp (32 KB per SM: well inside L1), warm-up pass first
    const int blocks = ctx->sm_count * 2, iters = 4096;
    CK(ctx->scratch.reserve((size_t)blocks * 8 * 2048 + 64, ctx->stream));
    CK(cudaMemsetAsync(ctx->scratch.p, 0x5a, (size_t)blocks * 8 * 2048 + 64, ctx->stream));
    uint32_t* sink = reinterpret_cast<uint32_t*>(ctx->scratch.as<uint8_t>() + (size_t)blocks * 8 * 2048);
    cudaEvent_t a = ctx->get_event(), b = ctx->get_event();
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(a, ctx->stream));
        k_probe_l1<<<blocks, 256, 0, ctx->stream>>>(ctx->scratch.as<uint4>(), iters, 0u, sink);
        LAUNCH_CHECK();
        CK(cudaEventRecord(b, ctx->stream));
        CK(cudaEventSynchronize(b));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (rep > 0 && ms > 0) best = std::max(best, (double)blocks * 8 * 2048.0 * iters / (ms * 1e-3) / 1e9);
    }
    ctx->ev_pool.push_back(a);
    ctx->ev_pool.push_back(b);
    *gb_per_s_out = best;
    return FA_OK;
}

int fa_pair_counts_device(fa_ctx* ctx, const int32_t** dev_matrix_out, int32_t* n_frequent_out) {
    if (!ctx || !dev_matrix_out || !n_frequent_out) return FA_EINVAL;
    if (ctx->levels.size() < 3 || !ctx->pair_counts.p) return ctx->fail(FA_EINVAL, "fa_pair_counts_device before fa_count_pairs");
    *dev_matrix_out = ctx->pair_counts.as<int32_t>();
    *n_frequent_out = ctx->F;
    return FA_OK;
}

int fa_reset_stats(fa_ctx* ctx) {
    if (!ctx) return FA_EINVAL;
    cudaStreamSynchronize(ctx->stream);
    ctx->resolve_timers();
    for (double& m : ctx->ms) m = 0;
    ctx->count_bytes = ctx->count_load_bytes = ctx->pair_ops = 0;
    ctx->count_launches = ctx->kernel_launches = 0;
    ctx->trims = 0;
    ctx->match_retries = 0;
    return FA_OK;
}

}  // extern "C"

// ---- allocation-free K4 (include/fastapriori_b200.h: fa_count_groups_noalloc) ---------------------
// Some hosts may not allocate device memory at all (see the header).  For them the scratch of K4 —
// the alive-row bitmap, plus one 64-bit cell for its population count — sits in this static device
// array.  In host code `g_noalloc_scratch` is only a handle for the runtime, never an address a kernel
// can use: the real device address is looked up once per device with cudaGetSymbolAddress, with that
// device current, and that pointer is what the kernels receive.
#include <mutex>

#define FA_NOALLOC_TAIL 64  // words after the bitmap; the first two hold the 64-bit alive-row total
static __device__ __align__(256) uint32_t g_noalloc_scratch[FA_NOALLOC_MAX_WORDS + FA_NOALLOC_TAIL];

static __global__ void k_noalloc_popc_total(const uint32_t* __restrict__ words, int64_t n, unsigned long long* __restrict__ total) {
    unsigned long long s = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += __popc(words[i]);
    for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(total, s);
}

namespace {
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_dev[64] = {nullptr};  // per device: address of g_noalloc_scratch on that device

int noalloc_fail(int code, char* err_out, int64_t err_cap, const char* fmt, long long a = 0, long long b = 0) {
    if (err_out && err_cap > 0) snprintf(err_out, (size_t)err_cap, fmt, a, b);
    return code;
}
}  // namespace

extern "C" int fa_count_groups_noalloc(int device, void* cuda_stream, const uint32_t* dev_bits, int64_t pitch_words,
                                       int32_t n_frequent, const int32_t* dev_prefix_sets, int32_t prefix_len,
                                       const int32_t* dev_group_rows, const int32_t* dev_group_offsets,
                                       const int32_t* dev_extensions, int64_t n_groups, int64_t n_candidates,
                                       int32_t* dev_counts_out, int64_t* alive_rows_out, char* err_out, int64_t err_cap) {
    if (err_out && err_cap > 0) err_out[0] = 0;
    if (device < 0 || device >= 64) return noalloc_fail(FA_EINVAL, err_out, err_cap, "device %lld out of range [0, 64)", device);
    if (pitch_words <= 0 || (pitch_words & 3)) return noalloc_fail(FA_EINVAL, err_out, err_cap, "pitch_words %lld must be a positive multiple of 4", pitch_words);
    if (pitch_words > FA_NOALLOC_MAX_WORDS)
        return noalloc_fail(FA_ELIMIT, err_out, err_cap, "pitch_words %lld exceeds the static scratch of %lld words (33554432 rows)", pitch_words, (long long)FA_NOALLOC_MAX_WORDS);
    if (n_frequent <= 0 || prefix_len < 1 || n_groups < 0 || n_candidates < 0 || n_candidates > 0x7fffffff)
        return noalloc_fail(FA_EINVAL, err_out, err_cap, "bad sizes (n_groups %lld, n_candidates %lld)", n_groups, n_candidates);
    if (prefix_len > FA_NOALLOC_MAX_PREFIX)  // the kernel stages 64 prefix items per group; beyond that it would AND the wrong rows
        return noalloc_fail(FA_ELIMIT, err_out, err_cap, "prefix_len %lld exceeds the limit of %lld items per prefix", prefix_len, (long long)FA_NOALLOC_MAX_PREFIX);
    if (!dev_bits || ((uintptr_t)dev_bits & 15) || (n_groups && (!dev_prefix_sets || !dev_group_rows || !dev_group_offsets)) ||
        (n_candidates && (!dev_extensions || !dev_counts_out)))
        return noalloc_fail(FA_EINVAL, err_out, err_cap, "null or misaligned device pointer");
    // the caller's current device is put back on every way out (a plugin must not change its host's device)
    struct DeviceGuard {
        int prev = -1;
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e == cudaSuccess && cur != device) {
        e = cudaSetDevice(device);
        if (e == cudaSuccess) guard.prev = cur;
    }
    if (e != cudaSuccess) return noalloc_fail(FA_ECUDA, err_out, err_cap, "selecting the device failed with CUDA error %lld", (long long)e);
    uint32_t* scratch = nullptr;
    {
        // set-up, once per device and with that device current: resolve the array's device address
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        if (!g_noalloc_dev[device]) {
            uint32_t* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) return noalloc_fail(FA_ECUDA, err_out, err_cap, "cudaGetSymbolAddress failed with CUDA error %lld", (long long)e);
            g_noalloc_dev[device] = p;
        }
        scratch = g_noalloc_dev[device];
    }
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    unsigned long long* total = reinterpret_cast<unsigned long long*>(scratch + FA_NOALLOC_MAX_WORDS);
    // the array still holds the previous call's bits: clear what this call reads or ORs into
    if ((e = cudaMemsetAsync(scratch, 0, 4 * (size_t)pitch_words, stream)) != cudaSuccess ||
        (e = cudaMemsetAsync(total, 0, 8, stream)) != cudaSuccess ||
        (n_candidates && (e = cudaMemsetAsync(dev_counts_out, 0, 4 * (size_t)n_candidates, stream)) != cudaSuccess))
        return noalloc_fail(FA_ECUDA, err_out, err_cap, "cudaMemsetAsync failed with CUDA error %lld", (long long)e);
    if (n_groups > 0) {  // also without candidates: the groups' prefix tidsets still define the alive rows
        const int64_t pitch_q = pitch_words / 4;
        const int64_t n_chunks = cdiv(pitch_q, 32 * COUNT_U);
        const int64_t units = cdiv(n_groups, NOALLOC_G) * n_chunks;
        const unsigned blocks = (unsigned)std::min<int64_t>(cdiv(units, 8), 0x7fffffff);
        if (COUNT_META_BYTES(prefix_len) > 48 * 1024 &&
            (e = cudaFuncSetAttribute(k_count_groups<COUNT_U, NOALLOC_G, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, COUNT_META_BYTES(64))) != cudaSuccess)
            return noalloc_fail(FA_ECUDA, err_out, err_cap, "cudaFuncSetAttribute failed with CUDA error %lld", (long long)e);
        k_count_groups<COUNT_U, NOALLOC_G, true><<<blocks, 256, COUNT_META_BYTES(prefix_len), stream>>>(reinterpret_cast<const uint4*>(dev_bits), pitch_q, dev_prefix_sets,
                                                                    prefix_len, dev_group_rows, dev_group_offsets, dev_extensions,
                                                                    n_groups, n_chunks, dev_counts_out, scratch);
        if ((e = cudaGetLastError()) != cudaSuccess) return noalloc_fail(FA_ECUDA, err_out, err_cap, "kernel launch failed with CUDA error %lld", (long long)e);
    }
    if (alive_rows_out) {
        k_noalloc_popc_total<<<(unsigned)std::min<int64_t>(cdiv(pitch_words, 256), 1024), 256, 0, stream>>>(scratch, pitch_words, total);
        unsigned long long h = 0;
        if ((e = cudaGetLastError()) != cudaSuccess || (e = cudaMemcpyAsync(&h, total, 8, cudaMemcpyDeviceToHost, stream)) != cudaSuccess ||
            (e = cudaStreamSynchronize(stream)) != cudaSuccess)
            return noalloc_fail(FA_ECUDA, err_out, err_cap, "reading the alive-row total failed with CUDA error %lld", (long long)e);
        *alive_rows_out = (int64_t)h;
    }
    return FA_OK;
}

#include "fa_text.inl"
