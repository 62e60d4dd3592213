// K3 on the 5th-generation tensor cores (FastApriori.scala:212-241 as a dense contraction).
//
// P = B^T B for the 0/1 transaction x item matrix B, exact in int32: P[i][j] = sum_r B[r][i]*B[r][j].
// Both operands come from the same bit-packed vertical database (items x rows, K = rows contiguous),
// i.e. a "TN" GEMM with K-major A and B.  One CTA owns a 128 x 128 tile of item pairs and a slice of
// the row axis:
//   * 8 producer warps (4 for the A tile, 4 for the B tile) read 128 row-bits per item and k-block (one
//     uint4), expand every bit to a byte and store the 128-byte row straight into the 128B-swizzled K-major shared-memory
//     layout the UMMA descriptors expect.  The bytes are SIGNED int8 0 / -1: (-1)*(-1) = +1, so the products are the same
//     0/1 as for a 0/1 matrix, and PRMT's sign-replicate mode makes 4 such bytes from 4 bits in one instruction (the bit
//     order inside a 32-row word is permuted identically for A and B, which a sum over rows does not see) (the expansion cannot be done by TMA, and a pre-expanded
//     int8 matrix would be 8x the traffic: 164 GB at BASELINE config 3);
//   * one thread issues tcgen05.mma.cta_group::1.kind::i8 (M=128, N=128, K=32 per instruction, four
//     per k-block) with the int32 accumulator in TMEM; tcgen05.commit releases the stage;
//   * after the last k-block the producer warps read the accumulator with tcgen05.ld (32 lanes per
//     warp) and add the upper-triangle entries into the pair matrix.
// The A operand does not go through shared memory: its producer warps write the expanded rows straight into TMEM
// (tcgen05.st, lane = row, 32 columns = 128 row-bytes per k-block) and the MMA takes A from TMEM.  With both operands
// in shared memory a k-block costs 32 KB of stores plus 32 KB of tensor-core reads per 256 MMA cycles, twice what a
// 128 B/clk shared memory delivers (that version measured 33 % tensor-pipe utilisation whatever the expansion cost).
// mbarrier full/empty pipeline, 4 stages (16 KB of B in shared memory + 32 TMEM columns of A each; 256 TMEM columns
// per CTA, two CTAs per SM, so one CTA's expansion overlaps the other's MMAs).  Every wait is bounded: a protocol
// error traps instead of hanging the GPU.
// Tried and dropped (all bit-exact, all slower on B200): a 256-entry shared-memory bit->byte table (1.6x slower: divergent
// LDS.64 lookups compete with the operand stores and the tensor core's own shared-memory reads); zeroed stages with one
// byte store per set bit for sparse rows (2x slower: the per-bit loops diverge across the warp); a multiply/mask
// nibble->4-byte expansion to 0/1 bytes (the first version, ~1.5x the instructions of the PRMT form).
#pragma once

#include "fa_common.cuh"

#define TC_TILE 128
#define TC_KB 128           // rows (K) per k-block = bytes per smem row = one 128B swizzle atom
#define TC_STAGES 4
#define TC_PRODUCERS 256
#define TC_THREADS 288      // 8 producer/epilogue warps + 1 MMA warp
#define TC_TILE_BYTES (TC_TILE * TC_KB)                  // 16 KB per operand per stage
#define TC_SMEM_BYTES (TC_STAGES * TC_TILE_BYTES + 1024 /*alignment*/ + 256 /*barriers*/)
#define TC_TMEM_COLS 256    // 128 accumulator columns + TC_STAGES * 32 columns of A

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin)
        if (spin > (1u << 28)) __trap();  // bounded: a broken pipeline must not hang the device
}

// K-major, 128B-swizzled operand tile at a 1024-byte aligned shared address (cute::UMMA::SmemDescriptor):
// start>>4 [0,14) | LBO=1 [16,30) | SBO=1024>>4 [32,46) | version=1 [46,48) | layout SWIZZLE_128B=2 [61,64)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fff) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::i8 instruction descriptor (cute::UMMA::InstrDescriptor): D=S32, A=B=signed 8-bit (format 1), both K-major
__device__ __forceinline__ constexpr uint32_t umma_idesc_i8(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | (0u << 15) | (0u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// same instruction with the A operand in TMEM (lane = M row, 4 int8 per 32-bit column)
__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// bytes 0x00 / 0xFF from the most significant bit of each byte of x (prmt, sign-replicate selectors 8..11)
__device__ __forceinline__ uint32_t msb_bytes(uint32_t x) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %1, 0xba98;" : "=r"(d) : "r"(x));
    return d;
}

// expand one item's 128 row bits into 32 words = 128 bytes of 0x00 / 0xFF: word 8*wi + s holds bits 7-s, 15-s, 23-s,
// 31-s of bits word wi (x << s brings them to the byte MSBs).  The same order is used for A and B.
__device__ __forceinline__ void expand_words(uint4 w, uint32_t (&v)[32]) {
    const uint32_t words[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int wi = 0; wi < 4; ++wi)
#pragma unroll
        for (int sft = 0; sft < 8; ++sft) v[8 * wi + sft] = msb_bytes(words[wi] << sft);
}
// B: the 128-byte row goes to shared memory, 8 chunks of 16 bytes, chunk index XOR row%8 (128B swizzle)
__device__ __forceinline__ void store_row_smem(uint8_t* tile, int r, const uint32_t (&v)[32]) {
    uint8_t* row = tile + r * TC_KB;
    const int sw = r & 7;
#pragma unroll
    for (int j = 0; j < 8; ++j)
        *reinterpret_cast<uint4*>(row + ((j ^ sw) << 4)) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
// A: the row goes to this thread's TMEM lane, 32 consecutive columns
__device__ __forceinline__ void store_row_tmem(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
          "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
          "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
          "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// the same store without the wait: the caller overlaps the TMEM write with the next row's expansion, then tmem_store_wait()
__device__ __forceinline__ void store_row_tmem_nowait(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
          "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
          "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
          "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void tmem_store_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

}  // namespace tc

__global__ void __launch_bounds__(TC_THREADS, 2) k_pair_tc(const uint32_t* __restrict__ bits, int64_t pitch_words, int F,
                                                           int64_t kblocks_per_split, const int2* __restrict__ tiles,
                                                           int32_t* __restrict__ P) {
    extern __shared__ uint8_t tc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_b = smem;                                   // [stage][128 rows][128 B]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + TC_STAGES * TC_TILE_BYTES);
    uint64_t* empty = full + TC_STAGES;
    uint64_t* acc_ready = empty + TC_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int2 tl = tiles[blockIdx.x];
    const int i0 = tl.x * TC_TILE, j0 = tl.y * TC_TILE;
    const int64_t n_kb = pitch_words / 4;  // 4 words = 128 rows per k-block (pitch is a multiple of 512 words)
    const int64_t kb0 = (int64_t)blockIdx.y * kblocks_per_split;
    const int64_t kb1 = min(n_kb, kb0 + kblocks_per_split);
    const int64_t nkb = kb1 - kb0;

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { tc::mbar_init(full + s, TC_PRODUCERS); tc::mbar_init(empty + s, 1); }
        tc::mbar_init(acc_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) {  // one warp allocates TMEM: 128 accumulator columns + the A stages
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(tmem_slot)), "r"((uint32_t)TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_acc = *tmem_slot;

    if (warp < 8) {
        // ---- producers: bits -> 0/1 bytes in swizzled smem; warps 0-3 fill the A tile, warps 4-7 the B tile ----
        const bool is_b = warp >= 4;
        const int r = tid & 127;  // row of the tile this thread expands
        const int item = (is_b ? j0 : i0) + r;
        const bool ok = item < F;
        const uint4* src = reinterpret_cast<const uint4*>(bits + (int64_t)(ok ? item : 0) * pitch_words) + kb0;
        const uint32_t a_lane_addr = tmem_acc + ((uint32_t)((warp & 3) * 32) << 16) + 128u;  // this warp's lanes, first A column
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        uint4 w = zero;
        if (nkb > 0 && ok) w = __ldg(src);
        for (int64_t kb = 0; kb < nkb; ++kb) {
            const int s = (int)(kb % TC_STAGES);
            const uint32_t ph = (uint32_t)((kb / TC_STAGES) & 1);
            uint4 nw = zero;  // next k-block's bits are in flight while this one is expanded
            if (kb + 1 < nkb && ok) nw = __ldg(src + kb + 1);
            tc::mbar_wait(empty + s, ph ^ 1);
            uint32_t v[32];
            tc::expand_words(w, v);
            if (is_b) {
                tc::store_row_smem(smem_b + s * TC_TILE_BYTES, r, v);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the MMA's async proxy
            } else {
                tc::store_row_tmem(a_lane_addr + (uint32_t)(s * 32), v);
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            }
            tc::mbar_arrive(full + s);
            w = nw;
        }
        // ---- epilogue: TMEM -> registers -> pair matrix (upper triangle); a warp may only touch TMEM lanes
        // 32*(warp%4)..+31, so warps 0-3 take accumulator columns 0-63 and warps 4-7 columns 64-127 ----
        if (nkb > 0) {
            tc::mbar_wait(acc_ready, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int lane_base = (warp & 3) * 32;
            const int i = i0 + lane_base + lane;
#pragma unroll 1
            for (int c = (is_b ? 2 : 0); c < (is_b ? 4 : 2); ++c) {
                uint32_t v[32];
                const uint32_t taddr = tmem_acc + ((uint32_t)lane_base << 16) + (uint32_t)(c * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    const int j = j0 + c * 32 + t;
                    if (i < j && j < F && v[t]) atomicAdd(P + (int64_t)i * F + j, (int32_t)v[t]);
                }
            }
        }
    } else if (lane == 0) {
        // ---- MMA issuer (one thread) ----
        const uint32_t idesc = tc::umma_idesc_i8(TC_TILE, TC_TILE);
        for (int64_t kb = 0; kb < nkb; ++kb) {
            const int s = (int)(kb % TC_STAGES);
            const uint32_t ph = (uint32_t)((kb / TC_STAGES) & 1);
            tc::mbar_wait(full + s, ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t ta = tmem_acc + 128u + (uint32_t)(s * 32);
            const uint64_t db = tc::umma_desc(tc::smem_u32(smem_b + s * TC_TILE_BYTES));
#pragma unroll
            for (int k = 0; k < TC_KB / 32; ++k)  // K = 32 bytes per instruction: 8 TMEM columns of A, 2 descriptor units of B
                tc::umma_i8_ts(tmem_acc, ta + 8 * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            tc::umma_commit(empty + s);           // frees the stage once these MMAs have read it
        }
        if (nkb > 0) tc::umma_commit(acc_ready);  // accumulator complete
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "r"((uint32_t)TC_TMEM_COLS) : "memory");
}

// ------------------------------------------------------------------------------------------
// K3 with a 256 x 128 item tile per CTA.  Same operand paths as k_pair_tc (A rows expanded straight into TMEM, B rows into
// 128B-swizzled shared memory, one thread issuing kind::i8 MMAs, int32 accumulators in TMEM), but the B tile a CTA
// expands and stores is used by TWO M=128 accumulators: per int8 MAC a quarter fewer rows are expanded and a third
// fewer bytes cross shared memory — the two resources the 128 x 128 kernel runs out of (ncu r1f: shared-memory data pipe
// 91 %, tensor pipe 52 %).  13 warps: 0-7 expand the 256 A rows (warp w: accumulator half w/4, TMEM lanes 32*(w%4)..),
// 8-11 the 128 B rows, warp 12 allocates TMEM and issues.  All 512 TMEM columns (2 x 128 accumulator + 4 stages x 2 x 32
// of A), hence one CTA per SM; the 4-stage ring keeps the producers ahead of the MMAs.
// Where it stands (ncu r2d/r2h): the producers spend ~40 % of their samples waiting for a free stage and the MMA warp never
// waits for a full one, i.e. the MMA issue rate of cta_group::1 is the limiter now (8 M128 N128 K32 instructions per
// k-block in ~930 cycles).  A second B producer group (even / odd k-blocks, 17 warps) was measured: 114 ms against 108 ms
// at 1 M x 16384 — more warps only add shared-memory pressure.  The next step is cta_group::2.
// Tiles: A block a covers items [256a, 256a+256), B block b items [128b, 128b+128), for every b >= 2a.
// ------------------------------------------------------------------------------------------
#define TC2_M 256
#define TC2_PRODUCERS 384
#define TC2_THREADS 416
#define TC2_TMEM_COLS 512

__global__ void __launch_bounds__(TC2_THREADS, 1) k_pair_tc2(const uint32_t* __restrict__ bits, int64_t pitch_words, int F,
                                                             int64_t kblocks_per_split, const int2* __restrict__ tiles,
                                                             int32_t* __restrict__ P) {
    extern __shared__ uint8_t tc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_b = smem;                                   // [stage][128 rows][128 B]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + TC_STAGES * TC_TILE_BYTES);
    uint64_t* empty = full + TC_STAGES;
    uint64_t* acc_ready = empty + TC_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int2 tl = tiles[blockIdx.x];
    const int i0 = tl.x * TC2_M, j0 = tl.y * TC_TILE;
    const int64_t n_kb = pitch_words / 4;
    const int64_t kb0 = (int64_t)blockIdx.y * kblocks_per_split;
    const int64_t kb1 = min(n_kb, kb0 + kblocks_per_split);
    const int64_t nkb = kb1 - kb0;

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { tc::mbar_init(full + s, TC2_PRODUCERS / 32); tc::mbar_init(empty + s, 1); }  // full: one arrival per producer warp
        tc::mbar_init(acc_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 12) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(tmem_slot)), "r"((uint32_t)TC2_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;   // columns: [0,128) accumulator 0, [128,256) accumulator 1, [256,512) A stages

    if (warp < 12) {
        const bool is_b = warp >= 8;
        const int half = warp >> 2;                           // A: which accumulator's rows (0/1); B: 2
        const int r = is_b ? (tid - 256) : (tid & 127);       // row inside the 128-row block this thread expands
        const int item = is_b ? j0 + r : i0 + half * 128 + r;
        const bool ok = item < F;
        const uint4* src = reinterpret_cast<const uint4*>(bits + (int64_t)(ok ? item : 0) * pitch_words) + kb0;
        // A rows of accumulator h, stage s: columns 256 + s*64 + h*32 .. +31, this warp's lane quarter
        const uint32_t a_lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 256u + (uint32_t)(half * 32);
        // Four k-blocks (64 bytes of the row: two whole sectors) are fetched per step and one step ahead: ncu showed the
        // producers waiting on their 16-byte row loads (long-scoreboard 3.9 per issue with one k-block of look-ahead),
        // with shared memory (65 %) and the tensor pipe (56 %) both unsaturated.  TC_STAGES == 4, so k-block kb + q uses
        // stage q: the ring indices are compile-time constants.
        static_assert(TC_STAGES == 4, "the producer loop maps k-block kb + q to stage q");
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        uint4 cur[4], nxt[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) cur[q] = (ok && q < nkb) ? __ldg(src + q) : zero;
        for (int64_t kb = 0; kb < nkb; kb += 4) {
            const uint32_t ph = (uint32_t)((kb >> 2) & 1);
#pragma unroll
            for (int q = 0; q < 4; ++q) nxt[q] = (ok && kb + 4 + q < nkb) ? __ldg(src + kb + 4 + q) : zero;
            if (is_b) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (kb + q < nkb) {
                        uint32_t v[32];
                        tc::expand_words(cur[q], v);
                        tc::mbar_wait(empty + q, ph ^ 1);
                        tc::store_row_smem(smem_b + q * TC_TILE_BYTES, r, v);
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) tc::mbar_arrive(full + q);
                    }
                }
            } else {
                // A rows: the TMEM store of k-block q is left in flight while row q+1 is expanded; its completion is awaited
                // (and the stage handed to the MMA warp) just before the next store is issued
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (kb + q < nkb) {
                        uint32_t v[32];
                        tc::expand_words(cur[q], v);
                        if (q > 0) {
                            tc::tmem_store_wait();
                            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                            __syncwarp();
                            if (lane == 0) tc::mbar_arrive(full + q - 1);
                        }
                        tc::mbar_wait(empty + q, ph ^ 1);
                        tc::store_row_tmem_nowait(a_lane_addr + (uint32_t)(q * 64), v);
                    }
                }
                {   // the group's last store
                    const int last = (int)min((int64_t)3, nkb - kb - 1);
                    tc::tmem_store_wait();
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(full + last);
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) cur[q] = nxt[q];
        }
        // epilogue: warps 0-7 read their accumulator (half) and lane quarter, 128 columns in four 32-column loads
        if (nkb > 0 && !is_b) {
            tc::mbar_wait(acc_ready, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int lane_base = (warp & 3) * 32;
            const int i = i0 + half * 128 + lane_base + lane;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(half * 128 + c * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    const int j = j0 + c * 32 + t;
                    if (i < j && j < F && v[t]) atomicAdd(P + (int64_t)i * F + j, (int32_t)v[t]);
                }
            }
        }
    } else if (lane == 0) {
        const uint32_t idesc = tc::umma_idesc_i8(TC_TILE, TC_TILE);
        for (int64_t kb = 0; kb < nkb; ++kb) {
            const int s = (int)(kb % TC_STAGES);
            const uint32_t ph = (uint32_t)((kb / TC_STAGES) & 1);
            tc::mbar_wait(full + s, ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t db = tc::umma_desc(tc::smem_u32(smem_b + s * TC_TILE_BYTES));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t ta = tmem_base + 256u + (uint32_t)(s * 64 + h * 32);
#pragma unroll
                for (int k = 0; k < TC_KB / 32; ++k)
                    tc::umma_i8_ts(tmem_base + (uint32_t)(h * 128), ta + 8 * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            }
            tc::umma_commit(empty + s);
        }
        if (nkb > 0) tc::umma_commit(acc_ready);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 12) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TC2_TMEM_COLS) : "memory");
}

// ------------------------------------------------------------------------------------------
// K3 on a CTA PAIR (tcgen05 cta_group::2): the two CTAs of a 2-cluster sit on the two SMs of a TPC and issue ONE MMA stream
// with M = 256 (128 rows of A from each CTA's own TMEM) and N = 128 (64 rows of B from each CTA's own shared memory).
// Per CTA that keeps the 256 x 128 kernel's two accumulators (512 A items per pair, 128 B items) while each SM expands
// and stores only half of the B tile, and the MMAs run in the 2-SM mode the tensor core reaches its full rate in.
//   threads per CTA: warps 0-7 expand 256 A rows into TMEM, warps 8-9 expand 64 B rows into shared memory, warp 10
//   allocates TMEM (both CTAs, cta_group::2) and — in the leader CTA (cluster rank 0) only — issues the MMAs.
//   barriers: full[s] lives in the leader (20 arrivals: every producer warp of both CTAs, the peer's through
//   mapa + mbarrier.arrive.shared::cluster); empty[s] and acc_ready exist in both CTAs and are signalled together by
//   tcgen05.commit ... multicast::cluster with mask 0b11.
// Tiles: pair (a, b) covers A items [512a, 512a + 512) (CTA rank r: [512a + 256r, +256)) and B items [128b, 128b + 128)
// (rank r expands rows [128b + 64r, +64)), for every b >= 4a.  All of it reads this GPU's memory only.
// ------------------------------------------------------------------------------------------
#define TC3_THREADS 352
#define TC3_B_ROWS 64
#define TC3_B_BYTES (TC3_B_ROWS * TC_KB)   // 8 KB per stage and CTA
#define TC3_SMEM_BYTES (TC_STAGES * TC3_B_BYTES + 1024 + 256)
#define TC3_FULL_ARRIVALS (2 * (8 + TC3_B_ROWS / 32))   // producer WARPS of both CTAs

namespace tc {
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in the CTA of cluster rank `rank`
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(bar)), "r"(rank));
    // (default semantics, as CUTLASS's ClusterBarrier::arrive(cta_id) uses: the .release.cluster form costs a GPU-scope
    // MEMBAR + ERRBAR per arrival — 4.1 membar stalls per issue in ncu r2l; what the arrival publishes is already complete:
    // shared-memory stores fenced to the async proxy, TMEM stores awaited with tcgen05.wait::st)
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void umma_i8_ts_2sm(uint32_t tmem_d, uint32_t tmem_a, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], [%1], %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {  // signals the barrier at this offset in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
}  // namespace tc

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC3_THREADS, 1)
k_pair_tc3(const uint32_t* __restrict__ bits, int64_t pitch_words, int F, int64_t kblocks_per_split, const int2* __restrict__ tiles,
           int32_t* __restrict__ P) {
    extern __shared__ uint8_t tc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_b = smem;                                   // [stage][64 rows][128 B]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + TC_STAGES * TC3_B_BYTES);
    uint64_t* empty = full + TC_STAGES;
    uint64_t* acc_ready = empty + TC_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int2 tl = tiles[blockIdx.x >> 1];
    const int i0 = tl.x * 512 + (int)rank * 256, j0 = tl.y * TC_TILE;
    const int64_t n_kb = pitch_words / 4;
    const int64_t kb0 = (int64_t)blockIdx.y * kblocks_per_split;
    const int64_t kb1 = min(n_kb, kb0 + kblocks_per_split);
    const int64_t nkb = kb1 - kb0;

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { tc::mbar_init(full + s, TC3_FULL_ARRIVALS); tc::mbar_init(empty + s, 1); }  // full: one arrival per producer warp of either CTA
        tc::mbar_init(acc_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 10) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(tmem_slot)), "r"((uint32_t)TC2_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    tc::cluster_sync_all();   // both CTAs' barriers and TMEM exist before anyone signals or issues
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;   // columns: [0,128) accumulator 0, [128,256) accumulator 1, [256,512) A stages

    if (warp < 10) {
        const bool is_b = warp >= 8;
        const int half = warp >> 2;                           // A: which accumulator's rows (0/1)
        const int r = is_b ? (tid - 256) : (tid & 127);       // row inside the block this thread expands (B: 0..63)
        const int item = is_b ? j0 + (int)rank * TC3_B_ROWS + r : i0 + half * 128 + r;
        const bool ok = item < F;
        const uint4* src = reinterpret_cast<const uint4*>(bits + (int64_t)(ok ? item : 0) * pitch_words) + kb0;
        const uint32_t a_lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 256u + (uint32_t)(half * 32);
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        uint4 cur[4], nxt[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) cur[q] = (ok && q < nkb) ? __ldg(src + q) : zero;
        for (int64_t kb = 0; kb < nkb; kb += 4) {
            const uint32_t ph = (uint32_t)((kb >> 2) & 1);
#pragma unroll
            for (int q = 0; q < 4; ++q) nxt[q] = (ok && kb + 4 + q < nkb) ? __ldg(src + kb + 4 + q) : zero;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (kb + q < nkb) {
                    uint32_t v[32];
                    tc::expand_words(cur[q], v);
                    tc::mbar_wait(empty + q, ph ^ 1);
                    if (is_b) {
                        tc::store_row_smem(smem_b + q * TC3_B_BYTES, r, v);
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    } else {
                        tc::store_row_tmem(a_lane_addr + (uint32_t)(q * 64), v);
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    }
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive_cluster(full + q, 0);     // the leader's barrier, from either CTA
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) cur[q] = nxt[q];
        }
        if (nkb > 0 && !is_b) {
            tc::mbar_wait(acc_ready, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int lane_base = (warp & 3) * 32;
            const int i = i0 + half * 128 + lane_base + lane;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(half * 128 + c * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    const int j = j0 + c * 32 + t;
                    if (i < j && j < F && v[t]) atomicAdd(P + (int64_t)i * F + j, (int32_t)v[t]);
                }
            }
        }
    } else if (rank == 0 && lane == 0) {
        const uint32_t idesc = tc::umma_idesc_i8(256, TC_TILE);
        for (int64_t kb = 0; kb < nkb; ++kb) {
            const int s = (int)(kb % TC_STAGES);
            const uint32_t ph = (uint32_t)((kb / TC_STAGES) & 1);
            tc::mbar_wait(full + s, ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t db = tc::umma_desc(tc::smem_u32(smem_b + s * TC3_B_BYTES));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint32_t ta = tmem_base + 256u + (uint32_t)(s * 64 + h * 32);
#pragma unroll
                for (int k = 0; k < TC_KB / 32; ++k)
                    tc::umma_i8_ts_2sm(tmem_base + (uint32_t)(h * 128), ta + 8 * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            }
            tc::umma_commit_2sm(empty + s);
        }
        if (nkb > 0) tc::umma_commit_2sm(acc_ready);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    tc::cluster_sync_all();   // the leader's MMAs read the peer's TMEM and shared memory: nobody leaves early
    if (warp == 10) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TC2_TMEM_COLS) : "memory");
}

// ------------------------------------------------------------------------------------------
// The CTA-pair kernel with ONE accumulator of N = 256: pair (a, b) covers A items [256a, +256) (128 per CTA, in TMEM)
// and B items [256b, +256) (128 per CTA, in shared memory), for every b >= a; four M256 N256 K32 MMAs per k-block
// instead of eight M256 N128 ones for the same MACs.  9 warps per CTA: 0-3 A rows, 4-7 B rows, 8 TMEM + (leader) MMAs.
// ------------------------------------------------------------------------------------------
#define TC4_THREADS 288
#define TC4_FULL_ARRIVALS 16   // producer warps of both CTAs

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC4_THREADS, 1)
k_pair_tc4(const uint32_t* __restrict__ bits, int64_t pitch_words, int F, int64_t kblocks_per_split, const int2* __restrict__ tiles,
           int32_t* __restrict__ P) {
    extern __shared__ uint8_t tc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_b = smem;                                   // [stage][128 rows][128 B]
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + TC_STAGES * TC_TILE_BYTES);
    uint64_t* empty = full + TC_STAGES;
    uint64_t* acc_ready = empty + TC_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int2 tl = tiles[blockIdx.x >> 1];
    const int i0 = tl.x * 256 + (int)rank * 128, j0 = tl.y * 256;
    const int64_t n_kb = pitch_words / 4;
    const int64_t kb0 = (int64_t)blockIdx.y * kblocks_per_split;
    const int64_t kb1 = min(n_kb, kb0 + kblocks_per_split);
    const int64_t nkb = kb1 - kb0;

    if (tid == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { tc::mbar_init(full + s, TC4_FULL_ARRIVALS); tc::mbar_init(empty + s, 1); }  // full: one arrival per producer warp of either CTA
        tc::mbar_init(acc_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc::smem_u32(tmem_slot)), "r"((uint32_t)TC2_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    tc::cluster_sync_all();   // both CTAs' barriers and TMEM exist before anyone signals or issues
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;   // columns: [0,256) accumulator (this CTA's 128 rows x 256), [256,384) A stages

    if (warp < 8) {
        const bool is_b = warp >= 4;
        const int r = tid & 127;                              // row inside the 128-row block this thread expands
        const int item = is_b ? j0 + (int)rank * 128 + r : i0 + r;
        const bool ok = item < F;
        const uint4* src = reinterpret_cast<const uint4*>(bits + (int64_t)(ok ? item : 0) * pitch_words) + kb0;
        const uint32_t a_lane_addr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + 256u;
        const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
        uint4 cur[4], nxt[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) cur[q] = (ok && q < nkb) ? __ldg(src + q) : zero;
        for (int64_t kb = 0; kb < nkb; kb += 4) {
            const uint32_t ph = (uint32_t)((kb >> 2) & 1);
#pragma unroll
            for (int q = 0; q < 4; ++q) nxt[q] = (ok && kb + 4 + q < nkb) ? __ldg(src + kb + 4 + q) : zero;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (kb + q < nkb) {
                    uint32_t v[32];
                    tc::expand_words(cur[q], v);
                    tc::mbar_wait(empty + q, ph ^ 1);
                    if (is_b) {
                        tc::store_row_smem(smem_b + q * TC_TILE_BYTES, r, v);
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    } else {
                        tc::store_row_tmem(a_lane_addr + (uint32_t)(q * 32), v);
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    }
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive_cluster(full + q, 0);     // the leader's barrier, from either CTA
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) cur[q] = nxt[q];
        }
        if (nkb > 0) {   // all eight warps: lane quarter warp % 4, warps 0-3 columns 0-127, warps 4-7 columns 128-255
            tc::mbar_wait(acc_ready, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int lane_base = (warp & 3) * 32;
            const int i = i0 + lane_base + lane;
#pragma unroll 1
            for (int c = (is_b ? 4 : 0); c < (is_b ? 8 : 4); ++c) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(c * 32);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                      "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                      "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int t = 0; t < 32; ++t) {
                    const int j = j0 + c * 32 + t;
                    if (i < j && j < F && v[t]) atomicAdd(P + (int64_t)i * F + j, (int32_t)v[t]);
                }
            }
        }
    } else if (rank == 0 && lane == 0) {
        const uint32_t idesc = tc::umma_idesc_i8(256, 256);
        for (int64_t kb = 0; kb < nkb; ++kb) {
            const int s = (int)(kb % TC_STAGES);
            const uint32_t ph = (uint32_t)((kb / TC_STAGES) & 1);
            tc::mbar_wait(full + s, ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t db = tc::umma_desc(tc::smem_u32(smem_b + s * TC_TILE_BYTES));
            const uint32_t ta = tmem_base + 256u + (uint32_t)(s * 32);
#pragma unroll
            for (int k = 0; k < TC_KB / 32; ++k)
                tc::umma_i8_ts_2sm(tmem_base, ta + 8 * k, db + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            tc::umma_commit_2sm(empty + s);
        }
        if (nkb > 0) tc::umma_commit_2sm(acc_ready);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    tc::cluster_sync_all();   // the leader's MMAs read the peer's TMEM and shared memory: nobody leaves early
    if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TC2_TMEM_COLS) : "memory");
}
