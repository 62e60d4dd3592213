"""GPU test of fa_count_groups_noalloc, the entry point for hosts that may not allocate device memory
(an application that owns the GPU through its own arena, a serving loop that cannot take cudaMalloc's
device-wide stall, a process at its memory limit).  It is called through the C-ABI on caller-owned device
buffers and must reproduce, exactly, the supports that the regular entry points (fa_count_candidates via
the level loop) produce — on two calls in a row, because every call re-uses the same static scratch array
and has to clear it itself.  Collected last on purpose (file name), so a -x run has already exercised
every other test before this one."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_WORDS = 1 << 20


def _pack_bitmaps(data, id_to_rank, F):
    """uint32 [F][pitch_words] in the library's layout: bit r&31 of word r>>5 of row i <=> row r holds rank i."""
    n = len(data)
    pitch = max(128, -(-(-(-n // 32)) // 128) * 128)
    bits = np.zeros((F, pitch), dtype=np.uint32)
    rows = np.repeat(np.arange(n), np.diff(data.offsets))
    ranks = id_to_rank[data.ids]
    ok = ranks >= 0
    np.bitwise_or.at(bits, (ranks[ok], rows[ok] >> 5), (np.uint32(1) << (rows[ok] & 31).astype(np.uint32)))
    return bits, pitch


def _recount(bits, items):
    """Support of an itemset straight from the packed rows: popcount of the AND of its item bitmaps."""
    w = np.bitwise_and.reduce(bits[list(items)], axis=0)
    return int(np.unpackbits(w.view(np.uint8)).sum())


def _call(lib, dev, bits_t, pitch, F, prefix_t, m, grow_t, goff_t, ext_t, n_groups, n_cands, out_t, want_alive=True, stream=None):
    alive = ctypes.c_int64(-1)
    err = ctypes.create_string_buffer(256)
    rc = lib.fa_count_groups_noalloc(dev, ctypes.c_void_p(stream) if stream else None, bits_t.data_ptr(), pitch, F, prefix_t.data_ptr(), m, grow_t.data_ptr(),
                                     goff_t.data_ptr(), ext_t.data_ptr(), n_groups, n_cands, out_t.data_ptr(),
                                     ctypes.byref(alive) if want_alive else None, err, 256)
    return rc, alive.value, err.value.decode()


def test_noalloc_matches_level_loop_twice_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from fastapriori_b200 import FastApriori, MiningContext, TokenizedLines, _native, quest

    lib = _native.load()
    sc = MiningContext(device=0)
    try:
        n_items = 14
        off, items = quest.generate(5000, 5, 3, 8, n_items, seed=31)
        data = TokenizedLines.from_int_csr(off, items, n_items)
        fa = FastApriori(1e-9)                                  # minCount = 1: every itemset that occurs at all is frequent
        fis, _, freq_items = fa.run(sc, data)                   # the existing entry points (fa_count_candidates per level)
        assert fa.counters["minCount"] == 1 and fis.max_k >= 4
        F = len(freq_items)
        want = fis.as_dict()
        bits, pitch = _pack_bitmaps(data, fa.id_to_rank, F)
        row_sets = [set(fa.id_to_rank[data.ids[data.offsets[r]:data.offsets[r + 1]]].tolist()) for r in range(len(data))]
        bits_t = torch.from_numpy(bits.view(np.int32)).cuda()

        def run_level(k):
            prefix, _ = fis.levels[k - 1]                       # every frequent (k-1)-set is a group; extensions: all larger ranks
            grow = np.arange(len(prefix), dtype=np.int32)
            ext = [np.arange(p[-1] + 1, F, dtype=np.int32) for p in prefix]
            goff = np.zeros(len(prefix) + 1, dtype=np.int32)
            np.cumsum([len(e) for e in ext], out=goff[1:])
            ext = np.concatenate(ext) if len(ext) else np.zeros(0, np.int32)
            keep = np.diff(goff) > 0                            # groups without extensions are dropped, like the level loop does
            grow, goff = grow[keep], np.concatenate(([0], np.cumsum(np.diff(goff)[keep]))).astype(np.int32)
            t = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (prefix, grow, goff, ext)]
            out = torch.full((max(len(ext), 1),), 12345, dtype=torch.int32, device="cuda")   # stale contents must not leak through
            free_before = torch.cuda.mem_get_info()[0]
            rc, alive, msg = _call(lib, 0, bits_t, pitch, F, t[0], k - 1, t[1], t[2], t[3], len(grow), len(ext), out)
            assert rc == 0, msg
            got = out.cpu().numpy()[:len(ext)]
            j = 0
            for g in grow.tolist():
                for y in range(prefix[g][-1] + 1, F):
                    cand = tuple(prefix[g].tolist()) + (y,)
                    assert got[j] == want.get(cand, 0), f"level {k}, group {g}, y {y}"   # the existing entry points' support
                    if j % 7 == 0:                                                       # and a direct recount from the rows
                        assert got[j] == _recount(bits, cand), f"level {k}, group {g}, y {y}: recount"
                    j += 1
            assert j == len(ext)
            groups = [set(prefix[g].tolist()) for g in grow.tolist()]
            assert alive == sum(1 for rs in row_sets if any(p <= rs for p in groups))
            # enqueue-only form (no alive total, so no wait inside the call) on a stream of the caller's
            st = torch.cuda.Stream()
            out2 = torch.full_like(out, -7)
            st.wait_stream(torch.cuda.current_stream())
            rc, _, msg = _call(lib, 0, bits_t, pitch, F, t[0], k - 1, t[1], t[2], t[3], len(grow), len(ext), out2, want_alive=False,
                               stream=st.cuda_stream)
            assert rc == 0, msg
            st.synchronize()
            assert np.array_equal(out2.cpu().numpy()[:len(ext)], got)
            if k == 4:  # no device memory was taken by either call (the first call of the process may load the module)
                assert torch.cuda.mem_get_info()[0] == free_before
            assert torch.cuda.current_device() == 0
            return len(ext)

        n3 = run_level(3)          # first call on this device: performs the symbol-address set-up
        n4 = run_level(4)          # second call: other sizes, same static scratch
        assert n3 > 0 and n4 > 0
        run_level(3)               # and back again

        # groups whose prefixes have no extension at all: nothing to count, but their tidsets still define the alive rows
        prefix, _ = fis.levels[2]
        t = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in
             (prefix, np.arange(len(prefix), dtype=np.int32), np.zeros(len(prefix) + 1, dtype=np.int32))]
        rc, alive, msg = _call(lib, 0, bits_t, pitch, F, t[0], 2, t[1], t[2], t[2], len(prefix), 0, t[2])
        assert rc == 0, msg
        assert alive == sum(1 for rs in row_sets if any(set(p) <= rs for p in prefix.tolist()))

        # long prefixes (more than 32 items: the kernel's second staging half, 64 KB of shared memory) and a group with
        # more than 32 extensions; far beyond what a level loop reaches here, so the check is the recount alone
        rng = np.random.default_rng(5)
        F2, n2 = 96, 3000
        dense = rng.random((n2, F2)) < 0.35
        dense[: n2 // 2, :41] = True                       # half of the rows hold items 0..40
        pitch2 = 128
        bits2 = np.zeros((F2, pitch2), dtype=np.uint32)
        rr, ii = np.nonzero(dense)
        np.bitwise_or.at(bits2, (ii, rr >> 5), (np.uint32(1) << (rr & 31).astype(np.uint32)))
        prefixes = np.stack([np.arange(40), np.concatenate((np.arange(39), [40]))]).astype(np.int32)   # 40 items each
        exts = [np.arange(41, F2, dtype=np.int32), np.arange(50, 60, dtype=np.int32)]                  # 55 and 10 extensions
        goff2 = np.array([0, len(exts[0]), len(exts[0]) + len(exts[1])], dtype=np.int32)
        t2 = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in
              (bits2.view(np.int32), prefixes, np.arange(2, dtype=np.int32), goff2, np.concatenate(exts))]
        out = torch.full((int(goff2[-1]),), 99, dtype=torch.int32, device="cuda")
        for _ in range(2):
            rc, alive, msg = _call(lib, 0, t2[0], pitch2, F2, t2[1], 40, t2[2], t2[3], t2[4], 2, int(goff2[-1]), out)
            assert rc == 0, msg
            got = out.cpu().numpy()
            j = 0
            for g in range(2):
                for y in exts[g].tolist():
                    assert got[j] == _recount(bits2, prefixes[g].tolist() + [y]), f"long prefix, group {g}, y {y}"
                    j += 1
            assert alive == int((dense[:, prefixes[0]].all(axis=1) | dense[:, prefixes[1]].all(axis=1)).sum())
            assert got[0] >= n2 // 2 * 0.3                 # the planted rows make these supports non-trivial
        rc, _, msg = _call(lib, 0, t2[0], pitch2, F2, t2[1], 65, t2[2], t2[3], t2[4], 2, int(goff2[-1]), out)
        assert rc == 4 and "prefix_len" in msg              # FA_ELIMIT: at most 64 items per prefix

        # sizes beyond the static scratch are refused with a clear error, before anything is launched
        dummy = torch.zeros(8, dtype=torch.int32, device="cuda")
        rc, _, msg = _call(lib, 0, bits_t, MAX_WORDS + 4, F, dummy, 2, dummy, dummy, dummy, 1, 1, dummy)
        assert rc == 4 and "exceeds the static scratch" in msg
        rc, _, msg = _call(lib, 0, bits_t, pitch + 2, F, dummy, 2, dummy, dummy, dummy, 1, 1, dummy)
        assert rc == 1 and "multiple of 4" in msg
    finally:
        sc.close()
