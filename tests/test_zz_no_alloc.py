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


def _call(lib, dev, bits_t, pitch, F, prefix_t, m, grow_t, goff_t, ext_t, n_groups, n_cands, out_t, want_alive=True):
    alive = ctypes.c_int64(-1)
    err = ctypes.create_string_buffer(256)
    rc = lib.fa_count_groups_noalloc(dev, None, bits_t.data_ptr(), pitch, F, prefix_t.data_ptr(), m, grow_t.data_ptr(),
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
            rc, alive, msg = _call(lib, 0, bits_t, pitch, F, t[0], k - 1, t[1], t[2], t[3], len(grow), len(ext), out)
            assert rc == 0, msg
            got = out.cpu().numpy()[:len(ext)]
            j = 0
            for g in grow.tolist():
                for y in range(prefix[g][-1] + 1, F):
                    assert got[j] == want.get(tuple(prefix[g].tolist()) + (y,), 0), f"level {k}, group {g}, y {y}"
                    j += 1
            assert j == len(ext)
            groups = [set(prefix[g].tolist()) for g in grow.tolist()]
            assert alive == sum(1 for rs in row_sets if any(p <= rs for p in groups))
            return len(ext)

        n3 = run_level(3)          # first call on this device: performs the symbol-address set-up
        n4 = run_level(4)          # second call: other sizes, same static scratch
        assert n3 > 0 and n4 > 0
        run_level(3)               # and back again

        # sizes beyond the static scratch are refused with a clear error, before anything is launched
        dummy = torch.zeros(8, dtype=torch.int32, device="cuda")
        rc, _, msg = _call(lib, 0, bits_t, MAX_WORDS + 4, F, dummy, 2, dummy, dummy, dummy, 1, 1, dummy)
        assert rc == 4 and "exceeds the static scratch" in msg
        rc, _, msg = _call(lib, 0, bits_t, pitch + 2, F, dummy, 2, dummy, dummy, dummy, 1, 1, dummy)
        assert rc == 1 and "multiple of 4" in msg
    finally:
        sc.close()
