"""TokenizedLines — the host-side stand-in for the reference's RDD[Array[String]]
(Utils.scala:19-27): lines of string tokens stored as CSR over a string vocabulary.
"""
from __future__ import annotations

import re

import numpy as np

_WS = re.compile(rb"[ \t\n\x0b\f\r]+")
_TRIM = bytes(range(0x21))


class TokenizedLines:
    """offsets[int64, n+1], ids[int32, nnz] index into vocab[list[str]] (dense ids)."""

    def __init__(self, offsets, ids, vocab):
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        # ids are int32; a uint16 array is kept as it is (compact(): the mining entry point then ships 2 bytes per id)
        self.ids = np.ascontiguousarray(ids, dtype=np.uint16 if getattr(ids, "dtype", None) == np.uint16 else np.int32)
        self.vocab = list(vocab)

    def compact(self, out=None):
        """The same lines with 16-bit ids (vocabularies of at most 65 536 tokens), optionally written into `out` (e.g. a
        pinned buffer); returns self when the vocabulary is larger."""
        if len(self.vocab) > 65536:
            return self
        ids16 = out if out is not None else np.empty(len(self.ids), dtype=np.uint16)
        np.copyto(ids16, self.ids, casting="unsafe")
        return TokenizedLines(self.offsets, ids16, self.vocab)

    def __len__(self):
        return len(self.offsets) - 1

    def token_index(self) -> dict:
        """token string -> dense id, built once per object (the rule stage maps the F frequent items into the users'
        vocabulary on every run; walking the whole vocabulary instead cost ~1 ms per run at 10 k items)."""
        idx = getattr(self, "_token_index", None)
        if idx is None or len(idx) != len(self.vocab):
            idx = self._token_index = {t: j for j, t in enumerate(self.vocab)}
        return idx

    def count(self) -> int:
        return len(self)

    def line(self, i: int):
        return [self.vocab[j] for j in self.ids[self.offsets[i]:self.offsets[i + 1]]]

    @staticmethod
    def from_lines(lines):
        """lines: iterable of iterables of str tokens."""
        vocab, index, ids, off = [], {}, [], [0]
        for line in lines:
            for t in line:
                j = index.get(t)
                if j is None:
                    j = index[t] = len(vocab)
                    vocab.append(t)
                ids.append(j)
            off.append(len(ids))
        return TokenizedLines(np.array(off, dtype=np.int64), np.array(ids, dtype=np.int32), vocab)

    @staticmethod
    def from_text(buf: bytes):
        """Native loader (fa_text_parse in libfastapriori_b200.so): same rules as from_text_py, ~100x faster."""
        import ctypes
        from . import _native
        lib = _native.load()
        h = ctypes.c_void_p()
        rc = lib.fa_text_parse(buf, len(buf), ctypes.byref(h))
        if rc != 0:
            raise _native.NativeError(rc, "fa_text_parse failed")
        try:
            n, nnz, nv = lib.fa_text_rows(h), lib.fa_text_nnz(h), lib.fa_text_vocab_size(h)
            off = np.empty(n + 1, dtype=np.int64)
            ids = np.empty(nnz, dtype=np.int32)
            vb = ctypes.create_string_buffer(max(int(lib.fa_text_vocab_bytes(h)), 1))
            lib.fa_text_copy(h, off.ctypes.data, ids.ctypes.data if nnz else None, vb)
            vocab = vb.raw[:int(lib.fa_text_vocab_bytes(h))].split(b"\n")[:nv] if nv else []
        finally:
            lib.fa_text_free(h)
        return TokenizedLines(off, ids, [v.decode("utf-8", "surrogateescape") for v in vocab])

    @staticmethod
    def from_text_py(buf: bytes):
        """Reference implementation in Python of the same rules (kept to cross-check the native loader).
        Hadoop TextInputFormat records + `_.trim().split("\\\\s+")` (Utils.scala:21): records end at
        \\n, \\r\\n or \\r, a trailing terminator adds no record, Java trim strips chars <= U+0020,
        and a blank line is the one-token line [""]."""
        if not buf:
            return TokenizedLines([0], [], [])
        recs = re.split(rb"\r\n|\n|\r", buf)
        if recs and recs[-1] == b"":
            recs.pop()
        vocab, index, ids, off = [], {}, [], [0]
        for r in recs:
            r = r.strip(_TRIM)
            toks = _WS.split(r) if r else [b""]
            for t in toks:
                j = index.get(t)
                if j is None:
                    j = index[t] = len(vocab)
                    vocab.append(t)
                ids.append(j)
            off.append(len(ids))
        return TokenizedLines(np.array(off, dtype=np.int64), np.array(ids, dtype=np.int32),
                              [v.decode("utf-8", "surrogateescape") for v in vocab])

    @staticmethod
    def from_int_csr(offsets, item_ints, n_items: int):
        """Rows of integer item ids in 1..n_items (the synthetic generator's output): dense id =
        item - 1, vocab = decimal strings.  No per-token Python work."""
        ids = np.ascontiguousarray(item_ints, dtype=np.int32) - 1
        return TokenizedLines(offsets, ids, [str(i) for i in range(1, n_items + 1)])
