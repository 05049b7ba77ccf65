"""Input grammar of the reference (main.c:76-108) via the native parser.

`Problem` owns a native `sa_problem*`; it is what every solver entry point takes.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Sequence

from ..ops import _native as N


class Problem:
    """Weights + Seq1 + N x Seq2, upper-cased exactly like main.c:83-87,102-106."""

    def __init__(self, handle: int):
        self._h = C.c_void_p(handle)

    # -- construction
    @classmethod
    def parse(cls, text: bytes | str) -> "Problem":
        data = text.encode("latin-1") if isinstance(text, str) else bytes(text)
        out = C.c_void_p()
        N.check(N.core().sa_parse(data, len(data), C.byref(out)))
        return cls(out.value)

    @classmethod
    def from_file(cls, path) -> "Problem":
        with open(path, "rb") as f:
            return cls.parse(f.read())

    @classmethod
    def from_parts(cls, weights: Sequence[int], seq1: bytes | str, seqs: Iterable[bytes | str]) -> "Problem":
        enc = lambda s: s.encode("latin-1") if isinstance(s, str) else bytes(s)
        s1 = enc(seq1)
        seq_list = [enc(s) for s in seqs]
        n = len(seq_list)
        w = (C.c_int32 * 4)(*[int(x) for x in weights])
        ptrs = (C.c_char_p * max(n, 1))(*seq_list) if n else (C.c_char_p * 1)()
        lens = (C.c_size_t * max(n, 1))(*[len(s) for s in seq_list]) if n else (C.c_size_t * 1)()
        out = C.c_void_p()
        N.check(N.core().sa_problem_create(w, s1, len(s1), n, ptrs, lens, C.byref(out)))
        return cls(out.value)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                N.core().sa_problem_free(h)
            except Exception:
                pass
            self._h = None

    # -- accessors
    @property
    def handle(self) -> C.c_void_p:
        return self._h

    @property
    def nseq(self) -> int:
        n = getattr(self, "_nseq", None)  # a parsed problem never changes: ask the library once
        if n is None:
            n = self._nseq = int(N.core().sa_problem_nseq(self._h))
        return n

    @property
    def weights(self) -> tuple[int, int, int, int]:
        w = (C.c_int32 * 4)()
        N.core().sa_problem_weights(self._h, w)
        return tuple(int(x) for x in w)

    @property
    def seq1(self) -> bytes:
        n = N.core().sa_problem_len1(self._h)
        return C.string_at(N.core().sa_problem_seq1(self._h), n)

    def seq2(self, i: int) -> bytes:
        if not 0 <= i < self.nseq:
            raise IndexError(i)
        n = N.core().sa_problem_len2(self._h, i)
        return C.string_at(N.core().sa_problem_seq2(self._h, i), n)

    def len2(self, i: int) -> int:
        return int(N.core().sa_problem_len2(self._h, i))

    def candidates(self, i: int | None = None) -> int:
        if i is None:
            return int(N.core().sa_problem_total_candidates(self._h))
        return int(N.core().sa_problem_candidates(self._h, i))

    def serialize(self) -> bytes:
        need = N.core().sa_problem_serialize(self._h, None, 0)
        buf = C.create_string_buffer(need)
        N.core().sa_problem_serialize(self._h, buf, need)
        return buf.raw[:need]

    def __repr__(self) -> str:
        return f"Problem(weights={self.weights}, len1={len(self.seq1)}, nseq={self.nseq}, candidates={self.candidates()})"


def parse_text(text: bytes | str) -> Problem:
    return Problem.parse(text)
