"""Result printer: `#i: score: S, n: N, k: K` per sequence, input order (main.c:199-211)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ..ops import _native as N


def format_results(results) -> bytes:
    """results: array-like [nseq, 3] of (score, n, k) -> the exact stdout bytes."""
    arr = np.ascontiguousarray(np.asarray(results, dtype=np.int32).reshape(-1, 3))
    n = arr.shape[0]
    ptr = arr.ctypes.data_as(C.POINTER(C.c_int32))
    need = N.core().sa_format(ptr, n, None, 0)
    buf = C.create_string_buffer(need + 1)
    N.core().sa_format(ptr, n, buf, need)
    return buf.raw[:need]
