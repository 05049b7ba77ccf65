"""CPU oracle entry points (race-free restatement of cudaFunctions.cu:63-176)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N


def solve_cpu(problem, method: str = "fast", threads: int = 0) -> np.ndarray:
    """Return int32[nseq, 3] = (score, n, k). method: 'fast' O(l1*l2) or 'naive' O(l1*l2^2)."""
    n = problem.nseq
    out = np.zeros((max(n, 1), 3), dtype=np.int32)
    fn = {"fast": N.core().sa_solve_cpu_fast, "naive": N.core().sa_solve_cpu_naive}[method]
    N.check(fn(problem.handle, out.ctypes.data_as(C.POINTER(C.c_int32)), int(threads)))
    return out[:n]
