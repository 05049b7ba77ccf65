"""Pair classes and the folded pair-value table (reference: main.c:14-44, 59-60).

The reference builds two 27x27 byte adjacency tables on every rank and uploads them to
constant memory (cudaFunctions.cu:35-44). Here a single 32x32 int32 table with the weights
already applied is built once by the packer and travels inside the blob.
"""
from __future__ import annotations

import numpy as np

from . import _native as N

TYPE1_GROUPS = ("NDEQ", "MILV", "FYW", "NEQK", "QHRK", "HY", "STA", "NHQK", "MILF")
TYPE2_GROUPS = ("SAG", "SGND", "NEQHRK", "HFY", "ATV", "STPA", "NDEQHK", "FVLIM", "CSA", "STNK", "SNDEQK")


def pair_class(a: str | int, b: str | int) -> int:
    """0 identical, 1 same type-1 group, 2 same type-2 group, 3 otherwise."""
    ia = ord(a) if isinstance(a, str) else int(a)
    ib = ord(b) if isinstance(b, str) else int(b)
    return int(N.core().sa_pair_class(ia, ib))


def pair_class_by_scan(a: str, b: str) -> int:
    """Group-membership scan straight from the problem statement (test oracle for pair_class)."""
    if a == b:
        return 0
    if any(a in g and b in g for g in TYPE1_GROUPS):
        return 1
    if any(a in g and b in g for g in TYPE2_GROUPS):
        return 2
    return 3


def value_table(weights) -> np.ndarray:
    """26x26 int32 table of +w1 / -w2 / -w3 / -w4 for letters A..Z."""
    w = [int(x) for x in weights]
    vals = (w[0], -w[1], -w[2], -w[3])
    t = np.empty((26, 26), dtype=np.int32)
    for i in range(26):
        for j in range(26):
            t[i, j] = vals[pair_class(65 + i, 65 + j)]
    return t
