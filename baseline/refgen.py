"""Pure-Python (numpy only) problem generator, candidate counter and oracle for the reference arm.

`bench.py --impl reference` must not load any native code of this repository, so everything the
arm needs besides the reference binary itself lives here: the input grammar (main.c:76-108), the
scaled-input generator (byte-identical to `mpi_openmp_cuda_b200.models.synthetic.scaled_input3`,
an independent implementation; tests/test_refgen.py holds the two together), the candidate
count (cudaFunctions.cu:116-118) and a numpy restatement of the scoring contract
used to count how many rows the reference binary gets right.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

_LETTERS = np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZ", dtype=np.uint8)
# The two group lists of the problem statement (main.c:59-60).
TYPE1 = ("NDEQ", "MILV", "FYW", "NEQK", "QHRK", "HY", "STA", "NHQK", "MILF")
TYPE2 = ("SAG", "SGND", "NEQHRK", "HFY", "ATV", "STPA", "NDEQHK", "FVLIM", "CSA", "STNK", "SNDEQK")
INT_MIN = -(1 << 31)


def parse_text(data: bytes):
    """-> (weights[4], seq1, [seq2...]) with ASCII a-z upper-cased; whitespace-separated tokens."""
    tok = data.split()
    if len(tok) < 6:
        raise ValueError("input: missing tokens")
    weights = [int(t) for t in tok[:4]]
    seq1 = tok[4].upper()
    n = int(tok[5])
    seqs = [t.upper() for t in tok[6:6 + n]]
    if len(seqs) != n:
        raise ValueError("input: fewer sequences than announced")
    return weights, seq1, seqs


def serialize(weights, seq1: bytes, seqs) -> bytes:
    head = b" ".join(str(int(w)).encode() for w in weights) + b"\n" + seq1 + b"\n" + str(len(seqs)).encode() + b"\n"
    return head + b"".join(s + b"\n" for s in seqs)


def scaled_parts(base_path, scale: int, seed: int = 0):
    """`scale` x the sequences of a shipped input: the originals first, then mutated copies
    (random letter substitutions) so that no two rows are identical work."""
    weights, seq1, seqs = parse_text(Path(base_path).read_bytes())
    rng = np.random.default_rng(seed)
    out = list(seqs)
    for _ in range(1, scale):
        for s in seqs:
            arr = bytearray(s)
            for i in rng.integers(0, len(arr), size=max(1, len(arr) // 16)):
                arr[i] = int(_LETTERS[rng.integers(0, 26)])
            out.append(bytes(arr))
    return weights, seq1, out


def scaled_text(base_path, scale: int, seed: int = 0) -> bytes:
    if scale == 1:
        return Path(base_path).read_bytes()
    return serialize(*scaled_parts(base_path, scale, seed))


def candidates(seq1: bytes, seqs) -> int:
    """(len1-len2)*len2 per sequence, 1 when the lengths are equal, 0 when Seq2 is longer."""
    l1, tot = len(seq1), 0
    for s in seqs:
        m = len(s)
        tot += 1 if m == l1 else (l1 - m) * m if m < l1 else 0
    return tot


def value_table(weights) -> np.ndarray:
    """int32[256][256]: +w1 identical, -w2 same type-1 group, -w3 same type-2 group, -w4 otherwise."""
    w1, w2, w3, w4 = (int(w) for w in weights)
    cls = np.full((256, 256), 3, dtype=np.int64)
    for groups, c in ((TYPE2, 2), (TYPE1, 1)):  # type 1 wins over type 2
        for g in groups:
            idx = np.frombuffer(g.encode(), dtype=np.uint8)
            cls[np.ix_(idx, idx)] = c
    cls[np.arange(256), np.arange(256)] = 0
    return np.array([w1, -w2, -w3, -w4], dtype=np.int64)[cls].astype(np.int32)


def solve(weights, seq1: bytes, seqs):
    """[(score, n, k)] per sequence: first maximum in n-major, k-minor order, k = 0 first."""
    lut = value_table(weights)
    s1 = np.frombuffer(seq1, dtype=np.uint8)
    l1 = len(s1)
    res = []
    for s in seqs:
        s2 = np.frombuffer(s, dtype=np.uint8)
        m = len(s2)
        if m > l1 or (m == 0 and l1 > 0):
            res.append((INT_MIN, 0, 0))
            continue
        if m == l1:
            res.append((int(lut[s2, s1].sum()), 0, 0))
            continue
        nd = l1 - m
        j = np.arange(m)[:, None] + np.arange(nd + 1)[None, :]
        d = lut[s2[:, None], s1[j]].astype(np.int64)  # d[i][n] = value(seq2[i], seq1[i+n])
        t = d.sum(axis=0)
        sc = np.empty((m, nd), dtype=np.int64)  # sc[k][n]
        sc[0] = t[:nd]
        if m > 1:
            sc[1:] = t[1:nd + 1][None, :] + np.cumsum(d[:-1, :nd] - d[:-1, 1:nd + 1], axis=0)
        flat = np.ascontiguousarray(sc.T).reshape(-1)  # n-major, k-minor
        a = int(np.argmax(flat))
        res.append((int(flat[a]), a // m, a % m))
    return res


def format_results(rows) -> bytes:
    return b"".join(b"#%d: score: %d, n: %d, k: %d\n" % (i, r[0], r[1], r[2]) for i, r in enumerate(rows))
