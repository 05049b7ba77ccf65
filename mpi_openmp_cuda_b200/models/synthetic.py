"""Synthetic workloads in the reference's input format (the format caps
sequence lengths at 3000/2000, so "scaled-up" means more sequences)."""
from __future__ import annotations

import numpy as np

from ..utils.parser import Problem

_LETTERS = np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZ", dtype=np.uint8)


def random_problem(nseq: int, len1: int = 1489, len2_range=(56, 1152), weights=(2, 2, 1, 10),
                   seed: int = 0, lowercase_fraction: float = 0.0) -> Problem:
    """Random letters; defaults mirror input3.txt's shape (len1 1489, Seq2 56..1152)."""
    rng = np.random.default_rng(seed)
    s1 = _LETTERS[rng.integers(0, 26, size=len1)].tobytes()
    lo, hi = len2_range
    lens = rng.integers(lo, hi + 1, size=nseq)
    seqs = []
    for l in lens:
        s = _LETTERS[rng.integers(0, 26, size=int(l))].tobytes()
        if lowercase_fraction > 0:
            arr = bytearray(s)
            for i in np.nonzero(rng.random(len(arr)) < lowercase_fraction)[0]:
                arr[i] |= 0x20
            s = bytes(arr)
        seqs.append(s)
    return Problem.from_parts(weights, s1, seqs)


def scaled_input3(scale: int, base_path, seed: int = 0) -> Problem:
    """`scale` x the sequences of a shipped input: the originals first, then mutated copies
    (random letter substitutions) so that no two rows are identical work."""
    base = Problem.from_file(base_path)
    rng = np.random.default_rng(seed)
    seqs = [base.seq2(i) for i in range(base.nseq)]
    out = list(seqs)
    for rep in range(1, scale):
        for s in seqs:
            arr = bytearray(s)
            for i in rng.integers(0, len(arr), size=max(1, len(arr) // 16)):
                arr[i] = int(_LETTERS[rng.integers(0, 26)])
            out.append(bytes(arr))
    return Problem.from_parts(base.weights, base.seq1, out)
