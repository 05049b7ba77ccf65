#!/usr/bin/env python
"""Tiny driver for ncu / compute-sanitizer: a few solves of one configuration, nothing else.

    python benchmarks/profile_target.py --scale 256 --iters 3      # throughput shape
    python benchmarks/profile_target.py --input input3.txt         # latency shape
"""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402

from mpi_openmp_cuda_b200 import Problem  # noqa: E402
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver  # noqa: E402
from mpi_openmp_cuda_b200.models.synthetic import scaled_input3  # noqa: E402
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--input", default="input3.txt")
ap.add_argument("--scale", type=int, default=1)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--check", action="store_true")
a = ap.parse_args()

path = ROOT / "data" / a.input
p = Problem.from_file(path) if a.scale == 1 else scaled_input3(a.scale, path, seed=1)
solver = AlignmentSolver("cuda:0")
blob = solver.pack(p)
for _ in range(a.iters):
    res = solver.solve_packed(blob)
if a.check:
    assert np.array_equal(res, solve_cpu(p, "fast"))
print("ok", p.nseq, p.candidates(), blob.nitems, int(blob.header.padded_cells))
