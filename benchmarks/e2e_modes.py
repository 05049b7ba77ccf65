#!/usr/bin/env python
"""End-to-end solve latency of the native session under its transfer policies."""
import statistics
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from mpi_openmp_cuda_b200 import Problem  # noqa: E402
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver  # noqa: E402
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu  # noqa: E402

solver = AlignmentSolver("cuda:0")
for name in ("input1.txt", "input3.txt", "input4.txt"):
    p = Problem.from_file(ROOT / "data" / name)
    ref = solve_cpu(p)
    for zc, direct in ((0, False), (0, True), (128 << 10, False), (128 << 10, True)):
        solver.session.configure(zc, direct)
        for _ in range(20):
            r = solver.solve(p)
        assert np.array_equal(r, ref)
        ts = []
        for _ in range(200):
            t0 = time.perf_counter()
            solver.solve(p)
            ts.append(time.perf_counter() - t0)
        print(f"{name} zero_copy_in={zc > 0} direct_out={direct}: median {statistics.median(ts) * 1e6:.1f} us  min {min(ts) * 1e6:.1f} us")
