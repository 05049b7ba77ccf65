#!/usr/bin/env python
"""Kernel-only timing of one configuration (CUDA events, L2 flushed between launches)."""
import argparse
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from mpi_openmp_cuda_b200 import Problem  # noqa: E402
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver  # noqa: E402
from mpi_openmp_cuda_b200.models.synthetic import scaled_input3  # noqa: E402
from mpi_openmp_cuda_b200.utils.timing import L2Flusher  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--input", default="input3.txt")
ap.add_argument("--scales", type=int, nargs="+", default=[1, 256, 1024])
ap.add_argument("--modes", nargs="+", default=["tile", "pair", "stream"])
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
solver = AlignmentSolver("cuda:0")
fl = L2Flusher("cuda:0")
for scale in a.scales:
    path = ROOT / "data" / a.input
    p = Problem.from_file(path) if scale == 1 else scaled_input3(scale, path, seed=1)
    for mode in a.modes:
        blob = solver.pack(p, mode=mode)
        solver.solve_packed(blob)
        ts = []
        for _ in range(a.iters):
            fl.flush()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            solver.exec.run(blob.header)
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = statistics.fmean(ts)
        print(f"{a.input} x{scale} mode={mode}: {ms * 1e3:.1f} us  min {min(ts) * 1e3:.1f} us  "
              f"cand/s {p.candidates() / ms / 1e6:.1f}e9  padded cells/s {blob.header.padded_cells / ms / 1e6:.1f}e9  items {blob.nitems}")
