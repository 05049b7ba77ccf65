#!/usr/bin/env python
"""BASELINE.json configs[1]: input1..6 on one B200 — kernel time, end-to-end time, byte-identical check."""
import json
import statistics
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from mpi_openmp_cuda_b200 import Problem, format_results  # noqa: E402
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver  # noqa: E402
from mpi_openmp_cuda_b200.utils.timing import L2Flusher  # noqa: E402

solver = AlignmentSolver("cuda:0")
fl = L2Flusher("cuda:0")
rows = []
for i in range(1, 7):
    path = ROOT / "data" / f"input{i}.txt"
    gold = (ROOT / "data" / "golden" / f"input{i}.out").read_bytes()
    text = path.read_bytes()
    p = Problem.parse(text)
    blob = solver.pack(p)
    ok = format_results(solver.solve(p)) == gold and format_results(solver.solve_packed(blob)) == gold  # also uploads
    ks = []
    for _ in range(30):
        fl.flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        solver.exec.run(blob.header)
        e1.record()
        e1.synchronize()
        ks.append(e0.elapsed_time(e1) * 1e3)
    es, ts = [], []
    for _ in range(100):
        fl.flush()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        solver.solve(p)
        es.append((time.perf_counter() - t0) * 1e6)
        t0 = time.perf_counter()
        format_results(solver.solve(Problem.parse(text)))
        ts.append((time.perf_counter() - t0) * 1e6)
    row = {"input": i, "nseq": p.nseq, "candidates": p.candidates(), "byte_identical": ok,
           "kernel_us": round(statistics.fmean(ks), 1), "kernel_us_min": round(min(ks), 1),
           "solve_us": round(statistics.median(es), 1), "parse_solve_format_us": round(statistics.median(ts), 1)}
    rows.append(row)
    print(json.dumps(row), flush=True)
out = ROOT / "gpurun_out" / "suite_ours.json"
out.parent.mkdir(exist_ok=True)
out.write_text(json.dumps(rows, indent=1))
