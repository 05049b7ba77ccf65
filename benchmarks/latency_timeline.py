#!/usr/bin/env python
"""Where do the microseconds of a small single-GPU solve go? Runs the TILE kernel with
SA_TIMELINE=1 (phase timestamps from %globaltimer, latest CTA per phase of its first item) with a
cold L2, next to the CUDA-event time of the same launch.

    SA_TIMELINE=1 python benchmarks/latency_timeline.py [input3.txt input5.txt ...]
"""
from __future__ import annotations

import json
import os
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
os.environ.setdefault("SA_TIMELINE", "1")

import torch  # noqa: E402

from mpi_openmp_cuda_b200 import Problem  # noqa: E402
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver  # noqa: E402
from mpi_openmp_cuda_b200.utils.timing import L2Flusher  # noqa: E402

NAMES = {0: "first_cta_start", 11: "first_item_known", 12: "first_tma_landed", 13: "first_scan_done",
         14: "first_key_merged", 15: "first_result_written", 2: "last_item_done", 16: "last_cta_exit"}


def main() -> int:
    inputs = sys.argv[1:] or ["input3.txt", "input4.txt", "input1.txt", "input5.txt"]
    solver = AlignmentSolver("cuda:0")
    flusher = L2Flusher(solver.device)
    lib = solver.exec.lib
    rows = []
    for name in inputs:
        p = Problem.from_file(ROOT / "data" / name)
        blob = solver.pack(p)
        solver.solve_packed(blob)
        ex = solver.exec
        off = int(lib.sa_cuda_scratch_bytes(blob.nseq)) - 256 + 64  # control block, timeline words
        tl_view = ex._scratch[off : off + 8 * 24].view(torch.int64)
        per_phase = {k: [] for k in NAMES}
        ev_ms = []
        for it in range(12):
            flusher.flush()
            tl_view.zero_()
            tl_view[0] = torch.iinfo(torch.int64).max
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ex.run(blob.header)
            e1.record()
            e1.synchronize()
            t = tl_view.cpu().tolist()
            tl_view.zero_()
            if it >= 2:
                ev_ms.append(e0.elapsed_time(e1))
                for k in NAMES:
                    per_phase[k].append((t[k] - t[0]) / 1e3 if t[k] else None)
        row = {"input": name, "items": blob.nitems, "event_us": round(statistics.median(ev_ms) * 1e3, 2)}
        for k, nm in NAMES.items():
            vals = [v for v in per_phase[k] if v is not None]
            row[nm + "_us"] = round(statistics.median(vals), 2) if vals else None
        rows.append(row)
        print(json.dumps(row), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
