"""Multi-GPU timing for bench.py (one process per GPU under torchrun).

Device-timed with CUDA events per rank, max over ranks per step. Rank 0's launch ends only when
every rank has delivered its results into rank 0's table, so its events already span broadcast +
scoring + gather; the max over ranks is still taken, as the contract asks.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist

from ..ops import plan
from ..ops.score import launch_count
from ..utils.clocks import ClockSampler
from ..utils.timing import L2Flusher
from .dist_solver import DistributedSolver
from .fabric import Fabric, init_distributed


class _Align:
    """Device-side alignment of the ranks before a timed step: a tiny all-reduce enqueued on the
    stream (no host synchronisation), so that every GPU reaches the timed launch within
    microseconds of the others and host launch skew is not billed to the solve."""

    def __init__(self):
        self.t = torch.zeros(1, device="cuda")

    def __call__(self):
        dist.all_reduce(self.t)


def _max_over_ranks(ms: list[float]) -> list[float]:
    t = torch.tensor(ms, dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.cpu()]


def time_ours_distributed(problem, steps: int, warmup: int, large_problem=None, strong_problem=None):
    rank, world, local = init_distributed("nccl")
    # capacity handshake (bootstrap, outside any timed region): only rank 0 knows the problem
    need = torch.zeros(3, dtype=torch.int64, device="cuda")
    if rank == 0:
        need[0] = max(plan.blob_size(problem, world), plan.blob_size(large_problem, world) if large_problem is not None else 0)
        need[1] = max(problem.nseq, large_problem.nseq if large_problem is not None else 0)
        need[2] = large_problem.nseq if large_problem is not None else 0
    dist.broadcast(need, 0)
    fabric = Fabric(max_seq=max(int(need[1].item()), 64), max_blob_bytes=int(need[0].item()) + (1 << 20))
    solver = DistributedSolver(fabric)
    flusher = L2Flusher(fabric.device)
    blob = solver.pack(problem) if rank == 0 else None
    nseq = problem.nseq if rank == 0 else 0
    t_n = torch.tensor([nseq], dtype=torch.int64, device="cuda")
    dist.broadcast(t_n, 0)
    nseq = int(t_n.item())
    align = _Align()

    def step_kernel_only():
        # blob already resident in rank 0's symmetric buffer: launch only
        if rank == 0:
            fabric.launch(blob.header)
        else:
            fabric.launch(None)

    # first solve (uploads the blob on rank 0) + correctness snapshot
    solver.launch_packed(blob)
    res = solver.fetch(nseq)
    res = res.copy() if rank == 0 else None
    solver.check()

    for _ in range(warmup):
        flusher.flush()
        align()
        step_kernel_only()
    torch.cuda.synchronize()
    sampler = ClockSampler(local).start()
    l0 = launch_count()
    kernel_ms = []
    dist.barrier()
    torch.cuda.synchronize()
    for _ in range(steps):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # The L2 flush (~100 us of GPU work) goes first so that the host has the alignment
        # all-reduce, the start event and the launch queued before the GPU gets to them: the
        # timed region then holds no host launch latency, only device time.
        flusher.flush()
        align()
        e0.record()
        step_kernel_only()
        e1.record()
        e1.synchronize()
        kernel_ms.append(e0.elapsed_time(e1))
    launches = launch_count() - l0
    torch.cuda.synchronize()
    solver.check()
    if fabric.timeline:
        import sys

        tl = [None] * world
        dist.all_gather_object(tl, fabric.read_timeline())
        if rank == 0:
            print("fabric timeline (us since each rank's first CTA; last timed step):", tl, file=sys.stderr)

    def e2e_step(prepacked: bool) -> float:
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        flusher.flush()
        align()
        e0.record()
        if rank == 0:
            b = blob if prepacked else solver.pack(problem)
            solver.launch_packed(b)
            solver.fetch(nseq)
        else:
            solver.launch_packed(None)
            solver.fetch(0)
        e1.record()
        e1.synchronize()
        return e0.elapsed_time(e1)

    for _ in range(warmup):
        e2e_step(False)
    e2e_ms = [e2e_step(False) for _ in range(steps)]
    pre_ms = [e2e_step(True) for _ in range(steps)]
    clocks = sampler.stop()
    all_clocks = [None] * world
    dist.all_gather_object(all_clocks, clocks)
    solver.check()
    # ---- optional: the same fabric path with milliseconds of work per GPU (fixed costs amortised)
    large = None
    if int(need[2].item()) > 0:
        lblob = solver.pack(large_problem) if rank == 0 else None
        solver.launch_packed(lblob)
        lres = solver.fetch(int(need[2].item()))
        lres = lres.copy() if rank == 0 else None
        lms = []
        lsampler = ClockSampler(local).start()
        for _ in range(max(3, min(steps, 10)) + 2):
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flusher.flush()
            align()
            e0.record()
            fabric.launch(lblob.header if rank == 0 else None, resident=os.environ.get("SA_BENCH_RESIDENT") == "1")
            e1.record()
            e1.synchronize()
            lms.append(e0.elapsed_time(e1))
        lclocks = [None] * world
        dist.all_gather_object(lclocks, lsampler.stop())
        own_ms = [round(x, 4) for x in lms[2:]]
        per_rank_ms = [None] * world
        dist.all_gather_object(per_rank_ms, sum(own_ms) / len(own_ms))
        lms = _max_over_ranks(lms[2:])
        solver.check()
        if fabric.timeline:
            import sys

            tl = [None] * world
            dist.all_gather_object(tl, fabric.read_timeline())
            if rank == 0:
                print("fabric timeline, large config (us since each rank's first CTA; last step):", tl, file=sys.stderr)
        if rank == 0:
            large = {"nseq": large_problem.nseq, "candidates": large_problem.candidates(), "kernel_ms": sum(lms) / len(lms),
                     "kernel_ms_min": min(lms), "result": lres, "h2d_bytes": lblob.nbytes,
                     "kernel_ms_per_rank": per_rank_ms,
                     "clocks_per_rank": [{"sm_mhz": c.get("sm_mhz"), "sm_mhz_min": c.get("sm_mhz_min"), "power_w_max": c.get("power_w_max"),
                                          "reasons": c.get("reasons")} for c in lclocks]}
        # put the headline problem back into rank 0's buffer for the final consistency check
        solver.launch_packed(blob)
        solver.fetch(nseq)
    # ---- strong scaling of ONE shipped input: its tiles dealt over all GPUs, so single sequences are split
    # and reduced with the cross-GPU argmax (atomicMax.sys on the home rank) — the reference's own mode of
    # scaling (the same input3.txt over more ranks, main.c:110-119)
    strong = None
    if strong_problem is not None:
        sblob = plan.plan_pack(strong_problem, world, mode="tile", split_tiles=True, out=solver._pinned_in(1 << 20)) if rank == 0 else None
        s_n = torch.tensor([strong_problem.nseq if rank == 0 else 0], dtype=torch.int64, device="cuda")
        dist.broadcast(s_n, 0)
        solver.launch_packed(sblob)
        sres = solver.fetch(int(s_n.item()))
        sres = sres.copy() if rank == 0 else None
        sms = []
        for _ in range(max(5, min(steps, 20)) + 3):
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            flusher.flush()
            align()
            e0.record()
            fabric.launch(sblob.header if rank == 0 else None)
            e1.record()
            e1.synchronize()
            sms.append(e0.elapsed_time(e1))
        sms = _max_over_ranks(sms[3:])
        solver.check()
        if rank == 0:
            final_s = solver.fetch(int(s_n.item())).copy()
            assert np.array_equal(final_s, sres), "strong-scaling results changed between steps"
            it = sblob.items()
            owners = {}
            for r in range(world):
                for q in it[sblob.header.rank_first_item[r]: sblob.header.rank_first_item[r + 1]]["seq"]:
                    owners.setdefault(int(q), set()).add(r)
            strong = {"kernel_ms": sum(sms) / len(sms), "kernel_ms_min": min(sms), "result": sres, "items": sblob.nitems,
                      "sequences_split_over_gpus": sum(1 for v in owners.values() if len(v) > 1), "nseq": strong_problem.nseq}
        else:
            solver.fetch(0)
        solver.launch_packed(blob)
        solver.fetch(nseq)
    kernel_ms, e2e_ms, pre_ms = _max_over_ranks(kernel_ms), _max_over_ranks(e2e_ms), _max_over_ranks(pre_ms)
    total_launches = torch.tensor([launches], dtype=torch.int64, device="cuda")
    dist.all_reduce(total_launches)
    dist.barrier()
    if rank != 0:
        dist.destroy_process_group()
        return None
    final = solver.fetch(nseq).copy()
    assert np.array_equal(final, res), "results changed between steps"
    out = {
        "kernel_ms": kernel_ms, "e2e_ms": e2e_ms, "e2e_prepacked_ms": pre_ms, "launches": int(total_launches.item()),
        "clocks": clocks, "clocks_all_ranks": all_clocks, "h2d": blob.nbytes, "d2h": nseq * 16, "result": res,
        "padded_cells": int(blob.header.padded_cells), "nitems": blob.nitems,
        "fabric": {"multicast": fabric.use_multicast, "world": world},
        "large": large,
        "strong": strong,
        "large_mode": lblob.mode if large is not None else None,
        "mode": blob.mode,
    }
    dist.destroy_process_group()
    return out
