#!/usr/bin/env python
"""BASELINE.json configs[4]: broadcast / argmax-reduce sweep, 1 KB - 1 GB, in-kernel fabric vs NCCL.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 benchmarks/fabric_sweep.py

Device-timed (CUDA events), a tiny all-reduce aligns the ranks before every timed op, max over
ranks, best and median of `--iters`. One JSON line per (op, impl, size) on rank 0.
"""
import argparse
import json
import os
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from mpi_openmp_cuda_b200.parallel.fabric import Fabric, init_distributed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--max-bytes", type=int, default=1 << 30)
ap.add_argument("--iters", type=int, default=7)
ap.add_argument("--out", default=None)
ap.add_argument("--only", default="", help="comma-separated subset of: bcast,pull,nccl,argmax (default: all)")
a = ap.parse_args()

only = set(filter(None, a.only.split(",")))
want = lambda name: not only or name in only  # noqa: E731
real = os.dup(1)
os.dup2(2, 1)
rank, world, local = init_distributed("nccl")
fab = Fabric(max_seq=64, max_blob_bytes=a.max_bytes)
dev = fab.device
tick = torch.zeros(1, device=dev)
lines = []


def timed(fn, pre=None):
    ts = []
    for _ in range(a.iters + 2):
        if pre is not None:
            pre()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tick.add_(1)           # keeps the host ahead of the device
        dist.all_reduce(tick)  # device-side alignment of the ranks
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    t = torch.tensor(ts[2:], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    v = t.cpu().tolist()
    return min(v), statistics.median(v)


def emit(op, impl, nbytes, best, med, ok=True):
    if rank == 0:
        row = {"op": op, "impl": impl, "world": world, "bytes": nbytes, "best_us": round(best, 2), "median_us": round(med, 2),
               "GBps_best": round(nbytes / best / 1e3, 2), "ok": ok}
        lines.append(row)
        os.write(real, (json.dumps(row) + "\n").encode())


size = 1024
while size <= a.max_bytes:
    # ---------------- broadcast
    region = fab.blob_region[:size]
    if rank == 0:
        region.copy_(torch.arange(size, device=dev, dtype=torch.int64).mul_(2654435761).bitwise_and_(0xFF).to(torch.uint8))
    else:
        region.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    for mc in (([True, False] if fab.mc_ptr else [False]) if want("bcast") else []):
        fab.use_multicast = mc and world > 1
        fab._desc.use_multicast = 1 if fab.use_multicast else 0
        if rank != 0:
            region.zero_()
        best, med = timed(lambda: fab.bcast(size))
        expect = torch.arange(size, device=dev, dtype=torch.int64).mul_(2654435761).bitwise_and_(0xFF).to(torch.uint8)
        ok = torch.tensor([int(torch.equal(region, expect))], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        emit("bcast", "in-kernel " + ("multimem.st (NVLS)" if mc else "P2P stores"), size, best, med, bool(ok.item()))
    if want("pull"):
        if rank != 0:
            region.zero_()
        best, med = timed(lambda: fab.pull(size))  # the barrier + all-reduce in timed() put rank 0's bytes in place first
        expect = torch.arange(size, device=dev, dtype=torch.int64).mul_(2654435761).bitwise_and_(0xFF).to(torch.uint8)
        ok = torch.tensor([int(torch.equal(region, expect))], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        emit("bcast", "in-kernel pull (LDG over NVLink, every rank its own copy)", size, best, med, bool(ok.item()))
    t = torch.empty(size, dtype=torch.uint8, device=dev)
    if want("nccl"):
        best, med = timed(lambda: dist.broadcast(t, 0))
        emit("bcast", "NCCL broadcast", size, best, med)
    if not want("argmax"):
        del t
        size *= 4
        continue
    # ---------------- argmax reduce (packed u64 keys)
    n = size // 8
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    keys = torch.randint(0, 1 << 62, (n,), dtype=torch.int64, device=dev, generator=g)
    dst = fab.blob_region[:size].view(torch.int64)
    best, med = timed(lambda: fab.argmax(keys), pre=(lambda: dst.zero_()) if rank == 0 else None)
    ref = keys.clone()
    dist.reduce(ref, 0, op=dist.ReduceOp.MAX)
    ok = bool(torch.equal(dst, ref)) if rank == 0 else True
    emit("argmax", "in-kernel atomicMax.sys over NVLink", size, best, med, ok)
    t2 = keys.clone()
    best, med = timed(lambda: dist.reduce(t2, 0, op=dist.ReduceOp.MAX))
    emit("argmax", "NCCL reduce(MAX)", size, best, med)
    del t, t2, keys, ref
    size *= 4

err = fab.error_word()
if rank == 0:
    out = a.out or str(ROOT / "gpurun_out" / f"fabric_sweep_{world}.json")
    Path(out).parent.mkdir(exist_ok=True)
    Path(out).write_text(json.dumps(lines, indent=1))
assert err == 0, f"fabric timeout code {err} on rank {rank}"
dist.barrier()
dist.destroy_process_group()
