"""`python -m mpi_openmp_cuda_b200.parallel.launcher --gpus N < input` — the multi-GPU (or, with
--cpu, multi-process CPU) form of `./final`: same stdin, same stdout bytes.

Analogue of `mpiexec -np N ./final` (makefile:11): spawns one worker process per GPU on this
node (127.0.0.1 rendezvous), feeds stdin to rank 0 only, prints rank 0's result lines.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time


def worker(cpu: bool, timing_path: str | None) -> int:
    import torch.distributed as dist

    from .. import Problem, format_results
    from .fabric import init_distributed

    # stdout carries the result lines and nothing else (byte parity with the reference), but
    # libraries such as NCCL may print banners to fd 1: park the real stdout, point fd 1 at stderr.
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    t0 = time.perf_counter()
    rank, world, local = init_distributed("gloo" if cpu else "nccl")
    problem = Problem.parse(sys.stdin.buffer.read()) if rank == 0 else None
    t1 = time.perf_counter()
    if cpu:
        from .cpu_world import solve_cpu_world

        res = solve_cpu_world(problem)
    else:
        from .dist_solver import DistributedSolver
        from .fabric import Fabric

        solver = DistributedSolver(Fabric())
        res = solver.solve(problem)
    t2 = time.perf_counter()
    if rank == 0:
        data = format_results(res)
        while data:
            data = data[os.write(real_stdout, data):]
        if timing_path:
            with open(timing_path, "w") as f:
                json.dump({"engine": "cpu-world" if cpu else "gpu-fabric", "world": world, "nseq": problem.nseq,
                           "candidates": problem.candidates(), "bootstrap_parse_s": t1 - t0, "solve_s": t2 - t1}, f)
    dist.barrier()
    dist.destroy_process_group()
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1, help="number of worker processes (one per GPU)")
    ap.add_argument("--cpu", action="store_true", help="CPU world: gloo + the CPU walk of the kernel logic")
    ap.add_argument("--timing-json", default=None)
    ap.add_argument("--worker", action="store_true", help=argparse.SUPPRESS)
    a = ap.parse_args()
    if a.worker:
        return worker(a.cpu, a.timing_json)
    data = sys.stdin.buffer.read()  # held by the parent so that a failed rendezvous can be retried
    rc = 1
    for attempt in range(2):
        # rendezvous through a fresh file store: no TCP port that another job could already hold
        rdzv_dir = tempfile.mkdtemp(prefix="sa_rdzv_")
        procs = []
        for r in range(a.gpus):
            env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(a.gpus), LOCAL_RANK=str(r),
                       SA_RDZV_FILE=os.path.join(rdzv_dir, "store"))
            env.pop("MASTER_PORT", None)
            cmd = [sys.executable, "-m", "mpi_openmp_cuda_b200.parallel.launcher", "--worker", "--gpus", str(a.gpus)]
            if a.cpu:
                cmd.append("--cpu")
            if a.timing_json:
                cmd += ["--timing-json", a.timing_json]
            procs.append(subprocess.Popen(cmd, env=env, stdin=subprocess.PIPE if r == 0 else subprocess.DEVNULL,
                                          stdout=subprocess.PIPE if r == 0 else subprocess.DEVNULL,
                                          stderr=subprocess.PIPE))
        out0, err0 = procs[0].communicate(data)  # stdin goes to rank 0 only (main.c:68-108)
        rc = procs[0].returncode
        errs = [err0]
        for q in procs[1:]:
            if rc != 0 and q.poll() is None:
                q.kill()  # a failed rank must not leave the others waiting
            _, e = q.communicate()
            errs.append(e)
            if q.returncode != 0 and rc == 0:
                rc = q.returncode
        if rc == 0:
            sys.stdout.buffer.write(out0)
            sys.stdout.buffer.flush()
            for e in errs:
                sys.stderr.buffer.write(e)
            return 0
        if not any(b"EADDRINUSE" in e or b"address already in use" in e for e in errs):
            break  # a real failure, not a rendezvous port collision
    for e in errs:
        sys.stderr.buffer.write(e)
    return rc


if __name__ == "__main__":
    sys.exit(main())
