#!/usr/bin/env python
"""Run / time the UNMODIFIED reference build (baseline/_ref, see build_ref.sh).

The reference has no timers and no API other than `mpiexec -np P ./final < input`
(makefile:11), so a "step" is one whole run of that command and the time is the wall clock
around it (process start + CUDA context + its per-row malloc/launch/sync loop). The fixed
start-up part is measured separately on the 9-candidate input5.txt and reported next to the
headline so it can be subtracted.

Used by `bench.py --impl reference` and standalone:
    python baseline/run_ref.py --inputs 1 2 3 4 5 6 --np 1 --repeat 3
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
REF = HERE / "_ref"
sys.path.insert(0, str(ROOT))


def ensure_built() -> str | None:
    """Returns None when usable, else a one-line reason."""
    need = [REF / "bin" / "final", REF / "bin" / "mpiexec", REF / "shim" / "libmpishim.so"]
    if all(p.exists() for p in need):
        return None
    if not Path("/root/reference").exists() and not (REF / "src" / "main.c").exists():
        return "baseline/_ref is not built and /root/reference is not present on this machine"
    proc = subprocess.run(["bash", str(HERE / "build_ref.sh")], capture_output=True, text=True)
    if proc.returncode != 0 or not all(p.exists() for p in need):
        return "building the reference failed: " + (proc.stderr.strip().splitlines() or ["?"])[-1]
    return None


def run_once(input_bytes: bytes, np_: int, variant: str = "final", timeout: float = 3600.0):
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = "1"  # the reference's parallel stdin loop scrambles rows otherwise (main.c:96-108)
    env["SHIM_GPUS"] = ",".join(str(i) for i in range(np_))  # it always opens device 0 (cudaFunctions.cu:209)
    env.pop("CUDA_VISIBLE_DEVICES", None)
    cmd = [str(REF / "bin" / "mpiexec"), "-np", str(np_), str(REF / "bin" / variant)]
    t0 = time.perf_counter()
    proc = subprocess.run(cmd, input=input_bytes, capture_output=True, env=env, timeout=timeout)
    dt = time.perf_counter() - t0
    return dt, proc.returncode, proc.stdout, proc.stderr


def expected_output(input_bytes: bytes) -> bytes:
    from mpi_openmp_cuda_b200 import Problem, format_results
    from mpi_openmp_cuda_b200.ops.oracle import solve_cpu

    return format_results(solve_cpu(Problem.parse(input_bytes), "fast"))


def diff_rows(got: bytes, want: bytes) -> dict:
    g, w = got.decode("latin-1").splitlines(), want.decode("latin-1").splitlines()
    bad = [i for i in range(max(len(g), len(w))) if i >= len(g) or i >= len(w) or g[i] != w[i]]
    return {"rows": len(w), "rows_matching_contract": len(w) - len(bad), "first_mismatches": bad[:8]}


def bench_reference(args, rank: int, world: int):
    """`bench.py --impl reference`: same input, same metric, the reference's own code path."""
    if rank != 0:
        return None
    why = ensure_built()
    if why:
        return {"impl": "reference", "unavailable": why}
    from bench import build_problem  # same config as our arm

    n_gpus = args.gpus
    problem = build_problem(args, n_gpus)
    cands = problem.candidates()
    data = problem.serialize() if args.scale * n_gpus > 1 else (ROOT / "data" / args.input).read_bytes()
    variant = os.environ.get("REF_VARIANT", "final")
    budget_s = float(os.environ.get("REF_BUDGET_S", "300"))
    floor_input = (ROOT / "data" / "input5.txt").read_bytes()
    try:
        t_first, rc, out, err = run_once(data, n_gpus, variant)
    except subprocess.TimeoutExpired:
        return {"impl": "reference", "unavailable": "the reference did not finish one run of this input in 3600 s"}
    if rc != 0:
        return {"impl": "reference", "unavailable": f"reference exited {rc} at np={n_gpus}: {err[-200:].decode('latin-1', 'replace')!r}"}
    check = diff_rows(out, expected_output(data))
    steps, warmup = args.steps, args.warmup
    if (steps + warmup) * t_first > budget_s:  # keep the arm inside the box's time budget; say so
        warmup = 1 if t_first * (steps + 1) <= budget_s else 0
        steps = max(1, min(steps, int(budget_s / max(t_first, 1e-3)) - warmup))
    for _ in range(max(0, warmup - 1)):  # the first run above already was one warm-up (JIT cache, page cache)
        run_once(data, n_gpus, variant)
    times = [run_once(data, n_gpus, variant)[0] for _ in range(steps)]
    floor = min(run_once(floor_input, n_gpus, variant)[0] for _ in range(3))
    ms = statistics.fmean(times) * 1e3
    return {
        "impl": "reference",
        "metric": "candidates/s (one full solve of the input; wall clock of the whole `mpiexec -np N ./final < input` run)",
        "value": cands / (ms * 1e-3), "unit": "candidates/s", "n_gpus": n_gpus,
        "steps": steps, "warmup": warmup, "steps_requested": args.steps, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32", "data": f"data/{args.input} x{args.scale * n_gpus}",
        "config": {"model": "exhaustive (n,k) alignment search", "input": args.input, "scale_per_gpu": args.scale,
                   "nseq": problem.nseq, "candidates": cands, "parallelism": f"mpi{n_gpus} (stand-in MPI, one rank per GPU)",
                   "variant": variant + (" (makefile flags: sm_52 SASS + PTX JIT)" if variant == "final" else " (-arch=sm_100)"),
                   "OMP_NUM_THREADS": 1},
        "first_run_ms": t_first * 1e3, "startup_floor_ms": floor * 1e3,
        "ms_per_step_minus_startup": ms - floor * 1e3,
        "output_check": check,
        "timing": "wall clock, whole process (the reference has no timers, SURVEY.md 5.1)",
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--inputs", type=int, nargs="+", default=[1, 2, 3, 4, 5, 6])
    ap.add_argument("--np", type=int, nargs="+", default=[1])
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--variants", nargs="+", default=["final", "final_sm100"])
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "ref_times.json"))
    a = ap.parse_args()
    why = ensure_built()
    if why:
        print(json.dumps({"unavailable": why}))
        return 0
    rows = []
    for variant in a.variants:
        for np_ in a.np:
            for i in a.inputs:
                data = (ROOT / "data" / f"input{i}.txt").read_bytes()
                want = expected_output(data)
                ts, rc, out, err = [], 0, b"", b""
                for _ in range(a.repeat):
                    try:
                        dt, rc, out, err = run_once(data, np_, variant, timeout=1800)
                    except subprocess.TimeoutExpired:
                        rc, dt = -9, 1800.0
                    ts.append(dt)
                    if rc != 0:
                        break
                row = {"variant": variant, "np": np_, "input": i, "rc": rc, "times_s": ts, "best_s": min(ts),
                       **diff_rows(out, want), "stderr_tail": err[-160:].decode("latin-1", "replace")}
                rows.append(row)
                print(json.dumps(row), flush=True)
                Path(a.out).parent.mkdir(parents=True, exist_ok=True)
                Path(a.out).write_text(json.dumps(rows, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
