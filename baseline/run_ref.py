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
import shutil
import statistics
import subprocess
import sys
import tempfile
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


def run_once(input_bytes: bytes, np_: int, variant: str = "final", timeout: float = 3600.0, probe: bool = True):
    """One whole `mpiexec -np N ./final < input` run. Returns (wall_s, rc, stdout, stderr, device) where
    `device` is the CUPTI probe's per-rank record list (baseline/cupti_inject) or None."""
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = "1"  # the reference's parallel stdin loop scrambles rows otherwise (main.c:96-108)
    env["SHIM_GPUS"] = ",".join(str(i) for i in range(np_))  # it always opens device 0 (cudaFunctions.cu:209)
    env.pop("CUDA_VISIBLE_DEVICES", None)
    inject = REF / "shim" / "libcuptiinject.so"
    tmpdir = None
    if probe and inject.exists():
        tmpdir = tempfile.mkdtemp(prefix="refcupti_")
        env["CUDA_INJECTION64_PATH"] = str(inject)
        env["REF_CUPTI_OUT"] = os.path.join(tmpdir, "rank")
    else:
        env.pop("CUDA_INJECTION64_PATH", None)
    cmd = [str(REF / "bin" / "mpiexec"), "-np", str(np_), str(REF / "bin" / variant)]
    t0 = time.perf_counter()
    proc = subprocess.run(cmd, input=input_bytes, capture_output=True, env=env, timeout=timeout)
    dt = time.perf_counter() - t0
    device = None
    if tmpdir:
        recs = []
        for f in sorted(os.listdir(tmpdir)):
            try:
                recs.append(json.loads(open(os.path.join(tmpdir, f)).read()))
            except Exception:
                pass
        shutil.rmtree(tmpdir, ignore_errors=True)
        recs = [r for r in recs if r.get("kernels", 0) > 0]
        device = recs or None
    return dt, proc.returncode, proc.stdout, proc.stderr, device


def device_ms(device) -> float | None:
    """Reference device time of one run: per rank the sum of its kernels' GPU durations (they are
    serialised by a cudaDeviceSynchronize each, cudaFunctions.cu:218), then the max over ranks."""
    if not device:
        return None
    return max(r["kernel_ns"] for r in device) * 1e-6


def expected_output(input_bytes: bytes) -> bytes:
    """The section 8.2 contract, restated in numpy (baseline/refgen.py): nothing of the repo's native code."""
    from baseline import refgen

    return refgen.format_results(refgen.solve(*refgen.parse_text(input_bytes)))


def diff_rows(got: bytes, want: bytes) -> dict:
    g, w = got.decode("latin-1").splitlines(), want.decode("latin-1").splitlines()
    bad = [i for i in range(max(len(g), len(w))) if i >= len(g) or i >= len(w) or g[i] != w[i]]
    return {"rows": len(w), "rows_matching_contract": len(w) - len(bad), "first_mismatches": bad[:8]}


def bench_reference(args, rank: int, world: int):
    """`bench.py --impl reference`: same input bytes, same metric, the reference's own code path
    (`mpiexec -np N ./final < input`, makefile:11). Loads no native code of this repository."""
    if rank != 0:
        return None
    why = ensure_built()
    if why:
        return {"impl": "reference", "unavailable": why}
    from baseline import refgen
    from baseline.contract import METRIC, MODEL, UNIT

    n_gpus = args.gpus
    total_scale = args.scale * n_gpus
    data = refgen.scaled_text(ROOT / "data" / args.input, total_scale, seed=1)
    weights, seq1, seqs = refgen.parse_text(data)
    cands = refgen.candidates(seq1, seqs)
    variant = os.environ.get("REF_VARIANT", "final")
    budget_s = float(os.environ.get("REF_BUDGET_S", "800"))
    floor_input = (ROOT / "data" / "input5.txt").read_bytes()
    t_begin = time.perf_counter()
    try:
        t_first, rc, out, err, dev_first = run_once(data, n_gpus, variant)
    except subprocess.TimeoutExpired:
        return {"impl": "reference", "unavailable": "the reference did not finish one run of this input in 3600 s"}
    if rc != 0:
        return {"impl": "reference", "unavailable": f"reference exited {rc} at np={n_gpus}: {err[-200:].decode('latin-1', 'replace')!r}"}
    check = diff_rows(out, expected_output(data))
    steps, warmup = args.steps, args.warmup
    if (steps + warmup) * t_first > budget_s:  # keep the arm inside the box's time budget; say so
        warmup = min(warmup, 1)
        steps = max(1, min(steps, int(budget_s / max(t_first, 1e-3)) - warmup))
    for _ in range(max(0, warmup - 1)):  # the first run above already was one warm-up (JIT cache, page cache)
        run_once(data, n_gpus, variant)
    runs = [run_once(data, n_gpus, variant) for _ in range(steps)]
    wall_ms = statistics.fmean(r[0] for r in runs) * 1e3
    dev = [device_ms(r[4]) for r in runs]
    have_dev = all(d is not None for d in dev)
    # start-up floor (process + CUDA context + JIT-cache lookup) on the 9-candidate input5.txt
    floor = run_once(floor_input, n_gpus, variant)[0]
    floor_plain = run_once(floor_input, n_gpus, variant, probe=False)[0]
    if have_dev:
        ms = statistics.fmean(dev)
        timing = ("GPU timestamps of the reference's own kernels (CUPTI activity records through CUDA_INJECTION64_PATH; "
                  "no replay, no source change), summed per rank, max over ranks")
    else:
        ms = wall_ms - floor * 1e3
        timing = "wall clock of the whole run minus the start-up floor measured on input5.txt (the CUPTI probe gave no records)"
    last = runs[-1][4] or []
    h2d = sum(r.get("h2d_bytes", 0) for r in last)
    d2h = sum(r.get("d2h_bytes", 0) for r in last)
    return {
        "impl": "reference",
        "metric": METRIC,
        "value": cands / (ms * 1e-3), "unit": UNIT, "n_gpus": n_gpus,
        "steps": steps, "warmup": warmup, "steps_requested": args.steps, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32", "data": f"shipped reference input data/{args.input}" + ("" if total_scale == 1 else f" replicated x{total_scale} with random substitutions (synthetic)"),
        "config": {"model": MODEL, "input": args.input, "scale_per_gpu": args.scale,
                   "nseq": len(seqs), "len1": len(seq1), "candidates": cands, "global_batch": len(seqs), "seq_len": len(seq1),
                   "parallelism": f"dp{n_gpus}",
                   "launcher": f"mpiexec -np {n_gpus} ./final < input (stand-in MPI, one rank per GPU)",
                   "variant": variant + (" (makefile flags: sm_52 SASS + PTX JIT)" if variant == "final" else " (-arch=sm_100)"),
                   "OMP_NUM_THREADS": 1},
        "timing": timing,
        "e2e": {"value": cands / (wall_ms * 1e-3), "unit": UNIT, "ms_per_step": wall_ms,
                "what": "wall clock of the whole `mpiexec -np N ./final < input` run: the only interface the reference has "
                        "(process start, CUDA context, stdin parse, H2D, its kernels, D2H, print)",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "startup_floor_ms": floor * 1e3, "ms_per_step_minus_startup": wall_ms - floor * 1e3},
        "gpu_launches": sum(r.get("kernels", 0) for r in last) * steps if last else None,
        "kernels_per_step": sum(r.get("kernels", 0) for r in last) if last else None,
        "device_ms_per_rank_last_step": [round(r["kernel_ns"] * 1e-6, 3) for r in last],
        "first_run_ms": t_first * 1e3, "startup_floor_ms": floor * 1e3,
        "probe_overhead_ms": (floor - floor_plain) * 1e3,
        "output_check": check,
        "arm_wall_s": time.perf_counter() - t_begin,
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
                ts, rc, out, err, dev = [], 0, b"", b"", None
                for _ in range(a.repeat):
                    try:
                        dt, rc, out, err, dev = run_once(data, np_, variant, timeout=1800)
                    except subprocess.TimeoutExpired:
                        rc, dt = -9, 1800.0
                    ts.append(dt)
                    if rc != 0:
                        break
                row = {"variant": variant, "np": np_, "input": i, "rc": rc, "times_s": ts, "best_s": min(ts),
                       "device_ms": device_ms(dev), "kernels": sum(r.get("kernels", 0) for r in dev) if dev else None,
                       **diff_rows(out, want), "stderr_tail": err[-160:].decode("latin-1", "replace")}
                rows.append(row)
                print(json.dumps(row), flush=True)
                Path(a.out).parent.mkdir(parents=True, exist_ok=True)
                Path(a.out).write_text(json.dumps(rows, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
