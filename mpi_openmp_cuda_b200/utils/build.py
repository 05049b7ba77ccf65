"""In-tree native build (reference: makefile:1-4 — three compiler invocations by hand).

`build_all()` runs the repo Makefile, which compiles
  * libseqalign_core.so  (g++ -O3; parser, pair table, CPU oracle, planner/packer)
  * libseqalign_cuda.so  (nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo)
  * ./final              (CLI host with the reference's stdin->stdout contract)
into the source tree so that the artefacts travel with a repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
from pathlib import Path

REPO_ROOT = Path(__file__).resolve().parents[2]
LIB_DIR = REPO_ROOT / "mpi_openmp_cuda_b200" / "_lib"


def lib_path(name: str) -> Path:
    return LIB_DIR / f"lib{name}.so"


def build_all(targets: tuple[str, ...] = ("build",), verbose: bool = False) -> None:
    """Run `make <targets>` at the repo root; raises on failure."""
    env = dict(os.environ)
    cuda_home = env.get("CUDA_HOME", "/usr/local/cuda")
    env["PATH"] = f"{cuda_home}/bin:" + env.get("PATH", "")
    proc = subprocess.run(["make", "-C", str(REPO_ROOT), *targets], env=env, capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout)
        print(proc.stderr)
    if proc.returncode != 0:
        raise RuntimeError(f"native build failed (make {' '.join(targets)}): exit {proc.returncode}")


def ensure_core() -> Path:
    p = lib_path("seqalign_core")
    if not p.exists():
        build_all(("core",))
    return p


def ensure_cuda() -> Path:
    p = lib_path("seqalign_cuda")
    if not p.exists():
        build_all(("cuda",))
    return p
