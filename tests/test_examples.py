"""The C / C++ examples compile against include/seqalign.h and the C one reproduces a golden output."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "mpi_openmp_cuda_b200" / "_lib"


def test_c_example_builds_and_matches_golden(tmp_path, shipped_inputs):
    exe = tmp_path / "solve_cpu"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Iinclude", "examples/solve_cpu.c", f"-L{LIB}", "-lseqalign_core",
                    f"-Wl,-rpath,{LIB}", "-o", str(exe)], cwd=ROOT, check=True)
    path, gold = shipped_inputs[1]
    out = subprocess.run([str(exe)], stdin=open(path, "rb"), capture_output=True, check=True)
    assert out.stdout == gold.read_bytes()


def test_cpp_gpu_example_compiles(tmp_path):
    if not (LIB / "libseqalign_cuda.so").exists():
        pytest.skip("CUDA library not built")
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Iinclude", "-c", "examples/solve_gpu.cpp", "-o", str(tmp_path / "solve_gpu.o")],
                   cwd=ROOT, check=True)
