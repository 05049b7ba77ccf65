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


@pytest.mark.parametrize("nranges", [1, 3, 40])
def test_range_packing_example_reproduces_goldens(tmp_path, shipped_inputs, nranges):
    """examples/pack_ranges.c: sa_plan_pack_range + the CPU walk, range by range (more ranges than sequences included)."""
    exe = tmp_path / "pack_ranges"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Iinclude", "examples/pack_ranges.c", f"-L{LIB}", "-lseqalign_core",
                    f"-Wl,-rpath,{LIB}", "-o", str(exe)], cwd=ROOT, check=True)
    for i in (1, 3, 6):
        path, gold = shipped_inputs[i]
        out = subprocess.run([str(exe), str(nranges)], stdin=open(path, "rb"), capture_output=True, check=True)
        assert out.stdout == gold.read_bytes()
