"""The native `final --gpus N` launcher (csrc/host/native_launch.cpp) without a GPU: real worker processes,
real Unix sockets and SCM_RIGHTS descriptor passing, with the CUDA worker replaced by the built-in stand-in
(SA_NATIVE_SELFTEST=1) that checks every descriptor and vote arrives at the right rank. The reference's
counterpart is `mpiexec -np N` + MPI_Init (makefile:11, main.c:62-64)."""
import os
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def final_bin():
    subprocess.run(["make", "-C", str(ROOT), "core"], check=True, capture_output=True)
    exe = ROOT / "build" / "final_launch_test"
    exe.parent.mkdir(exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-Iinclude", "-o", str(exe), "csrc/host/final_main.cpp", "csrc/host/native_launch.cpp",
                    "-Lmpi_openmp_cuda_b200/_lib", "-lseqalign_core", "-ldl", f"-Wl,-rpath,{ROOT}/mpi_openmp_cuda_b200/_lib"],
                   cwd=ROOT, check=True)
    return exe


def run(exe, args, path, **env):
    return subprocess.run([str(exe), *args], stdin=open(path, "rb"), capture_output=True, timeout=120,
                          env={**os.environ, "SA_NATIVE_SELFTEST": "1", **env})


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8, 16])
def test_descriptor_exchange_and_votes_reach_every_worker(final_bin, shipped_inputs, world):
    for i in (1, 3, 6):
        path, gold = shipped_inputs[i]
        out = run(final_bin, ["--gpus", str(world), "--fabric"], path)
        assert out.returncode == 0, out.stderr.decode()
        assert out.stdout == gold.read_bytes()
        assert out.stderr == b""


def test_a_worker_that_vanishes_fails_the_run_instead_of_hanging_it(final_bin, shipped_inputs):
    out = run(final_bin, ["--gpus", "4"], shipped_inputs[1][0], SA_NATIVE_SELFTEST_DIE="1")
    assert out.returncode == 1 and b"went away" in out.stderr


def test_world_size_is_bounded(final_bin, shipped_inputs):
    out = run(final_bin, ["--gpus", "17"], shipped_inputs[1][0])
    assert out.returncode == 2 and out.stdout == b""


def test_unsupported_input_falls_back_to_the_cpu_oracle(final_bin, tmp_path):
    p = tmp_path / "in.txt"
    p.write_bytes(b"1 2 3 4\nAB!#$%&*CD\n1\nB!#$%&*C\n")  # more than five distinct non-letter bytes
    out = run(final_bin, ["--gpus", "2"], p)
    assert out.returncode == 0 and out.stdout.startswith(b"#0: score:") and b"CPU oracle" in out.stderr


def test_parse_error(final_bin, tmp_path):
    p = tmp_path / "in.txt"
    p.write_bytes(b"1 2 3")
    out = run(final_bin, ["--gpus", "2"], p)
    assert out.returncode == 1 and out.stdout == b""
