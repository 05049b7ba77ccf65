"""`final` keeps the reference's stdin->stdout contract (main.c:46-244, makefile:11)."""
import json
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def final_bin():
    exe = ROOT / "final"
    if not exe.exists():
        subprocess.run(["make", "-C", str(ROOT), "core"], check=True, capture_output=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-Iinclude", "-o", "final", "csrc/host/final_main.cpp", "csrc/host/native_launch.cpp",
                        "-Lmpi_openmp_cuda_b200/_lib", "-lseqalign_core", "-ldl",
                        "-Wl,-rpath,$ORIGIN/mpi_openmp_cuda_b200/_lib"], cwd=ROOT, check=True)
    return exe


@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_cpu_oracle_mode_is_byte_identical(final_bin, shipped_inputs, i):
    path, gold = shipped_inputs[i]
    out = subprocess.run([str(final_bin), "--cpu-oracle"], stdin=open(path, "rb"), capture_output=True, check=True)
    assert out.stdout == gold.read_bytes()
    assert out.stderr == b""


def test_timing_json_goes_to_side_file(final_bin, shipped_inputs, tmp_path):
    path, gold = shipped_inputs[1]
    tj = tmp_path / "t.json"
    out = subprocess.run([str(final_bin), "--cpu-oracle", "--timing-json", str(tj)], stdin=open(path, "rb"),
                         capture_output=True, check=True)
    assert out.stdout == gold.read_bytes()  # stdout stays byte-identical
    info = json.loads(tj.read_text())
    assert info["nseq"] == 10 and info["candidates"] == 5006


def test_parse_error_is_nonzero_exit(final_bin):
    out = subprocess.run([str(final_bin), "--cpu-oracle"], input=b"1 2 3 4 ABC 3 AB", capture_output=True)
    assert out.returncode != 0 and out.stdout == b"" and b"expected 3 sequences" in out.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_gpu_mode_is_byte_identical(final_bin, shipped_inputs, i):
    path, gold = shipped_inputs[i]
    out = subprocess.run([str(final_bin)], stdin=open(path, "rb"), capture_output=True, check=True)
    assert out.stdout == gold.read_bytes()


@pytest.mark.gpu
def test_gpu_mode_across_two_gpus_is_byte_identical(final_bin, shipped_inputs):
    """`./final --gpus 2` = the reference's `mpiexec -np 2 ./final` (makefile:11): same bytes on stdout."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    for i in (1, 3):
        path, gold = shipped_inputs[i]
        for extra in ([], ["--python-launcher"]):  # native workers (default) and the torch-bootstrapped ranks
            out = subprocess.run([str(final_bin), "--gpus", "2", *extra], stdin=open(path, "rb"), capture_output=True, timeout=600)
            assert out.returncode == 0, out.stderr.decode()[-2000:]
            assert out.stdout == gold.read_bytes()


@pytest.mark.gpu
def test_native_fabric_path_on_one_gpu(final_bin, shipped_inputs, tmp_path):
    """--fabric forces the worker-process path (CUDA VMM allocation, descriptor exchange through the launcher,
    fabric kernel) even for one GPU, so the bring-up code is exercised on a single-GPU box too."""
    tj = tmp_path / "t.json"
    for i in (1, 2, 3, 4, 5, 6):
        path, gold = shipped_inputs[i]
        out = subprocess.run([str(final_bin), "--gpus", "1", "--fabric", "--timing-json", str(tj)], stdin=open(path, "rb"),
                             capture_output=True, timeout=300)
        assert out.returncode == 0, out.stderr.decode()[-2000:]
        assert out.stdout == gold.read_bytes()
    info = json.loads(tj.read_text())
    assert info["engine"] == "gpu-fabric-native" and info["gpus"] == 1 and info["kernel_ms"] > 0


@pytest.mark.gpu
def test_native_launcher_all_gpus_of_the_box(final_bin, shipped_inputs):
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    n = 8 if n >= 8 else 4 if n >= 4 else 2
    for i in (3, 4, 6):
        path, gold = shipped_inputs[i]
        for env_extra in ({}, {"SA_FABRIC_MULTICAST": "0"}):
            import os

            out = subprocess.run([str(final_bin), "--gpus", str(n)], stdin=open(path, "rb"), capture_output=True, timeout=600,
                                 env={**os.environ, **env_extra})
            assert out.returncode == 0, out.stderr.decode()[-2000:]
            assert out.stdout == gold.read_bytes()


def test_zero_sequences_prints_nothing(final_bin):
    out = subprocess.run([str(final_bin), "--cpu-oracle"], input=b"1 2 3 4 ABC 0\n", capture_output=True)
    assert out.returncode == 0 and out.stdout == b""


def test_python_module_entry_point(shipped_inputs):
    import sys

    path, gold = shipped_inputs[6]
    out = subprocess.run([sys.executable, "-m", "mpi_openmp_cuda_b200", "--cpu"], stdin=open(path, "rb"), capture_output=True,
                         cwd=ROOT, timeout=300)
    assert out.returncode == 0, out.stderr.decode()[-1000:]
    assert out.stdout == gold.read_bytes()
    bad = subprocess.run([sys.executable, "-m", "mpi_openmp_cuda_b200", "--cpu"], input=b"1 2 3", capture_output=True, cwd=ROOT, timeout=300)
    assert bad.returncode == 1 and bad.stdout == b""
