"""BASELINE.json configs[0]: multi-process CPU run (gloo, real processes) — plumbing + output diff."""
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.parametrize("world,i", [(2, 1), (2, 6), (3, 4), (2, 2)])
def test_cpu_world_output_is_byte_identical(shipped_inputs, world, i):
    path, gold = shipped_inputs[i]
    out = subprocess.run([sys.executable, "-m", "mpi_openmp_cuda_b200.parallel.launcher", "--cpu", "--gpus", str(world)],
                         stdin=open(path, "rb"), capture_output=True, cwd=ROOT, timeout=300)
    assert out.returncode == 0, out.stderr.decode()[-2000:]
    assert out.stdout == gold.read_bytes()


def test_output_does_not_depend_on_world_size(shipped_inputs):
    path, gold = shipped_inputs[3]
    outs = []
    for world in (1, 4):
        out = subprocess.run([sys.executable, "-m", "mpi_openmp_cuda_b200.parallel.launcher", "--cpu", "--gpus", str(world)],
                             stdin=open(path, "rb"), capture_output=True, cwd=ROOT, timeout=300)
        assert out.returncode == 0, out.stderr.decode()[-2000:]
        outs.append(out.stdout)
    assert outs[0] == outs[1] == gold.read_bytes()
