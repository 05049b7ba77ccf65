"""Multi-GPU fabric path: in-kernel broadcast + argmax reduce over NVLink (needs >= 2 GPUs for the
cross-GPU cases; the world=1 case runs on any GPU box)."""
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.gpu


def _ngpu():
    import torch

    return torch.cuda.device_count()


def _run(world, path, env_extra=None):
    import os

    env = dict(os.environ, **(env_extra or {}))
    return subprocess.run([sys.executable, "-m", "mpi_openmp_cuda_b200.parallel.launcher", "--gpus", str(world)],
                          stdin=open(path, "rb"), capture_output=True, cwd=ROOT, timeout=600, env=env)


@pytest.mark.parametrize("i", [1, 3])
def test_world_of_one_uses_the_fabric_kernel(shipped_inputs, i):
    path, gold = shipped_inputs[i]
    out = _run(1, path)
    assert out.returncode == 0, out.stderr.decode()[-2000:]
    assert out.stdout == gold.read_bytes()


@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_two_gpus_byte_identical(shipped_inputs, i):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    path, gold = shipped_inputs[i]
    out = _run(2, path)
    assert out.returncode == 0, out.stderr.decode()[-2000:]
    assert out.stdout == gold.read_bytes()


@pytest.mark.parametrize("mc", ["0", "1"])
def test_two_gpus_p2p_and_multicast_broadcast(shipped_inputs, mc):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    path, gold = shipped_inputs[4]
    out = _run(2, path, {"SA_FABRIC_MULTICAST": mc})
    assert out.returncode == 0, out.stderr.decode()[-2000:]
    assert out.stdout == gold.read_bytes()


@pytest.mark.parametrize("world", [4, 8])
def test_more_gpus_byte_identical(shipped_inputs, world):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    for i in (3, 4, 6):
        path, gold = shipped_inputs[i]
        out = _run(world, path)
        assert out.returncode == 0, out.stderr.decode()[-2000:]
        assert out.stdout == gold.read_bytes()


def test_two_gpus_stream_mode_through_the_fabric():
    """Large batch: the fabric kernel runs warp-autonomous STREAM items (whole pairs per rank)."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    import os

    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29733", str(ROOT / "tests" / "_fabric_stream_worker.py")],
                         capture_output=True, cwd=ROOT, timeout=600, env=dict(os.environ, PYTHONPATH=str(ROOT)))
    assert out.returncode == 0, out.stderr.decode()[-3000:]
    assert b"stream fabric ok" in out.stdout
