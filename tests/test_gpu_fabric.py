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


@pytest.mark.parametrize("world", [2, 4, 8])
def test_split_sequence_cross_gpu_argmax(world):
    """One sequence's offset tiles on several GPUs: keys reduced with atomicMax.sys on the home rank,
    finished by the last-arriving tile wherever it ran; ties, an exact substring, two solves in a row."""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    import os

    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                          "--master-addr", "127.0.0.1", "--master-port", str(29740 + world),
                          str(ROOT / "tests" / "_fabric_split_worker.py")],
                         capture_output=True, cwd=ROOT, timeout=600, env=dict(os.environ, PYTHONPATH=str(ROOT)))
    assert out.returncode == 0, out.stderr.decode()[-3000:]
    assert b"split fabric ok" in out.stdout


def test_dead_peer_times_out_instead_of_hanging():
    """Fault injection for SURVEY.md 5.3 / R11: a rank whose peer never broadcasts must come back with
    an error word, not spin forever (the reference's exit(1) leaves its peers inside MPI collectives).
    Needs one GPU only: 'rank 1' is launched alone, with two local buffers standing in for the two
    ranks' symmetric buffers; nobody ever plays rank 0."""
    import ctypes as C

    import torch

    from mpi_openmp_cuda_b200.ops import _native as N
    from mpi_openmp_cuda_b200.parallel.fabric import FabricDesc, FabricLayout

    lib = N.cuda()
    lib.sa_fabric_layout.argtypes = [C.c_uint32, C.c_size_t, C.POINTER(FabricLayout)]
    lib.sa_cuda_fabric_solve.argtypes = [C.POINTER(FabricDesc), C.POINTER(N.BlobHeader), C.c_void_p]
    lay = FabricLayout()
    N.check(lib.sa_fabric_layout(64, 1 << 20, C.byref(lay)))
    bufs = [torch.zeros(int(lay.total_bytes), dtype=torch.uint8, device="cuda:0") for _ in range(2)]
    d = FabricDesc()
    d.rank, d.world, d.epoch, d.use_multicast, d.timeout_ms = 1, 2, 1, 0, 50
    d.peer_base[0], d.peer_base[1] = bufs[0].data_ptr(), bufs[1].data_ptr()
    d.layout = lay
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    N.check(lib.sa_cuda_fabric_solve(C.byref(d), None, None))
    t1.record()
    torch.cuda.synchronize()
    err = int(bufs[1][lay.off_ctrl + 12 : lay.off_ctrl + 16].view(torch.int32).item())
    assert err == 1, "the waiting rank must report the timeout"
    assert 40.0 < t0.elapsed_time(t1) < 2000.0  # bounded by timeout_ms, not by luck


def test_rank0_times_out_when_a_peer_never_delivers(shipped_inputs):
    """The other direction: rank 0 of a world of two whose peer never runs. Rank 0 scores its own share,
    then waits (bounded) for the peer's result rows and must come back with error word 2."""
    import ctypes as C

    import torch

    from mpi_openmp_cuda_b200 import Problem
    from mpi_openmp_cuda_b200.ops import _native as N
    from mpi_openmp_cuda_b200.ops import plan
    from mpi_openmp_cuda_b200.parallel.fabric import FabricDesc, FabricLayout

    lib = N.cuda()
    lib.sa_fabric_layout.argtypes = [C.c_uint32, C.c_size_t, C.POINTER(FabricLayout)]
    lib.sa_cuda_fabric_solve.argtypes = [C.POINTER(FabricDesc), C.POINTER(N.BlobHeader), C.c_void_p]
    blob = plan.plan_pack(Problem.from_file(shipped_inputs[3][0]), 2)
    lay = FabricLayout()
    N.check(lib.sa_fabric_layout(64, 1 << 20, C.byref(lay)))
    bufs = [torch.zeros(int(lay.total_bytes), dtype=torch.uint8, device="cuda:0") for _ in range(2)]
    bufs[0][lay.off_blob : lay.off_blob + blob.nbytes].copy_(torch.from_numpy(blob.numpy()[: blob.nbytes]))
    d = FabricDesc()
    d.rank, d.world, d.epoch, d.use_multicast, d.timeout_ms = 0, 2, 1, 0, 50
    d.peer_base[0], d.peer_base[1] = bufs[0].data_ptr(), bufs[1].data_ptr()
    d.layout = lay
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    N.check(lib.sa_cuda_fabric_solve(C.byref(d), C.byref(blob.header), None))
    t1.record()
    torch.cuda.synchronize()
    err = int(bufs[0][lay.off_ctrl + 12 : lay.off_ctrl + 16].view(torch.int32).item())
    assert err == 2, "rank 0 must report that rows never arrived"
    assert 40.0 < t0.elapsed_time(t1) < 2000.0
    # the header packet did reach the (absent) peer's buffer: epoch in the upper half of every word
    pkt = bufs[1][lay.off_hdr_ll : lay.off_hdr_ll + 16].view(torch.int64).cpu().tolist()
    assert pkt[0] >> 32 == 1 and (pkt[0] & 0xFFFFFFFF) == 0x31424153
