"""Allocation-free entry point `sa_cuda_solve_blob_noalloc`, driven through the C ABI.

Some callers of this library cannot allocate device memory at the moment they need a solve:
a plug-in running inside a server that already owns all of HBM, or a custom op that must not go
around its framework's caching allocator. For them the library carries its scratch inside its
own image. This test plays such a caller: the blob and the result table are buffers it already
owns (torch tensors created up front stand in for that), and the solve itself must work
without any further allocation and must give exactly what the ordinary entry point gives —
on a first call, and again on a second call that meets the array as the first one left it.

The file name sorts last on purpose: with `pytest -x` every other test has already run if
this one fails. It does not depend on them and also passes alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

from mpi_openmp_cuda_b200 import Problem
from mpi_openmp_cuda_b200.models.synthetic import random_problem
from mpi_openmp_cuda_b200.ops import _native as N
from mpi_openmp_cuda_b200.ops import plan
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu

pytestmark = pytest.mark.gpu


def _device_buffers(blob):
    """What the caller owns before the solve: blob copy, result table, and (for the ordinary
    entry point only) a zeroed scratch area."""
    import torch

    lib = N.cuda()
    d_blob = torch.from_numpy(blob.numpy()[: blob.nbytes].copy()).to("cuda:0")
    d_out = torch.full((blob.nseq, 4), -7, dtype=torch.int32, device="cuda:0")
    d_scratch = torch.zeros(int(lib.sa_cuda_scratch_bytes(blob.nseq)), dtype=torch.uint8, device="cuda:0")
    return d_blob, d_out, d_scratch


def _solve_regular(blob, bufs):
    import torch

    d_blob, d_out, d_scratch = bufs
    d_out.fill_(-7)
    with torch.cuda.device(0):
        stream = torch.cuda.current_stream().cuda_stream
        N.check(N.cuda().sa_cuda_solve_blob(C.c_void_p(d_blob.data_ptr()), C.byref(blob.header),
                                             C.c_void_p(d_out.data_ptr()), C.c_void_p(d_scratch.data_ptr()),
                                             C.c_void_p(stream)))
        torch.cuda.synchronize()
    return d_out.cpu().numpy()[:, :3].copy()


def _solve_noalloc(blob, bufs):
    import torch

    d_blob, d_out, _ = bufs
    d_out.fill_(-7)
    with torch.cuda.device(0):
        stream = torch.cuda.current_stream().cuda_stream
        N.check(N.cuda().sa_cuda_solve_blob_noalloc(C.c_void_p(d_blob.data_ptr()), C.byref(blob.header),
                                                     C.c_void_p(d_out.data_ptr()), C.c_void_p(stream)))
        torch.cuda.synchronize()
    return d_out.cpu().numpy()[:, :3].copy()


def test_noalloc_matches_regular_entry_point_twice(shipped_inputs):
    p = Problem.from_file(shipped_inputs[3][0])
    blob = plan.plan_pack(p)
    bufs = _device_buffers(blob)
    expect = _solve_regular(blob, bufs)
    assert np.array_equal(expect, solve_cpu(p, "fast"))
    first = _solve_noalloc(blob, bufs)
    second = _solve_noalloc(blob, bufs)  # meets the built-in array exactly as call one left it
    # exact equality: integer scores, the same tolerance (none) every other GPU test uses
    assert np.array_equal(first, expect)
    assert np.array_equal(second, expect)


def test_noalloc_allocates_no_device_memory(shipped_inputs):
    """The reason the entry point exists: between the moment the caller owns its buffers and the
    moment the rows are back, the amount of free device memory must not move — neither cudaMalloc
    nor torch's allocator may have been asked for a byte."""
    import torch

    p = Problem.from_file(shipped_inputs[4][0])
    blob = plan.plan_pack(p)
    bufs = _device_buffers(blob)
    _solve_noalloc(blob, bufs)  # first call on this device: resolves the array's address (no allocation either)
    host_rows = torch.empty((blob.nseq, 4), dtype=torch.int32).pin_memory()
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(0)
    reserved0 = torch.cuda.memory_reserved(0)
    d_blob, d_out, _ = bufs
    with torch.cuda.device(0):
        stream = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            N.check(N.cuda().sa_cuda_solve_blob_noalloc(C.c_void_p(d_blob.data_ptr()), C.byref(blob.header),
                                                         C.c_void_p(d_out.data_ptr()), C.c_void_p(stream)))
        host_rows.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(0)
    assert free1 == free0 and torch.cuda.memory_reserved(0) == reserved0
    assert np.array_equal(host_rows.numpy()[:, :3], solve_cpu(p, "fast"))


def test_noalloc_stream_mode_and_a_dirty_array():
    """The warp-autonomous STREAM kernel through the same entry point (forced: AUTO picks it only for
    much larger batches), and a call that finds the built-in array full of another caller's leftovers —
    what a launch torn down half-way would leave behind. Every call clears what it uses first."""
    import torch

    p = random_problem(900, len1=500, len2_range=(5, 420), seed=24)
    blob = plan.plan_pack(p, mode="stream")
    assert blob.mode == "stream" and blob.nseq <= N.SA_NOALLOC_MAX_SEQ
    bufs = _device_buffers(blob)
    expect = _solve_regular(blob, bufs)
    assert np.array_equal(expect, solve_cpu(p, "fast"))
    assert np.array_equal(_solve_noalloc(blob, bufs), expect)
    # dirty the array through the ordinary entry point's eyes: run a TILE problem whose scratch layout
    # differs, then poison nothing ourselves (the array is private) but check that a different layout
    # followed by the stream blob still gives the right rows twice in a row
    other = plan.plan_pack(random_problem(33, len1=700, len2_range=(100, 600), seed=25))
    ob = _device_buffers(other)
    assert np.array_equal(_solve_noalloc(other, ob), _solve_regular(other, ob))
    assert np.array_equal(_solve_noalloc(blob, bufs), expect)
    assert np.array_equal(_solve_noalloc(blob, bufs), expect)
    torch.cuda.synchronize()


def test_noalloc_rejects_a_header_that_is_not_a_blob(shipped_inputs):
    blob = plan.plan_pack(Problem.from_file(shipped_inputs[1][0]))
    bufs = _device_buffers(blob)
    bad = N.BlobHeader.from_buffer_copy(bytes(blob.header))
    bad.magic = 0x12345678
    bad.nseq = N.SA_NOALLOC_MAX_SEQ + 5  # must still be reported as "not a blob", not as "too large"
    rc = N.cuda().sa_cuda_solve_blob_noalloc(C.c_void_p(bufs[0].data_ptr()), C.byref(bad), C.c_void_p(bufs[1].data_ptr()), None)
    assert rc == N.SA_ERR_ARG and "not a blob" in N.core().sa_last_error().decode()


def test_noalloc_switching_problem_sizes():
    """A smaller problem after a larger one reuses the front of the same array."""
    big = random_problem(1500, len1=600, len2_range=(10, 400), seed=21)
    small = random_problem(7, len1=300, len2_range=(1, 310), seed=22)
    for p in (big, small, big):
        blob = plan.plan_pack(p)
        bufs = _device_buffers(blob)
        assert np.array_equal(_solve_noalloc(blob, bufs), _solve_regular(blob, bufs))


def test_noalloc_rejects_problems_beyond_its_capacity():
    p = random_problem(N.SA_NOALLOC_MAX_SEQ + 1, len1=64, len2_range=(8, 16), seed=23)
    blob = plan.plan_pack(p)
    bufs = _device_buffers(blob)
    rc = N.cuda().sa_cuda_solve_blob_noalloc(C.c_void_p(bufs[0].data_ptr()), C.byref(blob.header),
                                             C.c_void_p(bufs[1].data_ptr()), None)
    assert rc == N.SA_ERR_TOO_LARGE
    msg = N.core().sa_last_error().decode()
    assert "exceed" in msg and str(N.SA_NOALLOC_MAX_SEQ) in msg
    # the ordinary entry point still takes it
    assert np.array_equal(_solve_regular(blob, bufs), solve_cpu(p, "fast"))
