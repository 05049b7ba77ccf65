"""Host-side fabric plumbing that needs no GPU: buffer layout, descriptor ABI, reference partition."""
import ctypes as C

import pytest

from mpi_openmp_cuda_b200.parallel.partition import reference_partition


def test_reference_partition_matches_main_c():
    # main.c:110-119: rows = N / P, remainder on root; P >= N -> one row each, no remainder
    assert reference_partition(32, 2) == (16, 0)
    assert reference_partition(30, 8) == (3, 6)
    assert reference_partition(10, 4) == (2, 2)
    assert reference_partition(1, 2) == (1, 0)
    assert reference_partition(5, 8) == (1, 0)


def test_descriptor_struct_sizes_match_the_header():
    from mpi_openmp_cuda_b200.ops import _native as N

    assert C.sizeof(N.BlobHeader) == 1024 and C.sizeof(N.SeqDesc) == 16
    try:
        from mpi_openmp_cuda_b200.parallel.fabric import FabricDesc, FabricLayout
    except Exception as e:  # torch.distributed missing would be an environment problem, not ours
        pytest.skip(str(e))
    assert C.sizeof(FabricLayout) == 48
    assert C.sizeof(FabricDesc) == 24 + 8 * N.SA_MAX_WORLD + 8 + 48
    assert FabricDesc.peer_base.offset == 24 and FabricDesc.layout.offset == 24 + 8 * N.SA_MAX_WORLD + 8
