"""CPU-only multi-process configuration (BASELINE.json configs[0]: "2 ranks on CPU, scoring kernel
swapped for its CPU reference; plumbing and output-diff check, no GPU").

Same roles as the GPU fabric, with gloo collectives standing in for the in-kernel ones: rank 0
plans + packs, the blob is broadcast, every rank scores its own item range into packed keys
with the CPU walk of the kernel's tile logic, keys are max-reduced, rank 0 finalises."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from ..ops import _native as N
from ..ops import plan


def solve_cpu_world(problem=None) -> np.ndarray | None:
    rank, world = dist.get_rank(), dist.get_world_size()
    if rank == 0:
        blob = plan.plan_pack(problem, world)
        size = torch.tensor([blob.nbytes], dtype=torch.int64)
    else:
        size = torch.zeros(1, dtype=torch.int64)
    dist.broadcast(size, 0)
    nbytes = int(size.item())
    data = torch.from_numpy(blob.numpy()[:nbytes]) if rank == 0 else torch.empty(nbytes, dtype=torch.uint8)
    dist.broadcast(data, 0)  # MPI_Bcast + MPI_Scatter of the reference in one message
    arr = data.numpy()
    hdr = N.BlobHeader.from_buffer_copy(arr[:256].tobytes())
    first, last = int(hdr.rank_first_item[rank]), int(hdr.rank_first_item[rank + 1])
    keys = np.zeros(max(int(hdr.nseq), 1), dtype=np.uint64)
    N.check(N.core().sa_blob_score_items_cpu(C.c_void_p(arr.ctypes.data), nbytes, first, last,
                                             keys.ctypes.data_as(C.POINTER(C.c_uint64))))
    # unsigned max over ranks: flip the top bit so that signed int64 MAX orders the keys correctly
    t = torch.from_numpy((keys ^ np.uint64(1 << 63)).view(np.int64).copy())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    keys = t.numpy().view(np.uint64) ^ np.uint64(1 << 63)
    out = np.zeros((max(int(hdr.nseq), 1), 4), dtype=np.int32)
    N.check(N.core().sa_blob_finalize_cpu(C.c_void_p(arr.ctypes.data), nbytes, keys.ctypes.data_as(C.POINTER(C.c_uint64)),
                                          out.ctypes.data_as(C.POINTER(C.c_int32))))
    return out[: int(hdr.nseq), :3].copy()
