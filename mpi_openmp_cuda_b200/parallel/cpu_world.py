"""CPU-only multi-process configuration (BASELINE.json configs[0]: "2 ranks on CPU, scoring kernel
swapped for its CPU reference; plumbing and output-diff check, no GPU").

Same roles and the same data movement as the GPU fabric, with gloo collectives standing in for the
in-kernel ones: rank 0 plans + packs; the 1 KB header is broadcast; every rank then receives ONLY
the byte ranges the header assigns to it (`sa_blob_rank_ranges`: the shared table/Seq1 plus its own
slices of the item list, the records and the Seq2 letters — the reference's MPI_Bcast + MPI_Scatter,
main.c:149-152,174) into an otherwise zero-filled blob; it scores its own item range into packed
keys with the CPU walk of the kernel's tile logic; keys are max-reduced; rank 0 finalises. Scoring
from the sparse blob is the proof that a rank's ranges cover everything its kernel reads."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from ..ops import _native as N
from ..ops import plan

_HDR = C.sizeof(N.BlobHeader)


def rank_ranges(hdr: N.BlobHeader, rank: int) -> list[tuple[int, int]]:
    lo, hi = (C.c_uint32 * 4)(), (C.c_uint32 * 4)()
    N.check(N.core().sa_blob_rank_ranges(C.byref(hdr), rank, lo, hi))
    return [(int(a), int(b)) for a, b in zip(lo, hi) if b > a]


def solve_cpu_world(problem=None) -> np.ndarray | None:
    rank, world = dist.get_rank(), dist.get_world_size()
    head = torch.zeros(_HDR, dtype=torch.uint8)
    if rank == 0:
        blob = plan.plan_pack(problem, world)
        full = blob.numpy()[: blob.nbytes]
        head.copy_(torch.from_numpy(full[:_HDR].copy()))
    dist.broadcast(head, 0)  # the header packet
    hdr = N.BlobHeader.from_buffer_copy(head.numpy().tobytes())
    nbytes = int(hdr.total_bytes)
    mine = rank_ranges(hdr, rank)
    if rank == 0:
        arr = full
        for r in range(1, world):  # the scatter: rank r gets its ranges, back to back, and nothing else
            parts = [torch.from_numpy(full[a:b].copy()) for a, b in rank_ranges(hdr, r)]
            dist.send(torch.cat(parts) if parts else torch.zeros(1, dtype=torch.uint8), r)
    else:
        got = torch.zeros(max(1, sum(b - a for a, b in mine)), dtype=torch.uint8)
        dist.recv(got, 0)
        arr = np.zeros(nbytes, dtype=np.uint8)  # everything this rank was NOT sent stays zero
        arr[:_HDR] = head.numpy()
        at = 0
        for a, b in mine:
            arr[a:b] = got.numpy()[at : at + (b - a)]
            at += b - a
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
