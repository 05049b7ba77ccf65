"""Worker for tests/test_gpu_fabric.py::test_split_sequence_cross_gpu_argmax (run under torchrun).

A few LONG sequences whose offset tiles are dealt to different GPUs (plan_pack(split_tiles=True)), so that
every sequence's best (score, n) is reduced across GPUs with atomicMax.sys on its home rank and the
last-arriving tile (on whichever GPU) finishes it — the path that replaces the reference's Gather
(main.c:195-197) when one row is scored by several ranks. Two solves in a row: the remote key / counter
slots must be left zeroed by the first."""
import numpy as np
import torch.distributed as dist

from mpi_openmp_cuda_b200 import Problem
from mpi_openmp_cuda_b200.ops import plan
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu
from mpi_openmp_cuda_b200.parallel.dist_solver import DistributedSolver
from mpi_openmp_cuda_b200.parallel.fabric import Fabric, init_distributed

rank, world, local = init_distributed("nccl")
solver = DistributedSolver(Fabric(max_seq=4096, max_blob_bytes=8 << 20))

rng = np.random.default_rng(7)
letters = np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZ", dtype=np.uint8)
seq1 = letters[rng.integers(0, 26, size=3000)].tobytes()
seqs = []
for m in (100, 700, 1500, 333, 1999):
    s = bytearray(letters[rng.integers(0, 26, size=m)].tobytes())
    seqs.append(bytes(s))
# ties: a periodic Seq1 stretch makes the same best score appear at several offsets (first one must win),
# and an exact substring pins a known winner far from offset 0
seqs.append(seq1[2100:2100 + 640])
tie1 = (b"ACDEFGHIKL" * 300)[:3000]
tie_seqs = [b"ACDEFGHIKL" * 20, b"GHIKLACDEF" * 31]

for s1, ss, oracle in ((seq1, seqs, "fast"), (tie1, tie_seqs, "fast"), (seq1[:400], [seqs[0], seq1[50:350]], "naive")):
    p = Problem.from_parts((5, 2, 3, 7), s1, ss)
    blob = plan.plan_pack(p, world, mode="tile", split_tiles=True, out=solver._pinned_in(1 << 20)) if rank == 0 else None
    # (a sequence of two or three small tiles may legitimately end up on one GPU: tiles go to the lightest rank)
    if rank == 0 and world > 1 and int(blob.seq_table()["ntiles"].max()) >= 4:
        h, items = blob.header, blob.items()
        owners = {}
        for r in range(world):
            for s in items[h.rank_first_item[r]: h.rank_first_item[r + 1]]["seq"]:
                owners.setdefault(int(s), set()).add(r)
        assert any(len(v) > 1 for v in owners.values()), "no sequence was split over GPUs"
    for _ in range(2):
        solver.launch_packed(blob)
        res = solver.fetch(len(ss))
        solver.check()
        if rank == 0:
            want = solve_cpu(p, oracle)
            assert np.array_equal(res, want), (res.tolist(), want.tolist())
dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print("split fabric ok")
