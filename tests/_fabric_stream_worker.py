"""Worker for tests/test_gpu_fabric.py::test_two_gpus_stream_mode_through_the_fabric (run under torchrun)."""
import numpy as np
import torch.distributed as dist

from mpi_openmp_cuda_b200.models.synthetic import random_problem
from mpi_openmp_cuda_b200.ops import plan
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu
from mpi_openmp_cuda_b200.parallel.dist_solver import DistributedSolver
from mpi_openmp_cuda_b200.parallel.fabric import Fabric, init_distributed

rank, world, local = init_distributed("nccl")
solver = DistributedSolver(Fabric(max_seq=4096, max_blob_bytes=8 << 20))
p = random_problem(1501, len1=900, len2_range=(1, 700), seed=4)
blob = plan.plan_pack(p, world, mode="stream") if rank == 0 else None
for _ in range(2):  # twice: scratch and flags must be reusable
    solver.launch_packed(blob)
    res = solver.fetch(1501)
    solver.check()
    if rank == 0:
        assert np.array_equal(res, solve_cpu(p, "fast"))
dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print("stream fabric ok")
