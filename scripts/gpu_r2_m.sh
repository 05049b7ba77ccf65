#!/bin/bash
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
timeout 900 python -m pytest tests/test_gpu_kernel.py -m gpu -x -q > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/m_pytest.log
tail -3 gpurun_out/m_pytest.log
python - > gpurun_out/m_pipeline.txt 2>&1 <<'PY'
import os, time, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
from mpi_openmp_cuda_b200.models.synthetic import scaled_input3
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu
s = AlignmentSolver("cuda:0")
for scale in (512, 1024, 2048):
    p = scaled_input3(scale, "data/input3.txt")
    ref = solve_cpu(p, "fast") if scale <= 1024 else None
    for pipe in (0, 16384):
        s.session.set_pipeline(pipe)
        for _ in range(3): r = s.solve(p)
        ts = []
        for _ in range(12):
            t = time.perf_counter(); r = s.solve(p); ts.append(time.perf_counter() - t)
        ok = "n/a" if ref is None else bool(np.array_equal(r, ref))
        print(f"input3 x{scale} ({p.nseq} seqs) pipeline_min_seq={pipe}: e2e ms min {min(ts)*1e3:.3f} med {sorted(ts)[6]*1e3:.3f} launches {s.session.last_launches()} equal_oracle {ok}", flush=True)
PY
cat gpurun_out/m_pipeline.txt
