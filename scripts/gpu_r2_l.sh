#!/bin/bash
# 1-GPU: concurrent range packers (tests + e2e), gpu tests.
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
timeout 900 python -m pytest tests/test_gpu_kernel.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/l_pytest.log
tail -3 gpurun_out/l_pytest.log
for mn in 1 2; do
  echo "== SA_PLAN_MAX_NARROW=$mn" | tee -a gpurun_out/l_narrow.txt
  SA_PLAN_MAX_NARROW=$mn python benchmarks/latency_timeline.py input3.txt input1.txt 2>&1 | tail -2 | tee -a gpurun_out/l_narrow.txt
  SA_PLAN_MAX_NARROW=$mn python bench.py --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench input3 ms_per_step', d['ms_per_step'], 'min', d.get('solve_ms_min'), 'e2e', d['e2e']['ms_per_step'])" | tee -a gpurun_out/l_narrow.txt
done
nproc | tee gpurun_out/l_pipeline.txt
python - >> gpurun_out/l_pipeline.txt 2>&1 <<'PY'
import os, time, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
from mpi_openmp_cuda_b200.models.synthetic import scaled_input3
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu
s = AlignmentSolver("cuda:0")
for scale in (512, 1024, 2048, 4096):
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
cat gpurun_out/l_pipeline.txt
timeout 600 python bench.py > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/l_bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e'], d.get('synthetic'), d.get('e2e_process'))"
timeout 120 build/cluster_probe > gpurun_out/l_cluster_probe.json 2> gpurun_out/l_cluster_probe.err; echo "cluster_probe rc=$?"; cat gpurun_out/l_cluster_probe.json
