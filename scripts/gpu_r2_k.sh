#!/bin/bash
# 1-GPU: pipelined large-batch solve (tests + e2e), pack phase times on the box's host, cold start of final.
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
timeout 900 python -m pytest tests/test_gpu_kernel.py tests/test_cli.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/k_pytest.log
tail -3 gpurun_out/k_pytest.log
nproc
for mn in 0 1; do
  echo "== SA_PLAN_MAX_NARROW=$mn" | tee -a gpurun_out/k_narrow.txt
  SA_PLAN_MAX_NARROW=$mn python benchmarks/latency_timeline.py input3.txt input1.txt input4.txt 2>&1 | tail -3 | tee -a gpurun_out/k_narrow.txt
  SA_PLAN_MAX_NARROW=$mn python bench.py --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench input3 ms_per_step', d['ms_per_step'], 'min', d.get('solve_ms_min'), 'e2e', d['e2e']['ms_per_step'])" | tee -a gpurun_out/k_narrow.txt
done
python - > gpurun_out/k_pipeline.txt 2>&1 <<'PY'
import os, time, subprocess, sys
import numpy as np, torch
sys.path.insert(0, os.getcwd())
from mpi_openmp_cuda_b200.models.synthetic import scaled_input3
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu
s = AlignmentSolver("cuda:0")
for scale in (256, 512, 1024, 2048):
    p = scaled_input3(scale, "data/input3.txt")
    ref = solve_cpu(p, "fast") if scale <= 1024 else None
    for pipe in (0, 16384):
        s.session.set_pipeline(pipe)
        for _ in range(3): r = s.solve(p)
        ts = []
        for _ in range(10):
            t = time.perf_counter(); r = s.solve(p); ts.append(time.perf_counter() - t)
        ok = "n/a" if ref is None else bool(np.array_equal(r, ref))
        print(f"input3 x{scale} ({p.nseq} seqs) pipeline_min_seq={pipe}: e2e ms min {min(ts)*1e3:.3f} med {sorted(ts)[5]*1e3:.3f} launches {s.session.last_launches()} equal_oracle {ok}", flush=True)
PY
cat gpurun_out/k_pipeline.txt
SA_PACK_TIMING=1 python - 2>&1 <<'PY' | tail -16 | tee gpurun_out/k_pack_phases.txt
import os, sys, time, torch
sys.path.insert(0, os.getcwd())
from mpi_openmp_cuda_b200.models.synthetic import scaled_input3
from mpi_openmp_cuda_b200.ops import plan
p = scaled_input3(1024, "data/input3.txt")
buf = torch.empty(64 << 20, dtype=torch.uint8, pin_memory=True)
for _ in range(4):
    t = time.perf_counter(); b = plan.plan_pack(p, 1, out=buf); print("pack ms %.3f" % ((time.perf_counter() - t) * 1e3), file=sys.stderr)
PY
for i in 1 2 3; do python - <<'PY'
import subprocess, time
t = time.perf_counter(); r = subprocess.run(["./final"], stdin=open("data/input3.txt", "rb"), capture_output=True); dt = time.perf_counter() - t
print(f"final cold wall_s={dt:.3f} rc={r.returncode} identical={r.stdout == open('data/golden/input3.out','rb').read()}")
PY
done | tee gpurun_out/k_cold.txt
build/startup_probe | tee -a gpurun_out/k_cold.txt
timeout 600 python bench.py > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/k_bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e'], d.get('synthetic'), d.get('e2e_process'))"
