#!/bin/bash
# compute-sanitizer over the three kernels (one tool per gpurun call, smallest inputs that
# exercise every path). Usage: gpu_sanitize.sh memcheck|racecheck|synccheck|initcheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
cat > /tmp/sanitize_target.py <<'PY'
import sys
sys.path.insert(0, ".")
import numpy as np
from mpi_openmp_cuda_b200 import Problem
from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver
from mpi_openmp_cuda_b200.models.synthetic import random_problem
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu
s = AlignmentSolver("cuda:0")
for name in ("input1.txt", "input6.txt", "input2.txt"):
    p = Problem.from_file("data/" + name)
    for mode in ("tile", "pair", "stream"):
        assert np.array_equal(s.solve_packed(s.pack(p, mode=mode)), solve_cpu(p)), (name, mode)
p = random_problem(40, len1=700, len2_range=(1, 703), seed=3)
for mode in ("tile", "pair", "stream"):
    assert np.array_equal(s.solve_packed(s.pack(p, mode=mode)), solve_cpu(p)), mode
assert np.array_equal(s.solve(p), solve_cpu(p))
print("sanitize target ok")
PY
python /tmp/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 &&
compute-sanitizer --tool $TOOL --print-limit 20 python /tmp/sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "exit $?"; tail -3 gpurun_out/sanitize_plain.log; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize target ok|=========.*(Error|hazard)" gpurun_out/sanitize_$TOOL.log | head -20
