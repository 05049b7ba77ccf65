#!/bin/bash
# STREAM profile build with letter pairs: numerics (all stream tests) and kernel time.
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
timeout 600 python -m pytest tests/test_gpu_kernel.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/q_pytest.log
tail -3 gpurun_out/q_pytest.log
python benchmarks/kernel_sweep.py 2>/dev/null | tee gpurun_out/q_kernel_sweep.txt | tail -12
