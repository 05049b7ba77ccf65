#!/bin/bash
# Final 1-GPU confirmation of the round: smoke, all gpu tests, bench line, ncu of the final TILE and STREAM kernels.
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/o_pytest.log
tail -3 gpurun_out/o_pytest.log
timeout 600 python bench.py > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/o_bench.json
timeout 300 bash scripts/gpu_ncu.sh 1 tile_final_r2
timeout 300 bash scripts/gpu_ncu.sh 1024 stream_final_r2
ls -la gpurun_out/*.ncu-rep
