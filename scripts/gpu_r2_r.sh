#!/bin/bash
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r_pytest.log
tail -2 gpurun_out/r_pytest.log
timeout 600 python bench.py --steps 50 --warmup 10 > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r_bench.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d.get('solve_ms_min'), d['e2e'], d.get('synthetic'), d.get('e2e_process'))"
