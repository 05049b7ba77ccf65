#!/bin/bash
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/s_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/s_pytest.log
tail -2 gpurun_out/s_pytest.log
for v in 0 4096; do
SA_FLAG_MAX_SEQ=$v timeout 300 python bench.py --steps 50 --warmup 10 2> /dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SA_FLAG_MAX_SEQ=$v', 'device', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])" | tee -a gpurun_out/s_flag.txt
done
python benchmarks/suite.py 2>/dev/null | tail -8 | tee gpurun_out/s_suite.txt
