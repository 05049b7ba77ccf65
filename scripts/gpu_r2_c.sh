#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fabric.py -m gpu -x -q > gpurun_out/pytest_fabric_2.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_fabric_2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2.json 2> gpurun_out/bench_2.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_2.json').read().strip().splitlines()[-1])
print('ms_per_step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'large',d['large']['kernel_ms'],'strong',d.get('strong'))
PY
tail -3 gpurun_out/bench_2.err
