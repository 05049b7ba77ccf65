#!/bin/bash
# 8 GPUs: cross-GPU tests (byte identity at 4/8, split-sequence argmax at 2/4/8) and the bench at N=8
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fabric.py -m gpu -x -q -k "split or more_gpus or stream_mode" > gpurun_out/pytest_fabric_8.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_fabric_8.log
SA_BENCH_PROC_RUNS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29618 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_8.json 2> gpurun_out/bench_8.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_8.json').read().strip().splitlines()[-1])
print('ms_per_step',d['ms_per_step'],'min',d['solve_ms_min'],'e2e',d['e2e']['ms_per_step'],'proc',d.get('e2e_process'),'large',d['large']['kernel_ms'],d['large']['kernel_ms_per_rank'],'strong',d.get('strong'))
PY
tail -3 gpurun_out/bench_8.err
