#!/bin/bash
# Round-2 call C (2 GPUs): fabric v2 (header packet + pull scatter + self-validating result rows).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fabric.py tests/test_gpu_kernel.py -m gpu -x -q > gpurun_out/pytest_fabric_2.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_fabric_2.log
SA_FABRIC_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2_tl.json 2> gpurun_out/bench_2_tl.err; echo "bench(timeline) rc=$?"
grep -a "fabric timeline" gpurun_out/bench_2_tl.err | tail -4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2.json 2> gpurun_out/bench_2.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_2.json; tail -5 gpurun_out/bench_2.err
