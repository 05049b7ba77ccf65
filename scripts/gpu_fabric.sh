#!/bin/bash
# Fabric checks on an N-GPU box: tests, then the bench at N GPUs through torchrun.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_fabric.py -x -q > gpurun_out/pytest_fabric_$N.log 2>&1
echo "pytest exit $?"; tail -12 gpurun_out/pytest_fabric_$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err
echo "bench exit $?"; cat gpurun_out/bench_$N.json; tail -5 gpurun_out/bench_$N.err
