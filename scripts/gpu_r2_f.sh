#!/bin/bash
# ncu: the fabric kernel as rank 1 on one GPU, fresh (pull + score) vs resident (score only)
mkdir -p gpurun_out
SA_STREAM_CTAS_PER_SM=1 timeout 300 python benchmarks/staging_probe.py --scale 512 --iters 1 > gpurun_out/plain_probe.log 2>&1 &&
SA_STREAM_CTAS_PER_SM=1 timeout 800 ncu --set full --clock-control none --import-source on -k regex:sa_score_fabric_kernel -s 0 -c 2 -f -o gpurun_out/prof_fabric python benchmarks/staging_probe.py --scale 512 --iters 1 > gpurun_out/ncu_probe.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_probe.log
