#!/bin/bash
# ncu capture of the scoring kernel (run via gpurun, one GPU). Plain run first, then ncu.
mkdir -p gpurun_out
SCALE=${1:-256}
TAG=${2:-tput}
python benchmarks/profile_target.py --scale $SCALE --iters 3 --check > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sa_score -s 2 -c 1 -f \
    -o gpurun_out/prof_$TAG python benchmarks/profile_target.py --scale $SCALE --iters 3 > gpurun_out/ncu_$TAG.log 2>&1
echo "exit $?"; tail -3 gpurun_out/plain_$TAG.log; tail -5 gpurun_out/ncu_$TAG.log
