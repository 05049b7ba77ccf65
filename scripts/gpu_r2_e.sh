#!/bin/bash
mkdir -p gpurun_out
run() { echo "== SA_DEBUG_SRC=$1 $2"; SA_DEBUG_SRC=$1 timeout 300 python benchmarks/staging_probe.py --scale 512 --iters 4 $2 2>> gpurun_out/staging_probe.err | python -c "import sys,json; d=json.load(sys.stdin); print({k:(d[k]['event_us'], d[k]['timeline']) for k in ('fresh','resident')})"; }
run 7 ""
run 3 "--no-flush"
run 0 "--no-flush"
