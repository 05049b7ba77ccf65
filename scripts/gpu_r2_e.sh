#!/bin/bash
mkdir -p gpurun_out
SA_STREAM_CTAS_PER_SM=1 timeout 300 python benchmarks/staging_probe.py --scale 512 --iters 2 2>> gpurun_out/staging_probe.err | python -c "import sys,json; d=json.load(sys.stdin); print({k:(d[k]['event_us'], d[k]['timeline']) for k in ('fresh','resident','fresh_again')})"
