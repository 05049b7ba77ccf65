#!/bin/bash
# Round-2 call D (1 GPU): full GPU tests after the blob v3 / fabric v2 change; staging probe.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python benchmarks/staging_probe.py --scale 512 > gpurun_out/staging_probe.json 2> gpurun_out/staging_probe.err; echo "probe rc=$?"; cat gpurun_out/staging_probe.json; tail -3 gpurun_out/staging_probe.err
timeout 300 python benchmarks/staging_probe.py --scale 1 --mode tile --iters 9 > gpurun_out/staging_probe_small.json 2>> gpurun_out/staging_probe.err; cat gpurun_out/staging_probe_small.json
