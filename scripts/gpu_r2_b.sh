#!/bin/bash
# Round-2 call B (1 GPU): GPU tests, pipe micro-benchmarks, latency timeline, bench (difference-form kernels).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 180 build/microbench > gpurun_out/MEASURED_PIPES.json 2> gpurun_out/microbench.err; echo "microbench rc=$?"
SA_TIMELINE=1 timeout 300 python benchmarks/latency_timeline.py > gpurun_out/latency_timeline.jsonl 2> gpurun_out/latency_timeline.err; echo "timeline rc=$?"; cat gpurun_out/latency_timeline.jsonl
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_1.json
timeout 300 python benchmarks/kernel_sweep.py > gpurun_out/kernel_sweep.log 2>&1; tail -15 gpurun_out/kernel_sweep.log
