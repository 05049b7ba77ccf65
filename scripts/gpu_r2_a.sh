#!/bin/bash
# Round-2 call A (1 GPU): pipe micro-benchmarks, reference arm with the CUPTI probe, parity ledger, own bench.
mkdir -p gpurun_out
timeout 180 build/microbench > gpurun_out/MEASURED_PIPES.json 2> gpurun_out/microbench.err; echo "microbench rc=$?"
timeout 400 python bench.py --impl reference --steps 1 --warmup 3 > gpurun_out/ref_arm.json 2> gpurun_out/ref_arm.err; echo "ref arm rc=$?"
tail -c 1500 gpurun_out/ref_arm.json
timeout 500 python baseline/parity_ledger.py --inputs 1 3 --repeat 3 > gpurun_out/parity.log 2>&1; echo "parity rc=$?"
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_1.json
