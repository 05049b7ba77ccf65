#!/bin/bash
# 4-GPU bench line exactly as the driver launches it (the one GPU count not measured after the narrow-tile change).
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
N=4
SA_BENCH_PROC_RUNS=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29615 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/p_bench_${N}gpu.json 2> gpurun_out/p_bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/p_bench_${N}gpu.json").read().strip().splitlines()[-1])
print("headline ms", d["ms_per_step"], "min", d.get("solve_ms_min"), "e2e", d["e2e"]["ms_per_step"], "launches", d.get("gpu_launches"), d.get("clocks"))
for k in ("large", "strong", "e2e_process"):
    print(k, json.dumps(d.get(k))[:400])
PY
tail -3 gpurun_out/p_bench_${N}gpu.err
