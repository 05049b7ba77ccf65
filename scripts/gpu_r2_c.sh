#!/bin/bash
# 2 GPUs: fabric v2 after the warp-uniform wait fix; large config with 2 and 1 CTAs per SM
mkdir -p gpurun_out
for cps in 2 1; do
SA_STREAM_CTAS_PER_SM=$cps timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2961$cps bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2_cps$cps.json 2> gpurun_out/bench_2_cps$cps.err; echo "bench cps=$cps rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_2_cps$cps.json').read().strip().splitlines()[-1])
print('ms_per_step',d['ms_per_step'],'min',d['solve_ms_min'],'e2e',d['e2e']['ms_per_step'],'proc',d.get('e2e_process',{}).get('wall_ms'),'large',d['large']['kernel_ms'],d['large']['kernel_ms_per_rank'])
PY
done
