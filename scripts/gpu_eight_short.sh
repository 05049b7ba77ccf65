#!/bin/bash
# Short 8-GPU rerun after a fabric change: benches at 2/4/8 (+timeline at 8) and the sweep at 8.
mkdir -p gpurun_out
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + n)) \
      bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/bench_$n.json 2> gpurun_out/bench_$n.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$n.json").read().strip().splitlines()[-1])
print("N=$n", "ms_per_step", round(d["ms_per_step"],4), "min", round(d["solve_ms_min"],4), "e2e", round(d["e2e"]["ms_per_step"],4), "large_ms", round(d["large"]["kernel_ms"],4))
PY
done
SA_FABRIC_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29899 \
    bench.py --gpus 8 --steps 5 --warmup 3 --large 0 2>&1 | grep timeline | cut -c1-1600 | tee gpurun_out/timeline_8.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29908 \
    benchmarks/fabric_sweep.py --max-bytes 1073741824 > gpurun_out/sweep_8.log 2> gpurun_out/sweep_8.err
echo "sweep 8 exit $?"; grep -c '"ok": true' gpurun_out/sweep_8.log; grep -E '"bytes": (1024|1048576|1073741824),' gpurun_out/sweep_8.log | cut -c1-200
