#!/bin/bash
# Everything that needs the 8-GPU box, in one call (8x GPU-minutes are charged).
mkdir -p gpurun_out
run_bench() {  # $1 = N
    SA_FABRIC_TIMELINE=${TL:-0} timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 \
        --master-port $((29800 + $1)) bench.py --gpus $1 --steps 20 --warmup 5 > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
    echo "bench $1 exit $?"; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$1.json").read().strip().splitlines()[-1])
print("N=$1", "ms_per_step", round(d["ms_per_step"],4), "min", round(d["solve_ms_min"],4), "e2e", round(d["e2e"]["ms_per_step"],4), "large", d.get("large"))
PY
    grep -h "timeline" gpurun_out/bench_$1.err | cut -c1-1200
}
echo "== fabric test at 8 ranks"; timeout 600 python -m pytest tests/test_gpu_fabric.py -x -q -k "more_gpus" 2>&1 | tail -3
for n in 2 4 8; do run_bench $n; done
TL=1 run_bench 8 > gpurun_out/timeline_8.txt 2>&1; grep timeline gpurun_out/timeline_8.txt | cut -c1-1500
echo "== fabric sweep"; for n in 4 8; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29900 + n)) \
        benchmarks/fabric_sweep.py --max-bytes 1073741824 > gpurun_out/sweep_$n.log 2> gpurun_out/sweep_$n.err
    echo "sweep $n exit $?"; grep -c '"ok": true' gpurun_out/sweep_$n.log; grep -c '"ok": false' gpurun_out/sweep_$n.log
done
echo "== MPI stand-in collectives (host memory)"; for n in 2 4 8; do
    SHIM_BYTES=1200000000 timeout 120 baseline/_ref/bin/mpiexec -np $n baseline/_ref/bin/bench_coll 1073741824 > gpurun_out/mpi_coll_$n.log 2>&1
    echo "mpi coll $n exit $?"; tail -1 gpurun_out/mpi_coll_$n.log
done
echo "== reference at 2/4/8 ranks"; timeout 900 python baseline/run_ref.py --inputs 4 3 --np 2 4 8 --repeat 1 --variants final \
    --out gpurun_out/ref_times_multi.json 2>&1 | cut -c1-330
echo "== reference bench arm N=2"; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
    --master-port 29950 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -1 | cut -c1-900
