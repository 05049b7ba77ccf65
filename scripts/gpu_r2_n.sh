#!/bin/bash
# N-GPU confirmation after the narrow-tile / launcher changes: fabric + CLI tests, bench line, launcher walls.
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
N=${1:-2}
wall() { python - "$@" <<'PY'
import subprocess, sys, time, os
tag, inp, gold = sys.argv[1], sys.argv[2], sys.argv[3]
cmd = sys.argv[4:]
t = time.perf_counter(); r = subprocess.run(cmd, stdin=open(inp, "rb"), capture_output=True, timeout=300); dt = time.perf_counter() - t
ok = r.stdout == open(gold, "rb").read()
print(f"{tag}: rc={r.returncode} wall_s={dt:.3f} identical={'yes' if ok else 'NO'} {r.stderr.decode()[-300:].strip()}", flush=True)
PY
}
{
for i in 1 2; do wall "native --gpus $N run$i" data/input3.txt data/golden/input3.out ./final --gpus $N --timing-json gpurun_out/n_native_$N.json; done
cat gpurun_out/n_native_$N.json
for i in 1; do SA_NATIVE_VISIBLE=own wall "native --gpus $N own-visible run$i" data/input3.txt data/golden/input3.out ./final --gpus $N --timing-json gpurun_out/n_native_own_$N.json; done
cat gpurun_out/n_native_own_$N.json
SA_FABRIC_MULTICAST=0 wall "native --gpus $N p2p" data/input3.txt data/golden/input3.out ./final --gpus $N
for k in 4 6; do wall "native --gpus $N input$k" data/input$k.txt data/golden/input$k.out ./final --gpus $N; done
SA_NATIVE_REPEAT=50 ./final --gpus $N --timing-json gpurun_out/n_native_rep_$N.json < data/input3.txt > /dev/null; cat gpurun_out/n_native_rep_$N.json
} 2>&1 | tee gpurun_out/n_native_${N}gpu.txt
if [ "$N" -le 2 ]; then
  timeout 900 python -m pytest tests/test_cli.py tests/test_gpu_fabric.py -m gpu -x -q > gpurun_out/n_pytest_${N}gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/n_pytest_${N}gpu.log
else
  timeout 600 python -m pytest tests/test_gpu_fabric.py -m gpu -x -q -k "(more_gpus_byte_identical or split_sequence) and $N" > gpurun_out/n_pytest_${N}gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/n_pytest_${N}gpu.log
fi
tail -4 gpurun_out/n_pytest_${N}gpu.log
SA_BENCH_PROC_RUNS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/n_bench_${N}gpu.json 2> gpurun_out/n_bench_${N}gpu.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/n_bench_${N}gpu.json").read().strip().splitlines()[-1])
print("headline ms", d["ms_per_step"], "min", d.get("solve_ms_min"), "e2e", d["e2e"]["ms_per_step"])
for k in ("large", "strong", "e2e_process"):
    print(k, json.dumps(d.get(k))[:500])
PY

timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29614 benchmarks/fabric_sweep.py --only pull,bcast --max-bytes $((64<<20)) --iters 5 --out gpurun_out/n_pull_sweep_$N.json > gpurun_out/n_pull_sweep_$N.txt 2> gpurun_out/n_pull_sweep_$N.err; echo "sweep rc=$?"
cat gpurun_out/n_pull_sweep_$N.txt | cut -c1-200
