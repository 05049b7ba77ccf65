#!/bin/bash
# 2-GPU round: native launcher vs python launcher (byte identity, process wall, kernel time), visibility
# and multicast variants, cold-start probe.
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
print(f"{tag}: rc={r.returncode} wall_s={dt:.3f} identical={'yes' if ok else 'NO'} {r.stderr.decode()[-300:].strip()}")
PY
}
{
for v in "" 0; do
  echo "== startup probe CUDA_VISIBLE_DEVICES='$v'"
  if [ -z "$v" ]; then build/startup_probe; build/startup_probe; else CUDA_VISIBLE_DEVICES=$v build/startup_probe; CUDA_VISIBLE_DEVICES=$v build/startup_probe; fi
done
for i in 1 2; do wall "final 1gpu run$i" data/input3.txt data/golden/input3.out ./final; done
for i in 1 2 3; do wall "native --gpus $N run$i" data/input3.txt data/golden/input3.out ./final --gpus $N --timing-json gpurun_out/j_native_$N.json; done
cat gpurun_out/j_native_$N.json
for i in 1 2; do SA_FABRIC_MULTICAST=0 wall "native --gpus $N p2p run$i" data/input3.txt data/golden/input3.out ./final --gpus $N --timing-json gpurun_out/j_native_p2p_$N.json; done
cat gpurun_out/j_native_p2p_$N.json
for i in 1 2; do SA_NATIVE_VISIBLE=own wall "native --gpus $N own-visible run$i" data/input3.txt data/golden/input3.out ./final --gpus $N --timing-json gpurun_out/j_native_own_$N.json; done
cat gpurun_out/j_native_own_$N.json
for k in 1 4 6; do wall "native --gpus $N input$k" data/input$k.txt data/golden/input$k.out ./final --gpus $N; done
SA_NATIVE_REPEAT=50 ./final --gpus $N --timing-json gpurun_out/j_native_rep_$N.json < data/input3.txt > /dev/null; cat gpurun_out/j_native_rep_$N.json
SA_NATIVE_REPEAT=50 SA_FABRIC_MULTICAST=0 ./final --gpus $N --timing-json gpurun_out/j_native_rep_p2p_$N.json < data/input3.txt > /dev/null; cat gpurun_out/j_native_rep_p2p_$N.json
for i in 1 2; do wall "python launcher --gpus $N run$i" data/input3.txt data/golden/input3.out ./final --gpus $N --python-launcher; done
} 2>&1 | tee gpurun_out/j_native_${N}gpu.txt
timeout 900 python -m pytest tests/test_cli.py tests/test_gpu_fabric.py -m gpu -x -q > gpurun_out/j_pytest_${N}gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/j_pytest_${N}gpu.log
tail -4 gpurun_out/j_pytest_${N}gpu.log
