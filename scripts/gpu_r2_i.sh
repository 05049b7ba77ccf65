#!/bin/bash
# 1-GPU round: gpu tests, bench line, native worker-process path on one GPU with process wall times.
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-/root/repo}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/i_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest.log
tail -5 gpurun_out/i_pytest.log
timeout 600 python bench.py > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/i_bench.json
for mode in plain fabric; do
  for i in 1 2 3; do
    s=$(date +%s.%N)
    if [ $mode = plain ]; then ./final < data/input3.txt > /tmp/o.txt 2> /tmp/e.txt; else ./final --gpus 1 --fabric --timing-json gpurun_out/i_native_timing_$i.json < data/input3.txt > /tmp/o.txt 2> /tmp/e.txt; fi
    rc=$?; e=$(date +%s.%N)
    echo "$mode run $i rc=$rc wall_s=$(echo "$e - $s" | bc) identical=$(cmp -s /tmp/o.txt data/golden/input3.out && echo yes || echo no)" | tee -a gpurun_out/i_cli_wall.txt
    head -3 /tmp/e.txt
  done
done
cat gpurun_out/i_native_timing_3.json
SA_NATIVE_REPEAT=20 ./final --gpus 1 --fabric --timing-json gpurun_out/i_native_rep.json < data/input3.txt > /dev/null; cat gpurun_out/i_native_rep.json
