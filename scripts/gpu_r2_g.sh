#!/bin/bash
mkdir -p gpurun_out
timeout 90 build/tc_scan 512 > gpurun_out/tc_scan.json 2> gpurun_out/tc_scan.err && timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 1 -c 1 -f -o gpurun_out/prof_tc_scan build/tc_scan 512 > gpurun_out/ncu_tc_scan.log 2>&1
echo "rc=$?"; cat gpurun_out/tc_scan.json; cat gpurun_out/tc_scan.err
