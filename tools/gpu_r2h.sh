#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_target.py scan 26 > gpurun_out/plain_scan.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 2 -c 1 -o gpurun_out/prof_scan2 -f python tools/ncu_target.py scan 26 > gpurun_out/ncu_scan2.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/plain_scan.log
