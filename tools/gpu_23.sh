#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 12 -c 1 -o gpurun_out/prof_scan_v8 python tools/bench_scan.py > gpurun_out/ncu_scan.log 2>&1; echo "ncu scan rc=$?"; tail -3 gpurun_out/ncu_scan.log
