#!/bin/bash
# 1-GPU: scan occupancy variants
mkdir -p gpurun_out
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-130 gpurun_out/bench_scan.jsonl
B200QUAD_LIB=$PWD/build/variants/scan3/libb200quad.so timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_occ3.jsonl 2>&1; echo "== 3 CTAs/SM"; cut -c1-130 gpurun_out/bench_scan_occ3.jsonl
