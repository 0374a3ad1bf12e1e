#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v8.jsonl 2>&1; cut -c1-330 gpurun_out/bench_scan_v8.jsonl
