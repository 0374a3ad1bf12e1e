#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v6.jsonl 2>&1; cut -c1-330 gpurun_out/bench_scan_v6.jsonl
timeout 300 python tools/bench_scan_world.py 2 > gpurun_out/bench_scan_world2.jsonl 2>&1; cat gpurun_out/bench_scan_world2.jsonl
timeout 300 python bench.py --steps 20 --warmup 3 2>/dev/null | cut -c1-400
