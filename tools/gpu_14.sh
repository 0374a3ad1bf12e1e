#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
B200QUAD_SCAN_CTAS=2 timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v3_occ2.jsonl 2>&1; cat gpurun_out/bench_scan_v3_occ2.jsonl
B200QUAD_SCAN_CTAS=3 timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v3_occ3.jsonl 2>&1; cat gpurun_out/bench_scan_v3_occ3.jsonl
timeout 300 python tools/sweep_profile.py > gpurun_out/sweep_profile_v6.jsonl 2>&1; cat gpurun_out/sweep_profile_v6.jsonl
bash tools/sanitize.sh
