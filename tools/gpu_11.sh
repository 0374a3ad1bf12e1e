#!/bin/bash
# scan v2 validation + occupancy sweep on one B200
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
for occ in 2 3 4; do
  B200QUAD_SCAN_CTAS=$occ timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v2_occ$occ.jsonl 2>&1; cat gpurun_out/bench_scan_v2_occ$occ.jsonl
done
timeout 120 ./bin/train4 --json; timeout 120 ./bin/riemann --json; timeout 60 ./bin/cintegrate --json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 2 -c 1 -o gpurun_out/prof_scan_v2 python tools/bench_scan.py > gpurun_out/ncu_scan.log 2>&1; echo "ncu scan rc=$?"
