#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep_quad.py 32 > gpurun_out/sweep32.jsonl 2>&1; tail -70 gpurun_out/sweep32.jsonl
timeout 120 ./bin/riemann --json; timeout 120 ./bin/cintegrate --json; timeout 120 ./bin/train4 --json
timeout 300 python bench.py --steps 3 --warmup 1 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:quad_kernel -s 1 -c 2 -o gpurun_out/prof_quad python bench.py --steps 3 --warmup 1 > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/ncu.log
