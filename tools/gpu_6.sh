#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sweep_n.py > gpurun_out/sweep_n_v2.jsonl 2>&1; cat gpurun_out/sweep_n_v2.jsonl
timeout 300 python bench.py --steps 3 --warmup 1 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:quad_kernel -s 1 -c 1 -o gpurun_out/prof_quad_v2 python bench.py --steps 3 --warmup 1 > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"
