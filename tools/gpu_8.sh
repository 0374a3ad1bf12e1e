#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 1 --log2n 18 > gpurun_out/plain18.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:quad_kernel -s 1 -c 1 -o gpurun_out/prof_quad_n18 python bench.py --steps 3 --warmup 1 --log2n 18 > gpurun_out/ncu18.log 2>&1
echo "ncu rc=$?"
