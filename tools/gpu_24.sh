#!/bin/bash
# reference arm at 1/2/4/8 workers (host CPUs of the GPU box) + nproc
mkdir -p gpurun_out
nproc > gpurun_out/ref_nproc.txt; lscpu | grep "Model name" >> gpurun_out/ref_nproc.txt
for g in 1 2 4 8; do
  timeout 600 python bench.py --impl reference --gpus $g --steps 2 --warmup 1 > gpurun_out/ref_g$g.json 2>/dev/null; cut -c1-2500 gpurun_out/ref_g$g.json
done
