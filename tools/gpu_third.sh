#!/bin/bash
mkdir -p gpurun_out
timeout 300 bin/rot_bench > gpurun_out/rot_bench.jsonl 2>&1; cat gpurun_out/rot_bench.jsonl
nproc; timeout 900 python bench.py --impl reference --gpus 1 --steps 2 --warmup 0 > gpurun_out/ref_arm_1.json 2> gpurun_out/ref_arm_1.err; cat gpurun_out/ref_arm_1.json; tail -3 gpurun_out/ref_arm_1.err
