#!/bin/bash
mkdir -p gpurun_out
ls -la cuda-v-mpi_b200 | head -3
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 3 2>/dev/null > gpurun_out/bench_g1.json; cat gpurun_out/bench_g1.json
timeout 300 python bench.py --steps 50 --warmup 5 --log2n 20 2>/dev/null > gpurun_out/bench_small_g1.json; cut -c1-1300 gpurun_out/bench_small_g1.json
