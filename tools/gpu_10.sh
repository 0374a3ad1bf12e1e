#!/bin/bash
# v6 validation on one B200
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; echo "bench rc=$?"; cat gpurun_out/bench_g1.json; tail -3 gpurun_out/bench_g1.err
timeout 600 python tools/sweep_n.py > gpurun_out/sweep_n_v6.jsonl 2>&1; grep '"eval": "auto"' gpurun_out/sweep_n_v6.jsonl
timeout 120 python tools/trace_quad.py 32 > gpurun_out/trace32_v6.json 2>&1; cat gpurun_out/trace32_v6.json
timeout 600 python tools/sweep_quad.py 32 > gpurun_out/sweep32_v6.jsonl 2>&1; grep -v '"rows": 16' gpurun_out/sweep32_v6.jsonl | grep rotate
timeout 600 ncu --set full --clock-control none --import-source on -k regex:quad_kernel -s 3 -c 1 -o gpurun_out/prof_quad_v6 python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_quad.log 2>&1; echo "ncu quad rc=$?"
