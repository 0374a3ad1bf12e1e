#!/bin/bash
# v5 validation on one B200: tests, bench, N sweep, scan bench, CTA trace, ncu of both kernels.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; echo "bench rc=$?"; cat gpurun_out/bench_g1.json; tail -3 gpurun_out/bench_g1.err
timeout 300 python bench.py --steps 50 --warmup 5 --log2n 20 > gpurun_out/bench_small_g1.json 2>/dev/null; cat gpurun_out/bench_small_g1.json
timeout 600 python tools/sweep_n.py > gpurun_out/sweep_n_v5.jsonl 2>&1; grep '"eval": "auto"' gpurun_out/sweep_n_v5.jsonl
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v1.jsonl 2>&1; cat gpurun_out/bench_scan_v1.jsonl
timeout 120 python tools/trace_quad.py 32 > gpurun_out/trace32_v5.json 2>&1; cat gpurun_out/trace32_v5.json
timeout 600 python tools/sweep_quad.py 32 > gpurun_out/sweep32_v5.jsonl 2>&1; grep -v '"rows": 16' gpurun_out/sweep32_v5.jsonl | grep rotate
timeout 600 ncu --set full --clock-control none --import-source on -k regex:quad_kernel -s 3 -c 1 -o gpurun_out/prof_quad_v5 python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_quad.log 2>&1; echo "ncu quad rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 2 -c 1 -o gpurun_out/prof_scan_v1 python tools/bench_scan.py > gpurun_out/ncu_scan.log 2>&1; echo "ncu scan rc=$?"
ls -la gpurun_out
