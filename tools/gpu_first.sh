#!/bin/bash
# First GPU pass: smoke, GPU tests, fp64 peak, bench (both evaluation modes).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/summary.txt
tail -15 gpurun_out/pytest_gpu.log
timeout 120 bin/peak_fp64 > gpurun_out/peak.log 2>&1; cat gpurun_out/peak.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_rotate.json 2> gpurun_out/bench_rotate.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench_rotate.json
timeout 300 python bench.py --steps 10 --warmup 3 --eval direct > gpurun_out/bench_direct.json 2> gpurun_out/bench_direct.err; cat gpurun_out/bench_direct.json
timeout 300 python bench.py --steps 10 --warmup 3 --dtype fp32 > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err; cat gpurun_out/bench_fp32.json
tail -5 gpurun_out/smoke.log gpurun_out/bench_rotate.err
