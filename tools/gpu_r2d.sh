#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_1.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_1.log
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-130 gpurun_out/bench_scan.jsonl
timeout 300 python tools/direct_rates.py 32 > gpurun_out/direct_rates.jsonl 2> gpurun_out/direct_rates.err; cut -c1-150 gpurun_out/direct_rates.jsonl
B200QUAD_LIB=$PWD/build/variants/unroll1/libb200quad.so timeout 300 python tools/direct_rates.py 32 > gpurun_out/direct_rates_unroll1.jsonl 2> gpurun_out/direct_rates_unroll1.err; echo "== unroll1"; grep '"direct"' gpurun_out/direct_rates_unroll1.jsonl | cut -c1-120
