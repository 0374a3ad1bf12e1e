#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep_quad.py 32 > gpurun_out/sweep32_v2.jsonl 2>&1; grep -v '"rows": 16' gpurun_out/sweep32_v2.jsonl | tail -40
