#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/sweep_n.py > gpurun_out/sweep_n_v3.jsonl 2>&1; grep '"eval": "auto"' gpurun_out/sweep_n_v3.jsonl
timeout 600 python tools/sweep_quad.py 32 > gpurun_out/sweep32_v3.jsonl 2>&1; grep -v '"rows": 16' gpurun_out/sweep32_v3.jsonl | grep rotate
