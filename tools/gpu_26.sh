#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_graph.py tests/test_torch_ops.py tests/test_gpu_quad.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/bench_graph.py 2>&1 | tail -1 | tee gpurun_out/bench_graph.json
timeout 120 python tools/trace_quad.py 16 2>&1 | tail -2
timeout 120 python tools/trace_quad.py 29 2>&1 | tail -2
timeout 300 python tools/sweep_quad.py 16 2>&1 | grep -v '"rows": 16' | grep '"fp64"' | grep -v direct
