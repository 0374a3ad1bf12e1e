#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_2_final.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_2_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
