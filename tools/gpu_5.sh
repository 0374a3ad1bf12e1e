#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/pytest_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "all rc=$?"; tail -5 gpurun_out/pytest_gpu.log
