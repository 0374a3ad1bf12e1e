#!/bin/bash
# 8-GPU session: multi-GPU tests, headline at 8, comparator, small-n sweep, the other BASELINE configs.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dist.py tests/test_gpu_scan.py tests/test_custom_integrand.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_gpu_8.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_8.log
timeout 600 python -m pytest tests/test_gpu_quad.py -m gpu -x -q -k "local_world or peer_timeout" >> gpurun_out/pytest_gpu_8.log 2>&1; tail -2 gpurun_out/pytest_gpu_8.log
PORT=29600
run() { # name args...
  name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 8 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "$name rc=$?"; cut -c1-420 gpurun_out/$name.json; PORT=$((PORT+1))
}
run bench_g8 --steps 20 --warmup 3
run bench_nccl_g8 --steps 20 --warmup 3 --combine nccl
run bench_2p36_fp32_trap_g8 --steps 5 --warmup 3 --log2n 36 --dtype fp32 --rule trap
run bench_2p34_mid_g8 --steps 10 --warmup 3 --log2n 34 --rule mid
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 tools/sweep_small_dist.py > gpurun_out/sweep_small_g8.jsonl 2> gpurun_out/sweep_small_g8.err; cat gpurun_out/sweep_small_g8.jsonl
timeout 300 python tools/bench_scan_world.py 8 > gpurun_out/bench_scan_world8.jsonl 2>&1; cat gpurun_out/bench_scan_world8.jsonl
