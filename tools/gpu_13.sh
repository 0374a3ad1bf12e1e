#!/bin/bash
# 8-GPU validation + scaling: tests, fused bench at 1/2/4/8, comparators at 8
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_8.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_8.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_8.log
run() { # G name args...
  G=$1; name=$2; shift; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $G "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; cat gpurun_out/$name.json; PORT=$((PORT+1))
}
PORT=29600
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; cat gpurun_out/bench_g1.json
run 2 bench_g2 --steps 20 --warmup 3
run 4 bench_g4 --steps 20 --warmup 3
run 8 bench_g8 --steps 20 --warmup 3
run 8 bench_nccl_g8 --steps 20 --warmup 3 --combine nccl
run 8 bench_small_g8 --steps 50 --warmup 5 --log2n 20
run 8 bench_small_nccl_g8 --steps 50 --warmup 5 --log2n 20 --combine nccl
run 8 bench_2p36_fp32_trap_g8 --steps 5 --warmup 3 --log2n 36 --dtype fp32 --rule trap
run 8 bench_2p34_mid_g8 --steps 10 --warmup 3 --log2n 34 --rule mid
timeout 120 ./bin/train4 --json --world 8
