#!/bin/bash
# usage: gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$N.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_$N.log
for g in 1 2 4 8; do
  if [ $g -le $N ]; then
    if [ $g -eq 1 ]; then
      timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; cat gpurun_out/bench_g1.json
    else
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2951$g bench.py --gpus $g --steps 20 --warmup 3 > gpurun_out/bench_g$g.json 2> gpurun_out/bench_g$g.err; cat gpurun_out/bench_g$g.json; tail -2 gpurun_out/bench_g$g.err
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2952$g bench.py --gpus $g --steps 20 --warmup 3 --combine nccl > gpurun_out/bench_nccl_g$g.json 2> gpurun_out/bench_nccl_g$g.err; cat gpurun_out/bench_nccl_g$g.json
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2953$g bench.py --gpus $g --steps 50 --warmup 5 --log2n 20 > gpurun_out/bench_small_g$g.json 2> gpurun_out/bench_small_g$g.err; cat gpurun_out/bench_small_g$g.json
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2954$g bench.py --gpus $g --steps 50 --warmup 5 --log2n 20 --combine nccl > gpurun_out/bench_small_nccl_g$g.json 2> gpurun_out/bench_small_nccl_g$g.err; cat gpurun_out/bench_small_nccl_g$g.json
    fi
  fi
done
timeout 120 ./bin/riemann --json --world $N; timeout 120 ./bin/train4 --json --world $N
