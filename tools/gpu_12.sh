#!/bin/bash
# multi-GPU validation: usage gpu_12.sh N
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$N.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_$N.log
run() { # name, extra args...
  name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $G "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; cat gpurun_out/$name.json; PORT=$((PORT+1))
}
PORT=29600
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; cat gpurun_out/bench_g1.json
for G in 2 4 8; do
  if [ $G -le $N ]; then
    run bench_g$G --steps 20 --warmup 3
    run bench_nccl_g$G --steps 20 --warmup 3 --combine nccl
    run bench_small_g$G --steps 50 --warmup 5 --log2n 20
    run bench_small_nccl_g$G --steps 50 --warmup 5 --log2n 20 --combine nccl
  fi
done
G=$N
run bench_2p36_fp32_trap_g$G --steps 5 --warmup 3 --log2n 36 --dtype fp32 --rule trap
run bench_2p34_mid_g$G --steps 5 --warmup 3 --log2n 34 --rule mid
timeout 120 ./bin/riemann --json --world $N; timeout 120 ./bin/train4 --json --world $N
