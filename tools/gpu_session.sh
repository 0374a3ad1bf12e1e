#!/bin/bash
# One measurement session on a B200 box (run under gpurun): usage  gpu_session.sh [N_GPUS]
#   tests, headline bench at 1..N GPUs (+ NCCL comparator and small-n at N), scan benches,
#   launch-bound loop, N sweep, per-CTA trace.  Everything lands in gpurun_out/.
N=${1:-1}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$N.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_$N.log
PORT=29600
run() { # G name args...
  G=$1; name=$2; shift; shift
  if [ $G -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $G "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
    PORT=$((PORT+1))
  fi
  echo "$name rc=$?"; cut -c1-700 gpurun_out/$name.json
}
for G in 1 2 4 8; do
  [ $G -le $N ] && run $G bench_g$G --steps 20 --warmup 3
done
if [ $N -gt 1 ]; then
  run $N bench_nccl_g$N --steps 20 --warmup 3 --combine nccl
  run $N bench_small_g$N --steps 50 --warmup 5 --log2n 20
  run $N bench_small_nccl_g$N --steps 50 --warmup 5 --log2n 20 --combine nccl
  timeout 300 python tools/bench_scan_world.py $N > gpurun_out/bench_scan_world$N.jsonl 2>&1; cat gpurun_out/bench_scan_world$N.jsonl
else
  timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-300 gpurun_out/bench_scan.jsonl
  timeout 300 python tools/bench_graph.py 2>/dev/null | tail -1 | tee gpurun_out/bench_graph.json
  timeout 600 python tools/sweep_n.py > gpurun_out/sweep_n.jsonl 2>&1; grep '"eval": "auto"' gpurun_out/sweep_n.jsonl
  timeout 120 python tools/trace_quad.py 32 > gpurun_out/trace32.json 2>&1; cat gpurun_out/trace32.json
fi
