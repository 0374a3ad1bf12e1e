#!/bin/bash
# Slim multi-GPU session (N = 4 or 8), sized for a small GPU-minute budget: multi-GPU tests, headline at
# N/2 and N, stamps, small-n p50/p99, comparator, scan, NVLS A/B.  Output: gpurun_out/sN_*.
N=${1:-8}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 python -m pytest tests/test_gpu_dist.py tests/test_gpu_stress.py tests/test_torch_ops.py tests/test_zz_no_alloc.py -m gpu -x -q -k "not 2p32" > gpurun_out/s${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s${N}_pytest.log
PORT=29700
for G in $((N/2)) $N; do
  timeout 200 $T --nproc-per-node $G --master-port $PORT bench.py --gpus $G --steps 20 --warmup 3 > gpurun_out/s${N}_bench_g$G.json 2> gpurun_out/s${N}_bench_g$G.err; echo "bench g$G rc=$?"; PORT=$((PORT+1))
  python -c "
import json; d=json.load(open('gpurun_out/s${N}_bench_g$G.json')); print(d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), d['ms_per_step'], d['kernel_stamps_us_rank0_last_step'], d['clocks']); print(json.dumps(d['extras'])[:900])"
  timeout 200 $T --nproc-per-node $G --master-port $PORT tools/sweep_small_dist.py > gpurun_out/s${N}_sweep_small_g$G.jsonl 2> gpurun_out/s${N}_sweep_small_g$G.err; PORT=$((PORT+1)); cut -c1-230 gpurun_out/s${N}_sweep_small_g$G.jsonl
done
timeout 200 $T --nproc-per-node $N --master-port $PORT tools/stamps.py 32 40 > gpurun_out/s${N}_stamps.jsonl 2> gpurun_out/s${N}_stamps.err; PORT=$((PORT+1)); cut -c1-420 gpurun_out/s${N}_stamps.jsonl
timeout 200 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --steps 20 --warmup 3 --combine nccl > gpurun_out/s${N}_bench_nccl.json 2> gpurun_out/s${N}_bench_nccl.err; PORT=$((PORT+1)); cut -c1-260 gpurun_out/s${N}_bench_nccl.json
timeout 200 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --steps 5 --warmup 3 --log2n 36 --dtype fp32 --rule trap --no-extras > gpurun_out/s${N}_bench_2p36_fp32_trap.json 2> gpurun_out/s${N}_bench_2p36_fp32_trap.err; PORT=$((PORT+1)); cut -c1-200 gpurun_out/s${N}_bench_2p36_fp32_trap.json; grep -o '"rel_err_vs_oracle": [^,]*' gpurun_out/s${N}_bench_2p36_fp32_trap.json
timeout 200 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --steps 10 --warmup 3 --log2n 34 --rule mid --no-extras > gpurun_out/s${N}_bench_2p34_mid.json 2> gpurun_out/s${N}_bench_2p34_mid.err; PORT=$((PORT+1)); cut -c1-200 gpurun_out/s${N}_bench_2p34_mid.json
timeout 300 python tools/nvls_ab.py $N > gpurun_out/s${N}_nvls_ab.jsonl 2> gpurun_out/s${N}_nvls_ab.err; cut -c1-900 gpurun_out/s${N}_nvls_ab.jsonl; tail -2 gpurun_out/s${N}_nvls_ab.err
timeout 300 python tools/bench_scan_world.py $N > gpurun_out/s${N}_bench_scan_world.jsonl 2>&1; cat gpurun_out/s${N}_bench_scan_world.jsonl
