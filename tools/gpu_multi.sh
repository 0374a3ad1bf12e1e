#!/bin/bash
# Multi-GPU session (N = 2, 4 or 8): usage gpu_multi.sh N.  Everything lands in gpurun_out/gN_*.
N=${1:-8}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/g${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/g${N}_pytest.log
PORT=29700
for G in 2 4 8; do
  [ $G -le $N ] || continue
  timeout 300 $T --nproc-per-node $G --master-port $PORT bench.py --gpus $G --steps 20 --warmup 3 > gpurun_out/g${N}_bench_g$G.json 2> gpurun_out/g${N}_bench_g$G.err; echo "bench g$G rc=$?"; PORT=$((PORT+1))
  python -c "
import json; d=json.load(open('gpurun_out/g${N}_bench_g$G.json')); print(d['n_gpus'], round(d['value'],1), round(d['e2e']['value'],1), d['ms_per_step'], d['kernel_stamps_us_rank0_last_step'], d['clocks'])"
  timeout 300 $T --nproc-per-node $G --master-port $PORT tools/sweep_small_dist.py > gpurun_out/g${N}_sweep_small_g$G.jsonl 2> gpurun_out/g${N}_sweep_small_g$G.err; PORT=$((PORT+1)); cat gpurun_out/g${N}_sweep_small_g$G.jsonl | cut -c1-260
done
timeout 300 $T --nproc-per-node $N --master-port $PORT tools/stamps.py 32 40 > gpurun_out/g${N}_stamps.jsonl 2> gpurun_out/g${N}_stamps.err; PORT=$((PORT+1)); cut -c1-420 gpurun_out/g${N}_stamps.jsonl
BQ_SYMM=1 BQ_NVLS=1 timeout 300 $T --nproc-per-node $N --master-port $PORT tools/stamps.py 20 40 > gpurun_out/g${N}_stamps_small_symm_nvls.jsonl 2> gpurun_out/g${N}_stamps_small_symm_nvls.err; PORT=$((PORT+1)); cut -c1-300 gpurun_out/g${N}_stamps_small_symm_nvls.jsonl | head -3
timeout 300 $T --nproc-per-node $N --master-port $PORT tools/stamps.py 20 40 > gpurun_out/g${N}_stamps_small.jsonl 2> gpurun_out/g${N}_stamps_small.err; PORT=$((PORT+1)); cut -c1-300 gpurun_out/g${N}_stamps_small.jsonl | head -3
timeout 300 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --steps 20 --warmup 3 --combine nccl > gpurun_out/g${N}_bench_nccl.json 2> gpurun_out/g${N}_bench_nccl.err; PORT=$((PORT+1)); cut -c1-300 gpurun_out/g${N}_bench_nccl.json
timeout 300 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --steps 10 --warmup 3 --log2n 34 --rule mid --no-extras > gpurun_out/g${N}_bench_2p34_mid.json 2> gpurun_out/g${N}_bench_2p34_mid.err; PORT=$((PORT+1)); cut -c1-220 gpurun_out/g${N}_bench_2p34_mid.json
timeout 300 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --steps 5 --warmup 3 --log2n 36 --dtype fp32 --rule trap --no-extras > gpurun_out/g${N}_bench_2p36_fp32_trap.json 2> gpurun_out/g${N}_bench_2p36_fp32_trap.err; PORT=$((PORT+1)); cut -c1-220 gpurun_out/g${N}_bench_2p36_fp32_trap.json; grep -o '"rel_err_vs_oracle": [^,]*' gpurun_out/g${N}_bench_2p36_fp32_trap.json
timeout 300 $T --nproc-per-node $N --master-port $PORT bench.py --gpus $N --workload scan --steps 10 --warmup 3 > gpurun_out/g${N}_bench_scan_18e6.json 2> gpurun_out/g${N}_bench_scan_18e6.err; PORT=$((PORT+1)); cut -c1-200 gpurun_out/g${N}_bench_scan_18e6.json
timeout 400 python tools/nvls_ab.py $N > gpurun_out/g${N}_nvls_ab.jsonl 2> gpurun_out/g${N}_nvls_ab.err; cat gpurun_out/g${N}_nvls_ab.jsonl | cut -c1-700; tail -2 gpurun_out/g${N}_nvls_ab.err
timeout 400 python tools/bench_scan_world.py $N > gpurun_out/g${N}_bench_scan_world.jsonl 2>&1; cat gpurun_out/g${N}_bench_scan_world.jsonl
timeout 300 python tools/cli_wall.py > gpurun_out/g${N}_cli_wall.jsonl 2> gpurun_out/g${N}_cli_wall.err; cut -c1-150 gpurun_out/g${N}_cli_wall.jsonl
nvidia-smi topo -m > gpurun_out/g${N}_topo.txt 2>&1
