#!/bin/bash
# Round-2 session A (2 GPUs): tests, launch floor, stamps at 1 and 2 GPUs, headline at 1 and 2,
# DIRECT rates, scan bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_start.csv 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_2.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_2.log
./bin/launch_floor > gpurun_out/launch_floor.json 2>&1; cat gpurun_out/launch_floor.json
timeout 200 python tools/stamps.py 32 40 > gpurun_out/stamps_g1.jsonl 2> gpurun_out/stamps_g1.err; cat gpurun_out/stamps_g1.jsonl
timeout 200 python tools/stamps.py 20 40 > gpurun_out/stamps_small_g1.jsonl 2> gpurun_out/stamps_small_g1.err; cat gpurun_out/stamps_small_g1.jsonl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29660 tools/stamps.py 32 40 > gpurun_out/stamps_g2.jsonl 2> gpurun_out/stamps_g2.err; cat gpurun_out/stamps_g2.jsonl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29661 tools/stamps.py 20 40 > gpurun_out/stamps_small_g2.jsonl 2> gpurun_out/stamps_small_g2.err; cat gpurun_out/stamps_small_g2.jsonl
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; echo "bench1 rc=$?"; cut -c1-600 gpurun_out/bench_g1.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29662 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "bench2 rc=$?"; cut -c1-600 gpurun_out/bench_g2.json
timeout 300 python tools/direct_rates.py 32 > gpurun_out/direct_rates.jsonl 2> gpurun_out/direct_rates.err; cat gpurun_out/direct_rates.jsonl
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-330 gpurun_out/bench_scan.jsonl
timeout 300 python tools/bench_scan_world.py 2 > gpurun_out/bench_scan_world2.jsonl 2>&1; cat gpurun_out/bench_scan_world2.jsonl
