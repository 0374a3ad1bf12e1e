#!/bin/bash
# Round-2 session B (2 GPUs): tests on the single-pass scan, scan benches, DIRECT variants, stamps.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_2.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_2.log
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-330 gpurun_out/bench_scan.jsonl
timeout 300 python tools/bench_scan_world.py 2 > gpurun_out/bench_scan_world2.jsonl 2>&1; cat gpurun_out/bench_scan_world2.jsonl
timeout 300 python tools/direct_rates.py 32 > gpurun_out/direct_rates.jsonl 2> gpurun_out/direct_rates.err; grep direct gpurun_out/direct_rates.jsonl
B200QUAD_LIB=$PWD/build/variants/magic/libb200quad.so timeout 300 python tools/direct_rates.py 32 > gpurun_out/direct_rates_magic.jsonl 2> gpurun_out/direct_rates_magic.err; grep direct gpurun_out/direct_rates_magic.jsonl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29660 tools/stamps.py 32 40 > gpurun_out/stamps_g2.jsonl 2> gpurun_out/stamps_g2.err; cat gpurun_out/stamps_g2.jsonl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29662 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "bench2 rc=$?"; cut -c1-300 gpurun_out/bench_g2.json
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; echo "bench1 rc=$?"; cut -c1-2500 gpurun_out/bench_g1.json; tail -3 gpurun_out/bench_g1.err
timeout 300 python bench.py --workload scan --gpus 1 --steps 10 --warmup 3 --n 268435456 > gpurun_out/bench_scan_2p28_g1.json 2> gpurun_out/bench_scan_2p28_g1.err; cut -c1-900 gpurun_out/bench_scan_2p28_g1.json
timeout 600 python tools/cli_wall.py > gpurun_out/cli_wall.jsonl 2> gpurun_out/cli_wall.err; cat gpurun_out/cli_wall.jsonl
