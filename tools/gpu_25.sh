#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/bench_graph.py 2>&1 | tail -3 | tee gpurun_out/bench_graph.json
timeout 300 python tools/bench_scan_world.py 2 > gpurun_out/bench_scan_world2.jsonl 2>&1; cat gpurun_out/bench_scan_world2.jsonl
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 3 2>/dev/null > gpurun_out/bench_g2.json; cat gpurun_out/bench_g2.json
