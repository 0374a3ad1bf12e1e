#!/bin/bash
# Round-2 session F (2 GPUs): full tests incl. stress, scan with 2 rounds of lookahead, headline bench.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu_2.log 2>&1; echo "pytest rc=$?"; tail -22 gpurun_out/pytest_gpu_2.log
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan.jsonl 2>&1; cut -c1-130 gpurun_out/bench_scan.jsonl
timeout 300 python tools/bench_scan_world.py 2 > gpurun_out/bench_scan_world2.jsonl 2>&1; cat gpurun_out/bench_scan_world2.jsonl
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/bench_g1.json 2> gpurun_out/bench_g1.err; echo "bench1 rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_g1.json')); print(d['value'], d['e2e']['value'], json.dumps(d['extras'])[:1500])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29662 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "bench2 rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_g2.json')); print(d['value'], d['e2e']['value'], json.dumps(d['extras'])[:1500])"
