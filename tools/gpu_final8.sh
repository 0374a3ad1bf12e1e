#!/bin/bash
# Final 8-GPU record: headline at 8 and 4 GPUs, NCCL comparator at 8.  Output: gpurun_out/f8_*.
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
timeout 200 $T --nproc-per-node 8 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/f8_bench8.json 2> gpurun_out/f8_bench8.err; echo "bench8 rc=$?"
timeout 150 $T --nproc-per-node 4 bench.py --gpus 4 --steps 20 --warmup 5 --no-extras > gpurun_out/f8_bench4.json 2> gpurun_out/f8_bench4.err; echo "bench4 rc=$?"
timeout 150 $T --nproc-per-node 8 bench.py --gpus 8 --steps 20 --warmup 5 --combine nccl --no-extras > gpurun_out/f8_nccl8.json 2> gpurun_out/f8_nccl8.err; echo "nccl8 rc=$?"
python - <<'PY'
import json
for f in ("f8_bench8", "f8_bench4", "f8_nccl8"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f))
        print(f, d["value"], d["ms_per_step"], d.get("e2e", {}).get("value"), d.get("kernel_stamps_us_rank0_last_step"), d.get("clocks"))
    except Exception as e:
        print(f, "unreadable", e)
PY
