"""Small-n latency sweep under torchrun (BASELINE.json config 4): device us per integrate() call,
n = 2^10 .. 2^24, fused in-kernel combine vs local kernel + NCCL all_reduce of one fp64.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29650 tools/sweep_small_dist.py
"""
import ctypes as C
import json
import math
import os
import sys

sys.path.insert(0, ".")
import torch
import torch.distributed as dist
import b200quad as bq
from b200quad import parallel

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
w = parallel.dist_world()
ctx = w.ctx
stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
d_value = torch.zeros(1, dtype=torch.float64, device=dev)
seg = torch.zeros(16, dtype=torch.float64, device=dev)
out = torch.zeros(1, dtype=torch.float64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
torch.cuda.synchronize(dev)
K = 200


def timed(step):
    for _ in range(5):
        step()
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for e0, e1 in evs:
        with torch.cuda.stream(stream):
            flush.zero_()
            e0.record(stream)
        step()
        e1.record(stream)
    torch.cuda.synchronize(dev)
    per = sorted(a.elapsed_time(b) for a, b in evs)
    # p50 and p99 of this rank's K calls, then the max over ranks of each
    t = torch.tensor([per[len(per) // 2] * 1e3, per[min(len(per) - 1, int(0.99 * len(per)))] * 1e3, per[-1] * 1e3],
                     dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [round(float(x), 2) for x in t.tolist()]


for log2n in range(10, 25, 2):
    n = 1 << log2n
    job = bq.make_job("sin", 0.0, math.pi, n)

    def fused():
        ctx.integrate_device_async(job, d_value.data_ptr())

    def nccl():
        rc = ctx.lib.bq_integrate_local_async(ctx.handle, C.byref(job), rank, world, C.c_void_p(seg.data_ptr()),
                                              C.c_void_p(ctx.stream))
        assert rc == 0
        with torch.cuda.stream(stream):
            torch.sum(seg, dim=0, keepdim=True, out=out)
            dist.all_reduce(out)

    f_us = timed(fused)
    c_us = timed(nccl)
    torch.cuda.synchronize(dev)
    if rank == 0:
        h = math.pi / n
        print(json.dumps({"log2n": log2n, "gpus": world, "calls": K, "fused_us_p50_p99_max": f_us,
                          "nccl_comparator_us_p50_p99_max": c_us,
                          "fused_value": float(d_value), "comparator_value": float(out) * h}), flush=True)
dist.barrier()
dist.destroy_process_group()
