"""Where the fixed per-call time of integrate() goes: per-rank table of the kernel's own %globaltimer
stamps (bq_result.stamp_ns) next to the CUDA-event time of the step, for the bench.py step shape
(L2 flush, event, ONE kernel, event).  Run alone (1 GPU) or under torchrun (N ranks):

  python tools/stamps.py [log2n] [steps]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29660 tools/stamps.py 32 40

Prints one JSON line per rank (medians over the steps, microseconds): event time, kernel span,
and the stamps as intervals: prologue | tiles | ->ticket | fold | publish | barrier wait | result.
"""
import json
import math
import os
import statistics
import sys

sys.path.insert(0, ".")
import torch
import b200quad as bq
from b200quad import parallel, _native

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
    ctx = parallel.dist_world().ctx
else:
    ctx = parallel.context(local)
stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
d_value = torch.zeros(1, dtype=torch.float64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
job = bq.make_job("sin", 0.0, math.pi, 1 << log2n)
rows = []
for it in range(steps + 5):
    if dist is not None and it == 5:
        torch.cuda.synchronize(dev)
        dist.barrier()
        torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        flush.zero_()
        e0.record(stream)
    ctx.integrate_device_async(job, d_value.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize(dev)          # one step at a time: the stamps of THIS step are read back
    r = ctx.fetch_device()
    if it >= 5:
        s = list(r.stamp_ns)
        rows.append([e0.elapsed_time(e1) * 1e3, (r.t_end_ns - r.t_start_ns) / 1e3] + [x / 1e3 for x in s[:7]] + [s[7]])
med = [statistics.median(c) for c in zip(*rows)]
ev, span, pro, tiles, ticket, fold, pub, bar, res, spins = med
line = {"log2n": log2n, "world": world, "rank": rank, "event_us": round(ev, 2), "kernel_span_us": round(span, 2),
        "outside_kernel_us": round(ev - span, 2),
        "prologue_us": round(pro, 2), "tiles_us": round(tiles - pro, 2), "to_ticket_us": round(ticket - tiles, 2),
        "fold_us": round(fold - ticket, 2), "publish_us": round(pub - fold, 2), "barrier_wait_us": round(bar - pub, 2),
        "result_us": round(res - bar, 2), "barrier_polls": spins,
        "note": "the finishing CTA's view; its tile time is the longest of the GPU by construction"}
if dist is not None:
    out = [None] * world
    dist.all_gather_object(out, line)
    if rank == 0:
        for l in out:
            print(json.dumps(l), flush=True)
    dist.barrier()
    dist.destroy_process_group()
else:
    print(json.dumps(line), flush=True)
