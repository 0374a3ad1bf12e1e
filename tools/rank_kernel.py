"""What one rank of an 8-GPU integrate(sin, 2^32) does, measured alone on one GPU: the kernel with
its cross-GPU stage switched off (rank r of world w owns 8/w of the 8 fixed segments).  Separates
the per-rank compute + in-kernel fixed cost from the exchange / arrival skew of a real 8-GPU step."""
import ctypes as C
import json
import math
import sys

sys.path.insert(0, ".")
import torch
import b200quad as bq
from b200quad import parallel

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ctx = parallel.context(0)
dev = torch.device("cuda", 0)
stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
seg = torch.zeros(16, dtype=torch.float64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
job = bq.make_job("sin", 0.0, math.pi, 1 << log2n)
one = None
for world in (1, 2, 4, 8):
    best_dev, best_k = 1e9, 1e9
    for it in range(8):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            flush.zero_()
            e0.record(stream)
        rc = ctx.lib.bq_integrate_local_async(ctx.handle, C.byref(job), 0, world, C.c_void_p(seg.data_ptr()),
                                              C.c_void_p(ctx.stream))
        assert rc == 0
        e1.record(stream)
        torch.cuda.synchronize()
        r = ctx.fetch_device()
        if it >= 2:
            best_dev = min(best_dev, e0.elapsed_time(e1) * 1e3)
            best_k = min(best_k, (r.t_end_ns - r.t_start_ns) / 1e3)
    if world == 1:
        one = best_k
    print(json.dumps({"log2n": log2n, "rank": 0, "world": world, "device_us": round(best_dev, 1),
                      "kernel_us": round(best_k, 1), "ideal_us_from_world1": round(one / world, 1),
                      "fixed_cost_us": round(best_k - one / world, 1)}), flush=True)
