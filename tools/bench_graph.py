"""Launch-bound regime: 64 integrals of n = 2^16 back to back, (a) one public-API call each,
(b) enqueued asynchronously on one stream, (c) captured once in a CUDA graph and replayed."""
import json
import math
import sys
import time

sys.path.insert(0, ".")
import torch
import b200quad as bq
from b200quad import parallel
from b200quad.ops import torch_ops

K, n = 64, 1 << 16
ctx = parallel.context(0)
job = bq.make_job("sin", 0.0, math.pi, n, "mid")
for _ in range(5):
    bq.integrate("sin", 0.0, math.pi, n, "mid")
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(K):
    bq.integrate("sin", 0.0, math.pi, n, "mid")
sync_us = (time.perf_counter() - t0) / K * 1e6

stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best_async = 1e9
for _ in range(5):
    e0.record(stream)
    for _ in range(K):
        ctx.integrate_async(job)
    e1.record(stream)
    torch.cuda.synchronize()
    best_async = min(best_async, e0.elapsed_time(e1) * 1e3 / K)

torch_ops.integrate("sin", 0.0, math.pi, n, "mid", device=0)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    outs = [torch_ops.integrate("sin", 0.0, math.pi, n, "mid", device=0) for _ in range(K)]
best_graph = 1e9
for _ in range(5):
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    best_graph = min(best_graph, e0.elapsed_time(e1) * 1e3 / K)
print(json.dumps({"bench": "launch-bound loop", "n": n, "calls": K,
                  "us_per_call_public_api_sync": round(sync_us, 2),
                  "us_per_call_async_stream_device": round(best_async, 2),
                  "us_per_call_cuda_graph_device": round(best_graph, 2),
                  "value": float(outs[-1])}))
