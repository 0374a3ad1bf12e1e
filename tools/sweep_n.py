"""Latency / throughput of integrate() over N on one GPU: device ms (events), kernel us (globaltimer)."""
import json
import math
import sys

sys.path.insert(0, ".")
import b200quad as bq
from b200quad import parallel

ctx = parallel.context(0)
for ev in ("auto", "direct", "rotate"):
    for log2n in list(range(10, 37, 2)):
        if ev == "direct" and log2n > 32:
            continue
        n = 1 << log2n
        job = bq.make_job("sin", 0.0, math.pi, n, "left", "fp64", ev)
        best, kbest = 1e9, 1e9
        for _ in range(6):
            r = ctx.integrate(job)
            best = min(best, float(r.device_ms))
            kbest = min(kbest, (r.t_end_ns - r.t_start_ns) / 1e3)
        ref = bq.utils.oracle_sin(0.0, math.pi, n, "left")
        print(json.dumps({"eval": ev, "used": {1: "direct", 2: "rotate"}[r.eval_used], "log2n": log2n,
                          "device_us": round(best * 1e3, 2), "kernel_us": round(kbest, 2),
                          "g_per_s": round(n / best / 1e6, 1),
                          "rel_err": abs(r.value - ref) / abs(ref)}), flush=True)
