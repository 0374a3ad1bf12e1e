"""integrate() of the table integrand (DIRECT mode, the cintegrate.cu workload) and of sin in
DIRECT mode on one GPU: device us, G intervals/s."""
import json
import math
import sys

sys.path.insert(0, ".")
import b200quad as bq
from b200quad import parallel

ctx = parallel.context(0)
for name, a, b, ev, dt in (("profile", 0.0, 1800.0, "auto", "fp64"), ("profile", 0.0, 1800.0, "auto", "fp32"),
                           ("dis", 0.0, 1800.0, "rotate", "fp64"), ("vel", 0.0, 1800.0, "rotate", "fp64")):
    for n in (18_000_000, 1 << 28, 1 << 32):
        job = bq.make_job(name, a, b, n, "left", dt, ev)
        best, kbest = 1e9, 1e9
        for _ in range(5):
            r = ctx.integrate(job)
            best = min(best, float(r.device_ms))
            kbest = min(kbest, (r.t_end_ns - r.t_start_ns) / 1e3)
        print(json.dumps({"integrand": name, "dtype": dt, "eval": {1: "direct", 2: "rotate"}[r.eval_used], "n": n,
                          "device_us": round(best * 1e3, 2), "kernel_us": round(kbest, 2),
                          "g_per_s": round(n / best / 1e6, 1), "value": r.value}), flush=True)
