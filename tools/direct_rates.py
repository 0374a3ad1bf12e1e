"""G intervals/s of the general (DIRECT) evaluation path and of ROTATE for every library integrand
at n = 2^LOG2N on one GPU: device ms (events, best of 5), relative error against the oracle.
DIRECT is the mode that does the reference's per-point work (cintegrate.cu:36-44,66-70): x formed and
f called for every sample.

  python tools/direct_rates.py [log2n]
"""
import json
import math
import sys

sys.path.insert(0, ".")
import b200quad as bq
from b200quad import parallel

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
n = 1 << log2n
ctx = parallel.context(0)
cases = [("sin", 0.0, math.pi, "fp64", "direct"), ("sin", 0.0, math.pi, "fp64", "rotate"),
         ("sin", 0.0, math.pi, "fp32", "direct"), ("sin", 0.0, math.pi, "fp32", "rotate"),
         ("profile", 0.0, 1800.0, "fp64", "direct"), ("profile", 0.0, 1800.0, "fp32", "direct"),
         ("dis", 0.0, 1800.0, "fp64", "direct"), ("dis", 0.0, 1800.0, "fp64", "rotate"),
         ("vel", 0.0, 1800.0, "fp64", "direct"), ("vel", 0.0, 1800.0, "fp64", "rotate"),
         ("acc", 0.0, 1800.0, "fp64", "direct")]
for name, a, b, dt, ev in cases:
    job = bq.make_job(name, a, b, n, "left", dt, ev)
    best = 1e9
    for _ in range(5):
        r = ctx.integrate(job)
        best = min(best, float(r.device_ms))
    ref = None
    if name == "sin":
        ref = bq.utils.oracle_sin(a, b, n, "left")
    elif name == "profile":
        ref = 122000.004 - 0.5 * (b - a) / n * (4e-15 - 0.0)     # piecewise linear, aligned grid: exact left sum
    print(json.dumps({"integrand": name, "dtype": dt, "eval": ev, "log2n": log2n, "device_ms": round(best, 4),
                      "g_per_s": round(n / best / 1e6, 1), "value": r.value,
                      "rel_err": abs(r.value - ref) / abs(ref) if ref else None}), flush=True)
