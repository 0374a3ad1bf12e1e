"""One small program per kernel for ncu (B200_PROFILING.md: profile the smallest case that shows it;
run the same command without ncu first).  3 launches of one kernel, nothing else.

  python tools/ncu_target.py scan 26        cumulative_integrate(profile, n = 2^26)
  python tools/ncu_target.py direct_sin 30  integrate(sin, DIRECT, fp64)
  python tools/ncu_target.py direct_profile 30 | direct_custom 28 | rotate_sin 32 | rotate_sin_fp32 32
"""
import math
import sys

sys.path.insert(0, ".")
import torch
import b200quad as bq
from b200quad import parallel

what = sys.argv[1]
log2n = int(sys.argv[2]) if len(sys.argv) > 2 else 28
n = 1 << log2n
ctx = parallel.context(0)
if what == "scan":
    for _ in range(3):
        r = bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, backend="cuda")
    print("scan", n, r.device_ms, r.p_last * 1800.0 / n)
else:
    if what == "direct_custom":
        f, a, b, ev, dt = bq.models.custom("exp(-x*x) * p0", params=(2.0,)), 0.0, 3.0, "direct", "fp64"
    else:
        ev, name = what.split("_", 1)
        dt = "fp32" if name.endswith("_fp32") else "fp64"
        f = name.replace("_fp32", "")
        a, b = (0.0, math.pi) if f == "sin" else (0.0, 1800.0)
    job = bq.make_job(f, a, b, n, "left", dt, ev)
    for _ in range(3):
        r = ctx.integrate(job)
    print(what, n, r.device_ms, r.value)
