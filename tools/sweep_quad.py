"""Launch-shape sweep of the fused quad kernel on one GPU (device ms, best of 5)."""
import json
import math
import sys

sys.path.insert(0, ".")
import b200quad as bq
from b200quad import parallel

ctx = parallel.context(0)
log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
n = 1 << log2n
rows_list = [0, 16]
for ev in ("rotate", "direct"):
    for dtype in ("fp64", "fp32"):
        for ctas in (1, 2, 3):
            for cluster in (1, 2):
                for rows in (rows_list if ev == "rotate" else [0]):
                    if ev == "direct" and dtype == "fp32" and ctas != 2:
                        continue
                    ctx.configure(ctas, cluster, rows)
                    job = bq.make_job("sin", 0.0, math.pi, n, "left", dtype, ev)
                    best = 1e9
                    for _ in range(4):
                        r = ctx.integrate(job)
                        best = min(best, float(r.device_ms))
                    print(json.dumps({"eval": ev, "dtype": dtype, "ctas_per_sm": ctas, "cluster": cluster,
                                      "rows": rows, "ms": round(best, 4),
                                      "g_per_s": round(n / best / 1e6, 1), "kernel_us": (r.t_end_ns - r.t_start_ns) / 1e3,
                                      "value": r.value}), flush=True)
ctx.configure(0, 0, 0)
