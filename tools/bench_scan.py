"""Device time of cumulative_integrate() on one GPU: ms (events), kernel us (globaltimer),
output GB/s against the measured HBM copy peak (MEASURED_PEAKS.json)."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, ".")
import torch
import b200quad as bq
from b200quad import parallel, _native, models

peaks = {}
if os.path.exists("MEASURED_PEAKS.json"):
    peaks = json.load(open("MEASURED_PEAKS.json"))
hbm = float(peaks.get("hbm_gbs", 6551.4))

ctx = parallel.context(0)
sampler = bq.utils.ClockSampler(0, 2).start()
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, n, dx in (("profile", 18_000_000, 1e-4), ("profile", 1 << 26, 1800.0 / (1 << 26)),
                    ("profile", 1 << 28, 1800.0 / (1 << 28)), ("sin", 1 << 26, 3.0 / (1 << 26))):
    integ, _ = models.get(name)
    job = _native.ScanJob(integ.native_id, 0, 0, 0, 0.0, dx, n, (C.c_double * 4)(0.0, 0.0, 0.0, 0.0))
    P = torch.empty(n, dtype=torch.float64, device=dev)
    Q = torch.empty(n, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    best, kbest = 1e9, 1e9
    for it in range(8):
        flush.zero_()
        torch.cuda.synchronize()
        r = ctx.cumulative(job, P.data_ptr(), Q.data_ptr())
        if it >= 2:
            best = min(best, float(r.device_ms))
            k = (r.t_end_ns - r.t_start_ns) / 1e3
            kbest = min(kbest, k)      # the single-pass kernel has no phases to stamp: start and end only
    gbs = 16.0 * n / (best * 1e-3) / 1e9
    print(json.dumps({"bench": "cumulative", "integrand": name, "n": n, "device_ms": round(best, 4),
                      "kernel_us": round(kbest, 1), "out_GBps": round(gbs, 1),
                      "frac_of_hbm_copy_peak": round(gbs / hbm, 3),
                      "gsamples_per_s": round(n / best / 1e6, 2),
                      "p_last_dx": float(r.p_last) * dx}), flush=True)
    del P, Q
print(json.dumps({"clocks": sampler.stop()}))
