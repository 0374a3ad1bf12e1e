"""NVLS (NVSwitch multicast, multimem.st) against per-peer stores, measured instead of argued:
a single-process world of N GPUs, same kernels, the publish path switched by bq_world_set_nvls.

  (1) small-n integrate(): device us per call, p50 / p99 of 200 calls, n = 2^10 .. 2^20
      (the exchange of the 8 partial sums is the whole cost at these sizes)
  (2) replicated cumulative_integrate(): ms and GB/s received per GPU for the fused all-gather,
      tables in plain device memory (peer stores), in symmetric memory (peer stores) and in
      symmetric memory through the multicast address (multimem.st)

  python tools/nvls_ab.py 2        (or 4, 8)
"""
import json
import math
import sys

sys.path.insert(0, ".")
import torch
import b200quad as bq
from b200quad import parallel, _native
import ctypes as C

world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
lw = parallel.local_world(world)
has_mc = lw.has_multicast
print(json.dumps({"world": world, "multicast_available": bool(has_mc)}), flush=True)


def pct(v, p):
    v = sorted(v)
    return v[min(len(v) - 1, int(p * len(v)))]


modes = [("peer-stores", False)] + ([("nvls-multimem", True)] if has_mc else [])
for log2n in (10, 14, 18, 20):
    job = bq.make_job("sin", 0.0, math.pi, 1 << log2n)
    row = {"bench": "integrate small n", "log2n": log2n, "world": world}
    for name, on in modes:
        lw.set_nvls(on) if has_mc else None
        for _ in range(20):
            lw.integrate(job)
        dev_us, kern_us = [], []
        for _ in range(200):
            res = lw.integrate(job)
            dev_us.append(max(float(r.device_ms) for r in res) * 1e3)
            kern_us.append(max((int(r.t_end_ns) - int(r.t_start_ns)) for r in res) / 1e3)
        row[name] = {"event_us_p50": round(pct(dev_us, 0.5), 2), "event_us_p99": round(pct(dev_us, 0.99), 2),
                     "kernel_us_p50": round(pct(kern_us, 0.5), 2), "kernel_us_p99": round(pct(kern_us, 0.99), 2),
                     "value": res[0].value}
    print(json.dumps(row), flush=True)

for n in (18_000_000, 1 << 26):
    job = _native.ScanJob(bq.models.REGISTRY["profile"].native_id, 1, 0, 0, 0.0, 1800.0 / n, n, (C.c_double * 4)(0, 0, 0, 0))
    row = {"bench": "cumulative replicate", "n": n, "world": world}
    variants = [("device-memory tables, peer stores", False, False), ("symmetric tables, peer stores", True, False)]
    if has_mc:
        variants.append(("symmetric tables, multimem.st", True, True))
    ref_sum = None
    for name, symm, on in variants:
        if has_mc:
            lw.set_nvls(on)
        if symm:
            Ps, Qs = lw.alloc_tables(n)
        else:
            tP = [torch.empty(n, dtype=torch.float64, device="cuda:%d" % d) for d in lw.devices]
            tQ = [torch.empty(n, dtype=torch.float64, device="cuda:%d" % d) for d in lw.devices]
            Ps, Qs = [t.data_ptr() for t in tP], [t.data_ptr() for t in tQ]
        best = 1e9
        for it in range(6):
            out = lw.cumulative(job, Ps, Qs)
            if it >= 1:
                best = min(best, max(float(o.device_ms) for o in out))
        # every rank must hold the whole table: checksum of the LAST rank's copy of P
        last = parallel.device_tensor(Ps[-1], n, lw.devices[-1]) if symm else tP[-1]
        chk = float(last.sum())
        if ref_sum is None:
            ref_sum = chk
        recv = 16.0 * n * (world - 1) / world          # bytes each GPU receives from its peers
        row[name] = {"ms": round(best, 4), "GBps_received_per_gpu": round(recv / best / 1e6, 1),
                     "tables_equal_to_first_variant": chk == ref_sum, "status": [int(o.status) for o in out]}
        if symm:
            del last
            lw.free_tables()
    print(json.dumps(row), flush=True)
