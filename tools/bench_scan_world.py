"""cumulative_integrate() on N GPUs driven by one process (LocalWorld): device ms as the max
over ranks, with and without the fused all-gather (replicate = the reference's MPI_Bcast).

  python tools/bench_scan_world.py 2
"""
import json
import sys

sys.path.insert(0, ".")
import b200quad as bq

world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for n in (18_000_000, 1 << 28):
    for replicate in (False, True):
        best = 1e9
        for it in range(5):
            r = bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, world=world, replicate=replicate,
                                        backend="cuda")
            if it >= 1:
                best = min(best, r.device_ms)
        out_bytes = 16.0 * n
        print(json.dumps({"bench": "cumulative", "world": world, "n": n, "replicate": replicate,
                          "device_ms_max_over_ranks": round(best, 4),
                          "gsamples_per_s": round(n / best / 1e6, 1),
                          "table_GBps_aggregate": round(out_bytes * (world if replicate else 1) / best / 1e6, 1),
                          "p_last_dx": r.p_last * 1800.0 / n}), flush=True)
        del r
