"""Per-CTA timeline of one integrate() call: spread of start/end times, pairing on SMs."""
import collections
import json
import math
import sys

sys.path.insert(0, ".")
import b200quad as bq
from b200quad import parallel

log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ctx = parallel.context(0)
ctx.trace(True)
job = bq.make_job("sin", 0.0, math.pi, 1 << log2n)
for _ in range(3):
    r = ctx.integrate(job)
tr = ctx.read_trace()
t0 = min(t[0] for t in tr)
starts = sorted(t[0] - t0 for t in tr)
ends = sorted(t[1] - t0 for t in tr)
durs = sorted(t[1] - t[0] for t in tr)
per_sm = collections.Counter(t[2] for t in tr)
q = lambda v, f: v[min(len(v) - 1, int(f * len(v)))]
print(json.dumps({"log2n": log2n, "ctas": len(tr), "kernel_us": (r.t_end_ns - r.t_start_ns) / 1e3,
                  "start_us": [q(starts, f) / 1e3 for f in (0, .5, .9, 1)],
                  "end_us": [q(ends, f) / 1e3 for f in (0, .1, .5, .9, 1)],
                  "dur_us": [q(durs, f) / 1e3 for f in (0, .1, .5, .9, 1)],
                  "ctas_per_sm_hist": dict(collections.Counter(per_sm.values())),
                  "tiles": [min(t[3] for t in tr), max(t[3] for t in tr)]}))
# rate per CTA: tiles / dur
rates = sorted((t[3] / max(1, (t[1] - t[0]))) for t in tr)
print(json.dumps({"tiles_per_us": [q(rates, f) * 1e3 for f in (0, .1, .5, .9, 1)]}))
ctx.trace(False)
