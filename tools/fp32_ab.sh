# fp32 ROTATE speed and error: the build in the tree ("tree") against other builds of the library under
# build/variants/<name>/ (tools/build_variant.py; e.g. one built from an older checkout).
#   VARIANTS="tree" bash tools/fp32_ab.sh
for v in ${VARIANTS:-tree}; do
  if [ "$v" = tree ]; then unset B200QUAD_LIB; else export B200QUAD_LIB=build/variants/$v/libb200quad.so; fi
  V=$v python - <<'PY'
import math, json, os
import b200quad as bq
from b200quad.utils import oracle_sin
cases = [(0.0, math.pi, 32, "left"), (0.0, math.pi / 2, 32, "mid"), (0.0, math.pi, 28, "left"), (0.0, math.pi, 26, "left"), (0.0, math.pi, 24, "left"), (0.0, math.pi, 20, "left"), (0.0, math.pi, 10, "left")]
for (a, b, lg, rule) in cases:
    n = 1 << lg
    ts, ks = [], []
    for i in range(14):
        r = bq.integrate_result("sin", a, b, n, rule, dtype="fp32", eval="rotate", backend="cuda")
        if i >= 2: ts.append(r.device_ms)
    ts.sort()
    ref = oracle_sin(a, b, n, rule)
    print(json.dumps({"variant": os.environ["V"], "b": round(b, 4), "log2n": lg, "rule": rule, "best_ms": round(ts[0], 4), "median_ms": round(ts[len(ts)//2], 4),
                      "g_per_s_best": round(n / ts[0] / 1e6, 1), "rel_err": abs(r.value - ref) / ref}))
PY
done
