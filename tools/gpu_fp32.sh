#!/bin/bash
# fp32 ROTATE with error-diffused row values: GPU == CPU bits, rates with 16- and 32-row runs, loop microbenchmark.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_quad.py tests/test_gpu_stress.py -m gpu -x -q -k "fp32 or bitwise or cpu or accuracy" 2>&1 | tail -3
for v in "" pr32; do
  if [ -n "$v" ]; then export B200QUAD_LIB=build/variants/$v/libb200quad.so; fi
  echo "== ${v:-default}"
  python - <<'PY'
import math, json
import b200quad as bq
from b200quad.utils import oracle_sin
import torch
for lg, rule in ((24, "left"), (28, "left"), (32, "left"), (32, "trap"), (34, "mid")):
    n = 1 << lg
    best = 1e9
    for i in range(6):
        r = bq.integrate_result("sin", 0.0, math.pi, n, rule, dtype="fp32", eval="rotate", backend="cuda")
        if i >= 2: best = min(best, r.device_ms)
    ref = oracle_sin(0.0, math.pi, n, rule)
    c = bq.integrate_result("sin", 0.0, math.pi, n, rule, dtype="fp32", eval="rotate", backend="cpu") if lg <= 28 else None
    print(json.dumps({"log2n": lg, "rule": rule, "device_ms": round(best, 4), "g_per_s": round(n / best / 1e6, 1),
                      "rel_err": abs(r.value - ref) / ref, "equals_cpu_tree": None if c is None else (c.value == r.value and c.value_lo == r.value_lo)}))
for name, a, b in (("vel", 0.0, 1800.0), ("acc", 5.0, 900.0)):
    n = 1 << 28
    r = bq.integrate_result(name, a, b, n, "mid", dtype="fp32", eval="rotate", backend="cuda")
    d = bq.integrate_result(name, a, b, n, "mid", dtype="fp64", eval="rotate", backend="cuda")
    c = bq.integrate_result(name, a, b, n, "mid", dtype="fp32", eval="rotate", backend="cpu")
    print(json.dumps({"integrand": name, "fp32": r.value, "fp64": d.value, "rel_diff": abs(r.value - d.value) / max(abs(d.value), 1e-300), "equals_cpu_tree": c.value == r.value}))
PY
done > gpurun_out/fp32_rates.jsonl 2>&1
cat gpurun_out/fp32_rates.jsonl
./bin/ffma_bench | grep '"ctas_per_sm": 2' | tee gpurun_out/ffma_bench.jsonl | cut -c1-170
