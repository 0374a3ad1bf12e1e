#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python tools/bench_scan.py > gpurun_out/bench_scan_v5.jsonl 2>&1; cut -c1-330 gpurun_out/bench_scan_v5.jsonl
for v in u2 main u8 u16; do
  if [ $v = main ]; then unset B200QUAD_LIB; else export B200QUAD_LIB=$PWD/build/variants/$v/libb200quad.so; fi
  echo "== unroll variant $v"
  timeout 300 python - <<'PY'
import sys, math, json
sys.path.insert(0, ".")
import b200quad as bq
from b200quad import parallel
ctx = parallel.context(0)
for dt in ("fp64", "fp32"):
    for log2n in (29, 32, 34):
        job = bq.make_job("sin", 0.0, math.pi, 1 << log2n, "left", dt, "rotate")
        best = 1e9
        for _ in range(6):
            r = ctx.integrate(job)
            best = min(best, float(r.device_ms))
        print(json.dumps({"dtype": dt, "log2n": log2n, "device_us": round(best * 1e3, 1), "g_per_s": round((1 << log2n) / best / 1e6, 1)}))
PY
done
