#!/bin/bash
mkdir -p gpurun_out
for v in base idx idxwt cell all; do
  if [ $v = all ]; then unset B200QUAD_LIB; else export B200QUAD_LIB=$PWD/build/variants/$v/libb200quad.so; fi
  echo "== variant $v"
  timeout 300 python tools/bench_scan.py 2>&1 | cut -c1-330
  timeout 300 python tools/sweep_profile.py 2>&1 | grep '"profile", "dtype": "fp64"'
done
unset B200QUAD_LIB
echo "== sanitizer"
/usr/local/cuda/bin/compute-sanitizer --tool memcheck ./bin/riemann --n 100000000 2>&1 | tail -8
