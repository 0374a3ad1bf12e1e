#!/bin/bash
for v in v6 main; do
  if [ $v = main ]; then unset B200QUAD_LIB; else export B200QUAD_LIB=$PWD/build/variants/$v/libb200quad.so; fi
  echo "== $v"; timeout 300 python tools/bench_scan.py 2>&1 | cut -c1-300
done
