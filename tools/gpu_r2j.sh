#!/bin/bash
mkdir -p gpurun_out
for v in default super8 poll20; do
  if [ $v = default ]; then unset B200QUAD_LIB; else export B200QUAD_LIB=$PWD/build/variants/$v/libb200quad.so; fi
  timeout 200 python tools/bench_scan.py > gpurun_out/bench_scan_$v.jsonl 2>&1; echo "== $v"; cut -c1-125 gpurun_out/bench_scan_$v.jsonl | head -4
done
