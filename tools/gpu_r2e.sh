#!/bin/bash
# ncu captures (one GPU): scan, DIRECT sin / table / custom, ROTATE sin.  Each target runs plainly first.
mkdir -p gpurun_out
cap() { # name regex args...
  name=$1; rx=$2; shift; shift
  python tools/ncu_target.py "$@" > gpurun_out/plain_$name.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -o gpurun_out/prof_$name -f python tools/ncu_target.py "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"; tail -2 gpurun_out/plain_$name.log
}
cap scan scan_kernel scan 26
cap direct_sin quad_kernel direct_sin 30
cap direct_profile quad_kernel direct_profile 30
cap direct_custom quad_kernel direct_custom 28
cap rotate_sin quad_kernel rotate_sin 32
ls -la gpurun_out/*.ncu-rep
