#!/bin/bash
# Sanitizer tier: compute-sanitizer over both fused kernels on one GPU,
# through the reference-compatible CLIs.  Output: gpurun_out/sanitizer.txt
out=gpurun_out/sanitizer.txt
mkdir -p gpurun_out; : > $out
run() {
  echo "== compute-sanitizer --tool $1 ${@:2}" >> $out
  timeout 600 compute-sanitizer --tool $1 --print-limit 5 "${@:2}" 2>&1 | grep -E "ERROR SUMMARY|RACECHECK SUMMARY|error|Error|hazard|seconds|distance|integral" | head -12 >> $out
}
run memcheck ./bin/riemann --n 100000000
run memcheck ./bin/riemann --n 3000001 --eval direct --rule trap
run memcheck ./bin/cintegrate
run memcheck ./bin/train4 --steps-per-sec 1000
run racecheck ./bin/riemann --n 20000000
run racecheck ./bin/train4 --steps-per-sec 100
run synccheck ./bin/riemann --n 20000000
run synccheck ./bin/train4 --steps-per-sec 100
run initcheck ./bin/train4 --steps-per-sec 100
cat $out
