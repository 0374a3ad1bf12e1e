#!/bin/bash
# Host sanitizer tier: AddressSanitizer + UndefinedBehaviorSanitizer builds of the CPU library and
# the three CLIs (python build.py --sanitize) under the CPU test tier and the CLI runs.  No GPU needed.
#   bash tools/sanitize_host.sh            -> profiles/r2/sanitize_host.log
set -u
cd "$(dirname "$0")/.."
out=${1:-profiles/r2/sanitize_host.log}
mkdir -p "$(dirname "$out")"
python build.py --sanitize > /dev/null || { echo "sanitizer build failed"; exit 1; }
SAN=build/sanitize
ASAN_RT=$(/usr/bin/g++ -print-file-name=libasan.so)
UBSAN_RT=$(/usr/bin/g++ -print-file-name=libubsan.so)
export ASAN_OPTIONS=detect_leaks=0:abort_on_error=0:halt_on_error=1
export UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
{
  echo "== host sanitizer tier: $(/usr/bin/g++ --version | head -1), -fsanitize=address,undefined"
  for cmd in "$SAN/riemann --backend cpu --n 30000001" "$SAN/riemann --backend cpu --n 1000003 --rule trap --eval direct" \
             "$SAN/riemann --backend cpu --n 2000000 --integrand profile --rule mid" "$SAN/riemann --backend cpu --n 5000000 --dtype fp32" \
             "$SAN/cintegrate --backend cpu" "$SAN/cintegrate --backend cpu --compat" "$SAN/train4 --backend cpu --steps-per-sec 1000" \
             "$SAN/train4 --backend cpu --steps-per-sec 997 --seconds 1800"; do
    echo "-- $cmd"
    $cmd 2>&1 | tail -4
    echo "   exit code ${PIPESTATUS[0]}"
  done
  echo "== CPU test tier on the sanitized library (B200QUAD_CPU_LIB=$SAN/libb200quad_cpu.so, runtimes preloaded into python)"
  LD_PRELOAD="$ASAN_RT $UBSAN_RT" B200QUAD_FORCE_CPU=1 B200QUAD_CPU_LIB=$PWD/$SAN/libb200quad_cpu.so \
      python -m pytest tests -q -m "not gpu" -x -p no:cacheprovider -k "not cli and not torch_ops" 2>&1 | tail -6
  echo "   exit code ${PIPESTATUS[0]}"
} > "$out" 2>&1
grep -c "runtime error\|ERROR: AddressSanitizer" "$out" | sed 's/^/sanitizer reports: /' >> "$out"
cat "$out"
