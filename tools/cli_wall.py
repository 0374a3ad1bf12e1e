"""Process-level time-to-solution of the three reference-compatible programs on this box: wall
clock of the whole process (fork to exit) and the "%lf seconds" line each one prints (the
reference's own timer placement), at --world 1 and every larger world the box has, best of 3.
Next to them, when baseline/_ref is built, the reference's binaries on the same box.

  python tools/cli_wall.py > gpurun_out/cli_wall.jsonl
"""
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from b200quad import _native

ngpu = _native.cuda_device_count()
worlds = [w for w in (1, 2, 4, 8) if w <= ngpu]


def run(cmd, reps=3, env=None):
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
        wall = time.perf_counter() - t0
        m = re.search(r"([0-9.]+) seconds", p.stdout)
        rec = {"rc": p.returncode, "wall_s": round(wall, 4), "printed_s": float(m.group(1)) if m else None,
               "last_line": p.stdout.strip().splitlines()[-1][:120] if p.stdout.strip() else ""}
        if best is None or rec["wall_s"] < best["wall_s"]:
            best = rec
    return best


for prog in ("riemann", "cintegrate", "train4"):
    exe = os.path.join(ROOT, "bin", prog)
    print(json.dumps({"program": prog, "args": "(none: default world)", **run([exe])}), flush=True)
    for w in worlds:
        print(json.dumps({"program": prog, "args": "--world %d" % w, **run([exe, "--world", str(w)])}), flush=True)
    print(json.dumps({"program": prog, "args": "--backend cpu", **run([exe, "--backend", "cpu"])}), flush=True)

ref = os.path.join(ROOT, "baseline", "_ref", "bin")
mpirun = os.path.join(ROOT, "baseline", "mpi_shim", "mpirun")
if os.path.exists(os.path.join(ref, "riemann")):
    for w in (1, 8):
        print(json.dumps({"program": "REFERENCE riemann", "args": "mpirun -n %d" % (w + 1),
                          **run([mpirun, "-n", str(w + 1), os.path.join(ref, "riemann")], reps=1)}), flush=True)
        print(json.dumps({"program": "REFERENCE 4main", "args": "mpirun -n %d" % w,
                          **run([mpirun, "-n", str(w), os.path.join(ref, "4main")], reps=1)}), flush=True)
    if ngpu:
        print(json.dumps({"program": "REFERENCE cintegrate", "args": "(as shipped)",
                          **run([os.path.join(ref, "cintegrate")], reps=2)}), flush=True)
