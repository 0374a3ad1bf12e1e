"""Reference arm of bench.py: runs the UNMODIFIED reference built by baseline/install.py.

Headline = the reference's own program for the benchmarked integral: riemann.cpp
(left Riemann sum of sin on [0, pi], STEPS hard-coded to 1e9, riemann.cpp:10) run
as `mpirun -n N+1 ./riemann` — N workers + its idle master rank (riemann.cpp:65) —
on the host CPUs, timed as the program itself prints it ("%lf seconds",
riemann.cpp:92).  The reference has no multi-GPU program and its GPU sin kernel
is never launched (cintegrate.cu:128), so "N GPUs" maps to "N MPI workers".

It cannot run the benchmark's N = 2^32: STEPS is a #define without a guard and
per-worker counts are `int` (riemann.cpp:34,55).  The metric is a rate
(G intervals/s), so its 1e9-step rate is reported, with the config saying so.

Extras (same JSON, "extra"): the reference CUDA program as shipped
(cintegrate: LUT integral, 18e6 samples, 64 threads, timer includes context
creation and 288 MB of PCIe copies), 4main at P = N ranks, and — labelled
MODIFIED — cintegrate with its own commented-out sin kernel enabled.
"""
import json
import os
import re
import subprocess
import time

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "_ref", "bin")
MPIRUN = os.path.join(HERE, "mpi_shim", "mpirun")
# the reference's own riemann (built as its Makefile does) measured on the B200 box's
# host CPUs at W = 1/2/4/8 workers, G intervals/s.  Shared with bench.py (vs_baseline).
BASELINE_G_INTERVALS_BY_WORKERS = {1: 0.139, 2: 0.249, 4: 0.422, 8: 0.820}
# The metric string of bench.py's own arm: both arms must report under the same name and unit or
# the driver cannot form their ratio.  What differs (n = 1e9 instead of 2^32) is in `config`.
METRIC = "G intervals/sec at N=2^%d (time-to-solution = ms_per_step)"
# A reference run takes 7.2 s / workers; the requested steps + warm-up are honoured unless they would
# exceed this many seconds, in which case the cap is stated in `config`.
TIME_CAP_S = 900.0


def _run(cmd, timeout=900):
    t0 = time.perf_counter()
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    wall = time.perf_counter() - t0
    m = re.search(r"([0-9.]+) seconds", p.stdout)
    return {"rc": p.returncode, "printed_s": float(m.group(1)) if m else None, "wall_s": wall,
            "stdout": p.stdout.strip().splitlines()[-3:]}


def _gpu_present():
    return os.path.exists("/dev/nvidiactl") or os.path.exists("/dev/nvidia0")


def run(gpus, steps, warmup, log2n=32):
    if not os.path.exists(os.path.join(BIN, "riemann")):
        return {"impl": "reference",
                "unavailable": "baseline/_ref is not built (python baseline/install.py needs /root/reference)"}
    workers = max(1, int(gpus))
    cmd = [MPIRUN, "-n", str(workers + 1), os.path.join(BIN, "riemann")]
    steps, warmup = max(1, int(steps)), max(0, int(warmup))
    first = _run(cmd)                         # first warm-up run doubles as the duration probe
    est = max(first["wall_s"], 1e-3)
    capped = None
    if warmup == 0:
        warm_done = 0
        runs = [first]
    else:
        warm_done = 1
        runs = []
    if est * (steps + warmup) > TIME_CAP_S:
        budget = max(3, int(TIME_CAP_S / est))
        new_warm = min(warmup, 1)
        new_steps = max(1, min(steps, budget - new_warm))
        capped = "requested %d steps + %d warm-up capped to %d + %d (one run takes %.1f s)" % (
            steps, warmup, new_steps, new_warm, est)
        steps, warmup = new_steps, new_warm
    while warm_done < warmup:
        _run(cmd)
        warm_done += 1
    while len(runs) < steps:
        runs.append(_run(cmd))
    ok = [r for r in runs if r["rc"] == 0 and r["printed_s"]]
    if not ok:
        return {"impl": "reference", "unavailable": "reference riemann failed: %s" % runs[-1]["stdout"]}
    sec = sum(r["printed_s"] for r in ok) / len(ok)
    n = 1_000_000_000
    value = n / sec / 1e9
    extra = {}
    try:
        extra["4main_P%d" % workers] = _run([MPIRUN, "-n", str(workers), os.path.join(BIN, "4main")])
        extra["riemann_O2"] = _run([MPIRUN, "-n", str(workers + 1), os.path.join(BIN, "riemann_O2")])
        if _gpu_present():
            _run([os.path.join(BIN, "cintegrate")])                       # context warm-up
            extra["cintegrate_as_shipped"] = _run([os.path.join(BIN, "cintegrate")])
            extra["cintegrate_sin_MODIFIED_launch_line"] = _run([os.path.join(BIN, "cintegrate_sin")])
    except Exception as e:  # extras never break the headline
        extra["error"] = repr(e)
    m = re.search(r"is ([0-9.eE+-]+)", " ".join(ok[-1]["stdout"]))
    return {
        "impl": "reference",
        "metric": METRIC % log2n,
        "value": value, "unit": "Gintervals/s", "n_gpus": gpus, "steps": len(ok), "warmup": warm_done,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": value / BASELINE_G_INTERVALS_BY_WORKERS.get(workers, 0.139), "dtype": "fp64",
        "data": "synthetic (analytic integrand)",
        "config": {"model": "riemann.cpp unmodified, mpirun -n %d (=%d workers + idle master), n=1e9, "
                            "built as its Makefile does (no -O flags), MPI shim over shared memory"
                            % (workers + 1, workers),
                   "global_batch": n, "seq_len": n // workers, "parallelism": "mpi%d (host CPUs)" % workers,
                   "n_note": "the reference cannot run N=2^%d: STEPS is hard-coded to 1e9 and its per-worker "
                             "counters are int (riemann.cpp:10,34,55); the metric is a rate, so its rate at "
                             "its own n=1e9 is what is reported; ms_per_step is the time of ITS 1e9-step "
                             "run (bench.py's own arm reports same_config_n1e9 for the like-for-like time)"
                             % log2n,
                   "steps_note": capped or "steps and warm-up as requested",
                   "timing": "as printed by the program (includes MPI_Init)"},
        "e2e": {"value": n / (sum(r["wall_s"] for r in ok) / len(ok)) / 1e9, "unit": "Gintervals/s",
                "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "note": "process wall clock incl. launch; the reference has no device path for this integral"},
        "gpu_launches": 0,
        "result": float(m.group(1)) if m else None,
        "extra": extra,
    }


if __name__ == "__main__":
    print(json.dumps(run(1, 1, 0)))
