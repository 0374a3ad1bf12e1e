#!/usr/bin/env python
"""Headline benchmark: G intervals/s and time-to-solution of integrate(sin, 0, pi, N=2^32)
in fp64 on 1/2/4/8 B200 (BASELINE.json "metric"), strong scaling (N is the total).

  python bench.py --gpus 1 --steps 20 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3
  python bench.py --impl reference --gpus 1        # the unmodified reference (baseline/_ref)
  python bench.py --workload scan                  # cumulative_integrate (the 4main.c workload)

A step = one integrate() call = ONE fused kernel per GPU (evaluation of this
rank's share of the N points, block/cluster/grid reduction, cross-GPU combine
over NVLink, result on every rank).

value      device-timed: CUDA events around each step on the launching stream,
           L2 flushed between steps outside the events, max over ranks.  A step
           is one kernel: its job description rides down in the launch's
           parameter buffer; the result (every rank's copy) is left in device
           memory, as in the NCCL comparator arm.
e2e.value  the public API bq.integrate() called K times, wall clock per call on
           the host: builds the job, launches (host -> device: the job in the
           parameter buffer), waits for the completion word, reads the float out
           of the result block the kernel wrote to pinned host memory.
extras     measured after the timed region, each device-timed the same way (not part of
           `value`): the general DIRECT evaluation path, the reference's own problem size
           (n = 1e9, what its binaries run) for a like-for-like time-to-solution, and the
           cumulative-integration (scan) kernel at the reference's 18e6 samples.
"""
import argparse
import ctypes as C
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# the reference's own `riemann` (unmodified, built as its Makefile does) measured on the
# host CPUs of the same B200 box at W = N workers (`mpirun -n N+1`), G intervals/s.  vs_baseline
# divides by the row of the matching worker count.
BASELINE_G_INTERVALS_BY_WORKERS = {1: 0.139, 2: 0.249, 4: 0.422, 8: 0.820}
METRIC = "G intervals/sec at N=2^%d (time-to-solution = ms_per_step)"


def baseline_for(gpus):
    return BASELINE_G_INTERVALS_BY_WORKERS.get(int(gpus), BASELINE_G_INTERVALS_BY_WORKERS[1])


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200quad", choices=["b200quad", "reference"])
    ap.add_argument("--workload", default="integrate", choices=["integrate", "scan"])
    ap.add_argument("--log2n", type=int, default=32)
    ap.add_argument("--n", type=int, default=0, help="scan workload: samples (default 18e6, the reference's)")
    ap.add_argument("--integrand", default="sin")
    ap.add_argument("--rule", default="left")
    ap.add_argument("--dtype", default="fp64", choices=["fp64", "fp32"])
    ap.add_argument("--eval", default="auto", choices=["auto", "direct", "rotate"])
    ap.add_argument("--replicate", type=int, default=0, help="scan workload: 1 = every rank ends with the whole tables")
    ap.add_argument("--no-extras", action="store_true", help="skip the untimed extra measurements")
    ap.add_argument("--combine", default="fused", choices=["fused", "nccl"],
                    help="fused: in-kernel NVLink combine (the product); nccl: local kernel + "
                         "torch.distributed all_reduce of one fp64 (the comparator)")
    ap.add_argument("--local-world", action="store_true",
                    help="drive --gpus GPUs from this one process instead of torchrun ranks")
    return ap.parse_args()


def run_reference(args):
    from baseline import reference_arm
    if int(os.environ.get("RANK", "0")) == 0:
        emit(reference_arm.run(args.gpus, args.steps, args.warmup, args.log2n))
    return 0


_REAL_STDOUT = None


def _quiet_stdout():
    """Everything except the final JSON line goes to stderr: libraries (NCCL's version banner,
    torchrun) write to fd 1 behind Python's back, and the driver wants exactly one line."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(obj):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(obj) + "\n")
    out.flush()


class Env:
    """Process group, device, context and the L2-flush buffer shared by every measurement."""

    def __init__(self, args):
        import torch
        import b200quad as bq
        from b200quad import _native, parallel
        self.torch, self.bq, self.native, self.parallel = torch, bq, _native, parallel
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            torch.cuda.set_device(self.local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist
        lib, cuda = _native.load()
        if not cuda or _native.cuda_device_count() < 1:
            raise SystemExit("bench.py needs a GPU (the CUDA library did not load or no device)")
        self.lib = lib
        self.dev = torch.device("cuda", self.local_rank)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)   # > 126 MB L2
        if self.world > 1:
            self.ctx = parallel.dist_world().ctx
        else:
            self.ctx = parallel.context(self.local_rank)
        self.stream = torch.cuda.ExternalStream(self.ctx.stream, device=self.dev)

    def barrier(self):
        self.torch.cuda.synchronize(self.dev)
        if self.dist is not None:
            self.dist.barrier()
            self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, values):
        t = self.torch.tensor([float(v) for v in values], dtype=self.torch.float64, device=self.dev)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.tolist()]

    def device_timed(self, launch, steps, warmup):
        """ms per step of `launch()` (enqueues on the context's stream): one event pair per step, the
        256 MiB L2 flush outside the pairs, every rank in lock step, max over ranks."""
        torch = self.torch
        for _ in range(warmup):
            launch()
            with torch.cuda.stream(self.stream):
                self.flush.zero_()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        self.barrier()
        for e0, e1 in evs:
            with torch.cuda.stream(self.stream):
                self.flush.zero_()                  # L2 flush, outside the timed events
                e0.record(self.stream)
            launch()
            e1.record(self.stream)
        self.barrier()
        ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
        return self.max_over_ranks([ms])[0] / steps

    def close(self):
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def main():
    args = parse()
    _quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)

    env = Env(args)
    if env.world != args.gpus:
        if env.world == 1 and args.gpus > 1:
            return bench_local_world(args, env)       # plain `python bench.py --gpus N`: single-process world
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, env.world))
    if args.workload == "scan":
        rc = bench_scan(args, env)
    elif args.combine == "nccl":
        rc = bench_nccl_comparator(args, env)
    else:
        rc = bench_integrate(args, env)
    env.close()
    return rc


# ---- the headline: integrate() ----------------------------------------------------------------

def bench_integrate(args, env):
    torch, bq, ctx = env.torch, env.bq, env.ctx
    n = 1 << args.log2n
    a, b = (0.0, math.pi) if args.integrand != "profile" else (0.0, 1800.0)
    job = bq.make_job(args.integrand, a, b, n, args.rule, args.dtype, args.eval)

    # ---- device-timed arm ---------------------------------------------------
    # The step's result stays in device memory here (a float64 tensor the kernel stores into),
    # exactly like the NCCL comparator arm; delivering it to the host is what the e2e arm times.
    d_value = torch.zeros(1, dtype=torch.float64, device=env.dev)
    torch.cuda.synchronize(env.dev)
    sampler = bq.utils.ClockSampler(env.local_rank, 2).start() if env.rank == 0 else None
    wall0 = time.perf_counter()
    ms_per_step = env.device_timed(lambda: ctx.integrate_device_async(job, d_value.data_ptr()),
                                   args.steps, args.warmup)
    wall1 = time.perf_counter()
    res = ctx.fetch_device()
    device_arm_value = float(d_value)
    kern_ns = env.max_over_ranks([int(res.t_end_ns) - int(res.t_start_ns)])[0]
    stamps = {k: round(v / 1e3, 2) for k, v in zip(env.native.STAMPS[:7], list(res.stamp_ns)[:7])}

    # ---- end-to-end arm: the public API, one synchronous call per step --------
    # bq.integrate() is the call a user makes: it returns the float the kernel wrote to host memory.
    kw = dict(dtype=args.dtype, eval=args.eval)
    for _ in range(args.warmup):
        bq.integrate(args.integrand, a, b, n, args.rule, **kw)
    env.barrier()
    e2e_s = 0.0
    e2e_value_last = None
    for _ in range(args.steps):
        t0 = time.perf_counter()
        e2e_value_last = bq.integrate(args.integrand, a, b, n, args.rule, **kw)
        e2e_s += time.perf_counter() - t0
        env.flush.zero_()
        env.barrier()          # untimed: L2 flushed and the ranks level before the next call
    env.barrier()
    clocks = sampler.stop() if sampler else None
    last = bq.integrate_result(args.integrand, a, b, n, args.rule, **kw)
    assert last.value == e2e_value_last == device_arm_value, (last.value, e2e_value_last, device_arm_value)
    env.barrier()
    e2e_ms = env.max_over_ranks([e2e_s * 1e3])[0] / args.steps

    extras = None if args.no_extras else measure_extras(args, env)

    if env.rank == 0:
        value = n / (ms_per_step * 1e-3) / 1e9
        e2e_value = n / (e2e_ms * 1e-3) / 1e9
        exact = bq.utils.oracle_sin(a, b, n, args.rule) if args.integrand == "sin" else None
        out = {
            "metric": METRIC % args.log2n,
            "value": value, "unit": "Gintervals/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": value / baseline_for(args.gpus),
            "dtype": args.dtype, "data": "synthetic (analytic integrand, no input tensors)",
            "impl": "b200quad",
            "config": {"model": "integrate(%s, %g, %g, n=2^%d, rule=%s)" % (
                           args.integrand, a, b, args.log2n, args.rule),
                       "global_batch": n, "seq_len": n // args.gpus,
                       "parallelism": "dp%d (interval decomposition, in-kernel NVLink combine)" % args.gpus,
                       "eval": last.eval, "l2": "flushed between steps (256 MiB memset, untimed)",
                       "baseline": "BASELINE.md: reference riemann on the same box, %d worker(s): %.3f Gintervals/s"
                                   % (args.gpus, baseline_for(args.gpus))},
            "e2e": {"value": e2e_value, "unit": "Gintervals/s",
                    "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": bytes_h2d(), "d2h_bytes_per_step": bytes_d2h(),
                    "h2d_path": "job description (sizeof Plan) + the 128 row constants of the fp64 loop, in the kernel-launch parameter buffer; "
                                "the workload has no input tensors",
                    "d2h_path": "result block written by the kernel into pinned mapped host memory; the host "
                                "watches the block's completion word and reads the float from it"},
            "gpu_launches": args.steps * args.gpus,
            "gpu_launches_per_rank": args.steps,
            "gpu_launches_note": "the %d timed steps of `value`; warm-up, the e2e arm and the extras launch the "
                                 "same kernels again outside it" % args.steps,
            "roofline": roofline(args, value),
            "kernel_us_last_step": kern_ns / 1e3,
            "kernel_stamps_us_rank0_last_step": stamps,
            "wall_s_timed_region": wall1 - wall0,
            "result": last.value,
            "rel_err_vs_oracle": abs(last.value - exact) / abs(exact) if exact else None,
            "clocks": clocks,
            "extras": extras,
        }
        emit(out)
    return 0


def roofline(args, value):
    """Fraction of the CUDA-core issue roofline for the headline's evaluation mode (ROTATE: two fused
    FMAs per interval).  The denominator is builder-measured (tools/peak_fp64.cu ->
    profiles/peak_fp64_fp32.jsonl): MEASURED_PEAKS.json carries no fp64 / fp32 CUDA-core entry."""
    if args.integrand not in ("sin", "acc", "vel") or args.eval == "direct":
        return None
    peak = 18563.0 if args.dtype == "fp64" else 35381.0          # G FMA instruction-lanes/s per GPU
    bound = peak / 2.0 * args.gpus
    return {"instr_per_interval": 2, "peak_g_fma_lanes_per_s_per_gpu": peak, "bound_g_intervals_per_s": bound,
            "frac_of_builder_measured_peak": value / bound,
            "note": "fp64: the two-FMA stream with uniform-register row operands reaches 97 % of this peak in isolation, "
                    "86 % with three register operands (tools/dfma_bench.cu)"}


def measure_extras(args, env):
    """Untimed-region measurements reported next to the headline (each: 5 steps, 3 warm-up,
    device-timed like `value`, max over ranks)."""
    torch, bq, ctx = env.torch, env.bq, env.ctx
    d_value = torch.zeros(1, dtype=torch.float64, device=env.dev)
    out = {}

    def quad(name, a, b, n, ev, dtype="fp64"):
        job = bq.make_job(name, a, b, n, "left", dtype, ev)
        ms = env.device_timed(lambda: ctx.integrate_device_async(job, d_value.data_ptr()), 5, 3)
        return {"ms_per_step": ms, "g_per_s": n / (ms * 1e-3) / 1e9, "value": float(d_value)}

    n32 = 1 << 32
    # DIRECT = x formed and f called for every sample: the path every non-trig / custom integrand takes
    out["direct_sin_2p32"] = quad("sin", 0.0, math.pi, n32, "direct")
    out["direct_profile_2p32"] = quad("profile", 0.0, 1800.0, n32, "direct")
    out["rotate_sin_fp32_2p32"] = quad("sin", 0.0, math.pi, n32, "rotate", "fp32")
    # the reference's own problem (riemann.cpp:10, cintegrate.cu:20: 1e9 steps), for time-to-solution
    same = quad("sin", 0.0, math.pi, 1_000_000_000, "auto")
    t0 = time.perf_counter()
    for _ in range(5):
        v = bq.integrate("sin", 0.0, math.pi, 1_000_000_000)
    same["e2e_ms_per_call"] = env.max_over_ranks([(time.perf_counter() - t0) / 5 * 1e3])[0]
    same["value"] = v
    same["note"] = "integrate(sin, 0, pi, n=1e9): the reference prints 7.2 s (1 worker) .. 1.2 s (8 workers) for it"
    out["same_config_n1e9"] = same
    # cumulative integration at the reference's size (4main.c: 18e6 samples, two running sums)
    out["scan_profile_18e6"] = scan_measure(env, "profile", 18_000_000, 1.0e-4, replicate=False, steps=5, warmup=3)
    return out


# ---- cumulative integration ---------------------------------------------------------------------

def scan_measure(env, name, n, dx, replicate, steps, warmup):
    torch, bq, ctx = env.torch, env.bq, env.ctx
    integ = bq.models.REGISTRY[name]
    job = env.native.ScanJob(integ.native_id, 1 if replicate else 0, 0, 0, 0.0, float(dx), int(n),
                             (C.c_double * 4)(0, 0, 0, 0))
    tiles = (n + 2047) // 2048
    segs = 8 // env.world
    lo = (env.rank * segs * tiles // 8) * 2048
    hi = min(n, ((env.rank + 1) * segs * tiles // 8) * 2048)
    opened = None
    if replicate and env.world > 1:
        P = torch.empty(n, dtype=torch.float64, device=env.dev)
        Q = torch.empty(n, dtype=torch.float64, device=env.dev)
        opened = env.parallel.dist_world().share_tables(P, Q)
    else:
        m = n if env.world == 1 else hi - lo
        P = torch.empty(m, dtype=torch.float64, device=env.dev)
        Q = torch.empty(m, dtype=torch.float64, device=env.dev)
    torch.cuda.synchronize(env.dev)

    def launch():
        rc = env.lib.bq_cumulative_async(ctx.handle, C.byref(job), C.c_void_p(P.data_ptr()),
                                         C.c_void_p(Q.data_ptr()), None, C.c_void_p(ctx.stream))
        assert rc == 0, env.native.last_error()
    ms = env.device_timed(launch, steps, warmup)
    res = env.native.ScanResult()
    env.native.check(env.lib.bq_scan_result_fetch(ctx.handle, C.byref(res)), "bq_scan_result_fetch")
    if opened is not None:
        env.parallel.dist_world().release_tables(opened)
    peaks = {}
    if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")):
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    hbm = float(peaks.get("hbm_gbs", 6551.4))
    out_bytes_per_gpu = 16.0 * (n if (replicate or env.world == 1) else n / env.world)
    gbs = out_bytes_per_gpu / (ms * 1e-3) / 1e9
    return {"n": n, "integrand": name, "replicate": bool(replicate), "ms_per_step": ms,
            "gsamples_per_s": n / (ms * 1e-3) / 1e9,
            "table_GBps_written_per_gpu": gbs,
            "frac_of_measured_hbm_copy_peak": gbs / hbm if not (replicate and env.world > 1) else None,
            "p_last_times_dx": float(res.p_last) * dx, "status": int(res.status)}


def bench_scan(args, env):
    bq = env.bq
    n = args.n or 18_000_000
    dx = 1800.0 / n
    sampler = bq.utils.ClockSampler(env.local_rank, 2).start() if env.rank == 0 else None
    r = scan_measure(env, "profile", n, dx, bool(args.replicate), args.steps, args.warmup)
    clocks = sampler.stop() if sampler else None
    if env.rank == 0:
        emit({"metric": "G samples/sec of cumulative_integrate(profile, n=%d) (two running sums, 16 B out per sample)" % n,
              "value": r["gsamples_per_s"], "unit": "Gsamples/s", "n_gpus": args.gpus, "steps": args.steps,
              "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
              "vs_baseline": None, "dtype": "fp64", "data": "synthetic (the ex4vel.h table, lerped on the fly)",
              "impl": "b200quad",
              "config": {"model": "cumulative_integrate(profile, x0=0, dx=1800/n, n=%d)" % n, "global_batch": n,
                         "seq_len": n // args.gpus, "replicate": bool(args.replicate),
                         "parallelism": "sp%d (blockwise scan, in-kernel NVLink carry exchange)" % args.gpus,
                         "l2": "flushed between steps (256 MiB memset, untimed)",
                         "reference": "4main.c prints 0.279 s (1 rank) .. 0.193 s (8 ranks) for n = 18e6 on the same box"},
              "gpu_launches": args.steps * args.gpus, "detail": r, "clocks": clocks})
    return 0


# ---- comparator ------------------------------------------------------------------------------------

def bench_nccl_comparator(args, env):
    """Baseline comparator (NOT the product path): the same kernel with its cross-GPU stage off
    leaves this rank's segment pairs in a device buffer; torch.distributed (NCCL) all_reduces one
    fp64 per rank.  Device-timed exactly like the fused arm."""
    torch, bq, ctx, dist = env.torch, env.bq, env.ctx, env.dist
    n = 1 << args.log2n
    a, b = (0.0, math.pi) if args.integrand != "profile" else (0.0, 1800.0)
    job = bq.make_job(args.integrand, a, b, n, args.rule, args.dtype, args.eval)
    seg = torch.zeros(16, dtype=torch.float64, device=env.dev)
    out = torch.zeros(1, dtype=torch.float64, device=env.dev)

    def step():
        rc = env.lib.bq_integrate_local_async(ctx.handle, C.byref(job), env.rank, env.world,
                                              C.c_void_p(seg.data_ptr()), C.c_void_p(ctx.stream))
        assert rc == 0
        with torch.cuda.stream(env.stream):
            torch.sum(seg, dim=0, keepdim=True, out=out)
            if dist is not None:
                dist.all_reduce(out)
    per = env.device_timed(step, args.steps, args.warmup)
    if env.rank == 0:
        h = (job.b - job.a) / n
        emit({"metric": "G intervals/sec at N=2^%d, NCCL comparator" % args.log2n,
              "impl": "b200quad local kernel + NCCL all_reduce (comparator)",
              "value": n / (per * 1e-3) / 1e9, "unit": "Gintervals/s", "n_gpus": args.gpus,
              "steps": args.steps, "warmup": args.warmup, "ms_per_step": per,
              "higher_is_better": True, "scaling": "strong", "dtype": args.dtype,
              "result": float(out) * h, "gpu_launches": args.steps * args.gpus})
    return 0


def bytes_h2d():
    # sizeof(bq::Plan): the job description every launch carries from the host to the device
    from b200quad import _native
    return int(_native.load()[0].bq_sizeof_params())


def bytes_d2h():
    from b200quad import _native
    return C.sizeof(_native.Result)


def bench_local_world(args, env):
    """Single process driving N GPUs (used by the CLIs; not the driver's launch mode)."""
    bq, torch = env.bq, env.torch
    n = 1 << args.log2n
    a, b = (0.0, math.pi) if args.integrand != "profile" else (0.0, 1800.0)
    job = bq.make_job(args.integrand, a, b, n, args.rule, args.dtype, args.eval)
    w = env.parallel.local_world(args.gpus)
    for _ in range(args.warmup):
        w.integrate(job)
    for d in range(args.gpus):
        torch.cuda.synchronize(d)
    ms = 0.0
    t_wall = 0.0
    for _ in range(args.steps):
        t0 = time.perf_counter()
        res = w.integrate(job)
        t_wall += time.perf_counter() - t0
        ms += max(float(r.device_ms) for r in res)
    value = n / (ms / args.steps * 1e-3) / 1e9
    emit({
        "metric": METRIC % args.log2n,
        "value": value, "unit": "Gintervals/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": value / baseline_for(args.gpus), "dtype": args.dtype,
        "data": "synthetic (analytic integrand, no input tensors)", "impl": "b200quad",
        "config": {"model": "integrate(%s, n=2^%d, rule=%s)" % (args.integrand, args.log2n, args.rule),
                   "global_batch": n, "seq_len": n // args.gpus,
                   "parallelism": "dp%d single-process world" % args.gpus},
        "e2e": {"value": n / (t_wall / args.steps) / 1e9, "unit": "Gintervals/s",
                "h2d_bytes_per_step": bytes_h2d() * args.gpus,
                "d2h_bytes_per_step": bytes_d2h() * args.gpus},
        "gpu_launches": args.steps * args.gpus, "result": float(res[0].value)})
    return 0


if __name__ == "__main__":
    sys.exit(main())
