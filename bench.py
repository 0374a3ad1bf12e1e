#!/usr/bin/env python
"""Headline benchmark: G intervals/s and time-to-solution of integrate(sin, 0, pi, N=2^32)
in fp64 on 1/2/4/8 B200 (BASELINE.json "metric"), strong scaling (N is the total).

  python bench.py --gpus 1 --steps 20 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 3
  python bench.py --impl reference --gpus 1        # the unmodified reference (baseline/_ref)

A step = one integrate() call = ONE fused kernel per GPU (evaluation of this
rank's share of the N points, block/cluster/grid reduction, cross-GPU combine
over NVLink, result on every rank).

value      device-timed: CUDA events around each step on the launching stream,
           L2 flushed between steps outside the events, max over ranks.  A step
           is one kernel: its 152-byte job description rides down in the launch's
           parameter buffer; the result (every rank's copy) is left in device
           memory, as in the NCCL comparator arm.
e2e.value  the public API bq.integrate() called K times, wall clock per call on
           the host: builds the job, launches (host -> device: the job in the
           parameter buffer), waits for the completion word, reads the float out
           of the result block the kernel wrote to pinned host memory.
"""
import argparse
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


def baseline_for(gpus):
    return BASELINE_G_INTERVALS_BY_WORKERS.get(int(gpus), BASELINE_G_INTERVALS_BY_WORKERS[1])


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200quad", choices=["b200quad", "reference"])
    ap.add_argument("--log2n", type=int, default=32)
    ap.add_argument("--integrand", default="sin")
    ap.add_argument("--rule", default="left")
    ap.add_argument("--dtype", default="fp64", choices=["fp64", "fp32"])
    ap.add_argument("--eval", default="auto", choices=["auto", "direct", "rotate"])
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


def main():
    args = parse()
    _quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import b200quad as bq
    from b200quad import _native, parallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if world != args.gpus and not (world == 1 and args.local_world):
        if world == 1 and args.gpus > 1:
            args.local_world = True     # plain `python bench.py --gpus N`: single-process world
        elif world != args.gpus:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))

    n = 1 << args.log2n
    a, b = (0.0, math.pi) if args.integrand != "profile" else (0.0, 1800.0)
    job = bq.make_job(args.integrand, a, b, n, args.rule, args.dtype, args.eval)
    lib, cuda = _native.load()
    if not cuda or _native.cuda_device_count() < 1:
        raise SystemExit("bench.py needs a GPU (the CUDA library did not load or no device)")

    dev = torch.device("cuda", local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    if args.local_world and args.gpus > 1:
        return bench_local_world(args, bq, parallel, job, n, torch)

    if world > 1:
        w = parallel.dist_world()
        ctx = w.ctx
    else:
        ctx = parallel.context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize(dev)

    if args.combine == "nccl":
        return bench_nccl_comparator(args, bq, ctx, job, n, dist, dev, stream, flush, rank, world, torch)

    # ---- device-timed arm ---------------------------------------------------
    # The step's result stays in device memory here (a float64 tensor the kernel stores into),
    # exactly like the NCCL comparator arm; delivering it to the host is what the e2e arm times.
    d_value = torch.zeros(1, dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    for _ in range(args.warmup):
        ctx.integrate_device_async(job, d_value.data_ptr())
        with torch.cuda.stream(stream):
            flush.zero_()
    res = ctx.fetch_device()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    sampler = bq.utils.ClockSampler(local_rank, 2).start() if rank == 0 else None
    barrier()
    wall0 = time.perf_counter()
    for e0, e1 in evs:
        with torch.cuda.stream(stream):
            flush.zero_()                      # L2 flush, outside the timed events
            e0.record(stream)
        ctx.integrate_device_async(job, d_value.data_ptr())   # ONE kernel: params in, compute, combine
        e1.record(stream)
    barrier()
    wall1 = time.perf_counter()
    res = ctx.fetch_device()
    device_arm_value = float(d_value)
    clocks = sampler.stop() if sampler else None
    dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
    kern_ns = int(res.t_end_ns) - int(res.t_start_ns)

    # ---- end-to-end arm: the public API, one synchronous call per step --------
    # bq.integrate() is the call a user makes: it returns the float the kernel wrote to host memory.
    for _ in range(args.warmup):
        bq.integrate(args.integrand, a, b, n, args.rule, dtype=args.dtype, eval=args.eval)
    barrier()
    e2e_s = 0.0
    e2e_value_last = None
    for _ in range(args.steps):
        t0 = time.perf_counter()
        e2e_value_last = bq.integrate(args.integrand, a, b, n, args.rule, dtype=args.dtype, eval=args.eval)
        e2e_s += time.perf_counter() - t0
        flush.zero_()
        torch.cuda.synchronize(dev)
    barrier()
    last = bq.integrate_result(args.integrand, a, b, n, args.rule, dtype=args.dtype, eval=args.eval)
    assert last.value == e2e_value_last == device_arm_value, (last.value, e2e_value_last, device_arm_value)
    barrier()

    t = torch.tensor([dev_ms, e2e_s * 1e3, float(kern_ns)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kern_ns = [float(x) for x in t.tolist()]

    if rank == 0:
        ms_per_step = dev_ms / args.steps
        value = n / (ms_per_step * 1e-3) / 1e9
        e2e_value = n / (e2e_ms / args.steps * 1e-3) / 1e9
        exact = bq.utils.oracle_sin(a, b, n, args.rule) if args.integrand == "sin" else None
        out = {
            "metric": "G intervals/sec at N=2^%d (time-to-solution = ms_per_step)" % args.log2n,
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
                    "ms_per_step": e2e_ms / args.steps,
                    "h2d_bytes_per_step": bytes_h2d(), "d2h_bytes_per_step": bytes_d2h(),
                    "h2d_path": "job description (sizeof Plan) in the kernel-launch parameter buffer; "
                                "the workload has no input tensors",
                    "d2h_path": "result block written by the kernel into pinned mapped host memory; the host "
                                "watches the block's completion word and reads the float from it"},
            "gpu_launches": args.steps * args.gpus,
            "gpu_launches_per_rank": args.steps,
            "kernel_us_last_step": kern_ns / 1e3,
            "wall_s_timed_region": wall1 - wall0,
            "result": last.value,
            "rel_err_vs_oracle": abs(last.value - exact) / abs(exact) if exact else None,
            "clocks": clocks,
        }
        emit(out)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def bench_nccl_comparator(args, bq, ctx, job, n, dist, dev, stream, flush, rank, world, torch):
    """Baseline comparator (NOT the product path): the same kernel with its cross-GPU stage off
    leaves this rank's segment pairs in a device buffer; torch.distributed (NCCL) all_reduces one
    fp64 per rank.  Device-timed exactly like the fused arm."""
    import ctypes as C
    lib = ctx.lib
    seg = torch.zeros(16, dtype=torch.float64, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)

    def step():
        rc = lib.bq_integrate_local_async(ctx.handle, C.byref(job), rank, world,
                                          C.c_void_p(seg.data_ptr()), C.c_void_p(ctx.stream))
        assert rc == 0
        with torch.cuda.stream(stream):
            torch.sum(seg, dim=0, keepdim=True, out=out)
            if dist is not None:
                dist.all_reduce(out)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize(dev)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    for e0, e1 in evs:
        with torch.cuda.stream(stream):
            flush.zero_()
            e0.record(stream)
        step()
        e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = torch.tensor([sum(e0.elapsed_time(e1) for e0, e1 in evs)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        per = float(ms) / args.steps
        h = (job.b - job.a) / n
        emit({"metric": "G intervals/sec at N=2^%d, NCCL comparator" % args.log2n,
                          "impl": "b200quad local kernel + NCCL all_reduce (comparator)",
                          "value": n / (per * 1e-3) / 1e9, "unit": "Gintervals/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": per,
                          "higher_is_better": True, "scaling": "strong", "dtype": args.dtype,
                          "result": float(out) * h, "gpu_launches": args.steps * args.gpus})
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def bytes_h2d():
    # sizeof(bq::Plan): the job description every launch carries from the host to the device
    from b200quad import _native
    return int(_native.load()[0].bq_sizeof_params())


def bytes_d2h():
    import ctypes
    from b200quad import _native
    return ctypes.sizeof(_native.Result)


def bench_local_world(args, bq, parallel, job, n, torch):
    """Single process driving N GPUs (used by the CLIs; not the driver's launch mode)."""
    w = parallel.local_world(args.gpus)
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
        "metric": "G intervals/sec at N=2^%d (time-to-solution = ms_per_step)" % args.log2n,
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
