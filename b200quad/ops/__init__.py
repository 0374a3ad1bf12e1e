"""User-facing operations: integrate() and cumulative_integrate().

integrate(f, a, b, n, rule, world) is the API BASELINE.json asks for.  With a
library integrand it is one fused sm_100a kernel per GPU (evaluation, block /
cluster / grid reduction and the cross-GPU combine in the same launch); the
reference needs an MPI job (riemann.cpp) or a 64-thread kernel plus a host loop
(cintegrate.cu:127-138) for the same thing.
"""
import ctypes as C
import math
import os

from .. import _native, models, parallel
from .._native import Job, ScanJob, NativeError

__all__ = ["integrate", "integrate_result", "cumulative_integrate", "IntegrateResult", "backend"]


class IntegrateResult:
    """Value plus what the run looked like (mirrors bq_result)."""

    def __init__(self, res, backend, world):
        self.value = float(res.value)
        self.value_lo = float(res.value_lo)
        self.raw_sum = (float(res.raw_sum_hi), float(res.raw_sum_lo))
        self.segments = [(float(res.seg_hi[i]), float(res.seg_lo[i])) for i in range(_native.SEGMENTS)]
        self.n = int(res.n)
        self.status = int(res.status)
        self.eval = _native.EVAL_NAMES.get(int(res.eval_used), "?")
        self.world = world
        self.rank = int(res.rank)
        self.device_ms = float(res.device_ms)
        self.kernel_ns = int(res.t_end_ns) - int(res.t_start_ns) if res.t_end_ns else 0
        self.launches = int(res.launches)
        self.backend = backend

    def __float__(self):
        return self.value

    def __repr__(self):
        return ("IntegrateResult(value=%.17g, n=%d, eval=%s, backend=%s, world=%d, device_ms=%.4f)"
                % (self.value, self.n, self.eval, self.backend, self.world, self.device_ms))


_backend_cache = None


def backend():
    """'cuda' when the CUDA library loaded and a device exists, else 'cpu'."""
    global _backend_cache
    if _backend_cache is None:
        _backend_cache = "cuda" if _native.cuda_device_count() > 0 else "cpu"
    return _backend_cache


def make_job(f, a, b, n, rule="left", dtype="fp64", eval="auto", params=None):
    integ, _ = models.get(f)
    if integ is None:
        raise TypeError("make_job needs a library integrand")
    native_id = integ.native_id
    if getattr(integ, "expr", None) is not None:
        from ..models import jit as _custom
        native_id = _custom.integrand_code(integ.expr)   # builds / loads / registers its kernel on first use
        if params is None:
            params = integ.params
    try:
        job = Job(native_id, _native.RULES[rule], _native.DTYPES[dtype], _native.EVALS[eval],
                  float(a), float(b), int(n), (C.c_double * 4)(*(params or (0.0, 0.0, 0.0, 0.0))))
    except KeyError as e:
        raise ValueError("unknown option %s" % e)
    return job


def _dist_state():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist, dist.get_world_size(), dist.get_backend()
    except Exception:
        pass
    return None, 1, None


_fast_jobs = {}


def _dist_initialised():
    try:
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized()
    except Exception:
        return False


def _callable_integrate(fn, a, b, n, rule, dtype, chunk=1 << 22):
    """Slow path for arbitrary Python callables: chunked torch evaluation, Neumaier across chunks."""
    import torch
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    tdt = torch.float64 if _native.DTYPES[dtype] == 0 else torch.float32
    h = (b - a) / n
    off = {"left": 0.0, "right": 1.0, "mid": 0.5, "midpoint": 0.5, "trap": 0.0, "trapezoid": 0.0}[rule]
    s, c = 0.0, 0.0
    for start in range(0, n, chunk):
        m = min(chunk, n - start)
        i = torch.arange(start, start + m, device=dev, dtype=torch.float64)
        x = (a + (i + off) * h).to(tdt)
        part = float(fn(x).to(torch.float64).sum())
        t = s + part
        c += (s - t) + part if abs(s) >= abs(part) else (part - t) + s
        s = t
    val = (s + c) * h
    if rule in ("trap", "trapezoid"):
        ends = fn(torch.tensor([a, b], device=dev, dtype=tdt)).to(torch.float64)
        val += 0.5 * h * float(ends[1] - ends[0])
    return val


def integrate_result(f, a, b, n, rule="left", world=None, *, dtype="fp64", eval="auto",
                     backend=None, params=None, threads=None):
    """Like integrate() but returns an IntegrateResult."""
    integ, fn = models.get(f)
    if integ is None:
        raise TypeError("integrate_result needs a library integrand (got a Python callable)")
    job = make_job(integ, a, b, n, rule, dtype, eval, params)
    dist, dsize, dbackend = _dist_state()
    be = backend or globals()["backend"]()
    if world is not None and int(world) < 1:
        raise ValueError("world must be >= 1")

    if be == "cpu":
        if dist is not None and (world is None or int(world) == dsize) and dsize > 1:
            w = parallel.CpuWorld(threads=threads)
            return IntegrateResult(w.integrate(job), "cpu", dsize)
        w = parallel.CpuWorld.__new__(parallel.CpuWorld)
        w.lib, _ = _native.load()
        w.dist, w.rank, w.size, w.group = None, 0, 1, None
        w.threads = threads or max(1, os.cpu_count() or 1)
        return IntegrateResult(w.integrate(job), "cpu", 1)

    if be != "cuda":
        raise ValueError("backend must be 'cuda' or 'cpu'")
    if dist is not None and dsize > 1 and (world is None or int(world) == dsize):
        if dbackend != "nccl" and "nccl" not in str(dbackend):
            raise NativeError("CUDA backend under torch.distributed needs the nccl process group")
        return IntegrateResult(parallel.dist_world().integrate(job), "cuda", dsize)
    w = 1 if world is None else int(world)
    if w == 1:
        return IntegrateResult(parallel.context().integrate(job), "cuda", 1)
    res = parallel.local_world(w).integrate(job)
    return IntegrateResult(res[0], "cuda", w)


def integrate(f, a, b, n, rule="left", world=None, *, dtype="fp64", eval="auto", backend=None,
              params=None, threads=None):
    """Definite integral of f over [a, b] with n sub-intervals.

    f      'sin' | 'acc' | 'vel' | 'dis' | 'profile' | models.Integrand | callable(tensor)->tensor
    rule   'left' (the reference's rule, riemann.cpp:36-39) | 'right' | 'mid' | 'trap'
    world  None: every rank of the initialised process group, else this process's
           current GPU; k: k GPUs (driven by this process when no process group exists)
    dtype  'fp64' | 'fp32' (compensated)
    eval   'auto' | 'direct' | 'rotate'
    Returns a Python float; every rank of the world gets the same bits.
    """
    # Repeated calls with the same arguments (a solver loop, a benchmark) skip the argument
    # normalisation: the prepared job and its context are remembered per argument tuple.
    key = None
    if isinstance(f, str) and params is None:
        key = (f, a, b, n, rule, world, dtype, eval, backend, _dist_initialised())
        hit = _fast_jobs.get(key)
        if hit is not None:
            return hit[0].integrate_value(hit[1])
    integ, fn = models.get(f)
    if integ is None:
        return _callable_integrate(fn, float(a), float(b), int(n), rule, dtype)
    if (backend or globals()["backend"]()) == "cuda":
        # Fast path: this process's GPU alone, or its rank of the initialised process group.
        # One kernel launch; the float is read from the pinned block the kernel wrote.
        dist, dsize, dbackend = _dist_state()
        ctx = None
        if dist is not None and dsize > 1 and (world is None or int(world) == dsize):
            if "nccl" in str(dbackend):
                ctx = parallel.dist_world().ctx
        elif world is None or int(world) == 1:
            ctx = parallel.context()
        if ctx is not None:
            job = make_job(integ, a, b, n, rule, dtype, eval, params)
            if key is not None:
                if len(_fast_jobs) > 256:
                    _fast_jobs.clear()
                _fast_jobs[key] = (ctx, job)
            return ctx.integrate_value(job)
    return integrate_result(integ, a, b, n, rule, world, dtype=dtype, eval=eval, backend=backend,
                            params=params, threads=threads).value


class CumulativeResult:
    def __init__(self, P, Q, samples, res, backend, world, slice_):
        self.P, self.Q, self.samples = P, Q, samples
        self.p_last = float(res.p_last)
        self.q_last = float(res.q_last)
        self.p_second_last = float(res.p_second_last)
        self.slice = slice_
        self.status = int(res.status)
        self.device_ms = float(res.device_ms)
        self.kernel_ns = int(res.t_end_ns) - int(res.t_start_ns) if res.t_end_ns else 0
        self.backend, self.world = backend, world


def cumulative_integrate(f="profile", x0=0.0, dx=1.0e-4, n=18_000_000, world=None, *,
                         replicate=True, want_samples=False, backend=None, params=None,
                         threads=None):
    """Running sums on a uniform grid (capability of 4main.c:76-225).

    P[i] = sum_{k<=i} f(x0 + k dx)   ("default_sum"), Q[i] = sum_{k<=i} P[k]
    ("default_sum_of_sums").  Defaults are the reference's: the lerped velocity
    profile, 10000 steps per second over 1800 s.  Multiply P by dx (and Q by
    dx**2) for position / its integral.

    Returns a CumulativeResult with torch tensors P, Q (float64).  With several
    ranks and replicate=True every rank holds the complete tables (the
    reference's MPI_Bcast, 4main.c:157,221); with replicate=False each rank
    holds its slice `result.slice`.
    """
    import torch
    integ, fn = models.get(f)
    if integ is None:
        raise TypeError("cumulative_integrate needs a library integrand")
    n = int(n)
    job = ScanJob(integ.native_id, 1 if replicate else 0, 1 if want_samples else 0, 0, float(x0),
                  float(dx), n, (C.c_double * 4)(*(params or (0.0, 0.0, 0.0, 0.0))))
    res = _native.ScanResult()
    be = backend or globals()["backend"]()
    dist, dsize, dbackend = _dist_state()
    lib, _ = _native.load()
    if be == "cpu" and dist is not None and dsize > 1 and (world is None or int(world) == dsize):
        # one CPU rank per process (gloo): every rank scans its segments, 32 doubles are exchanged
        w = parallel.CpuWorld(threads=threads)
        P, Q, S, res, sl = w.cumulative(job, want_samples=want_samples, replicate=replicate)
        return CumulativeResult(P, Q, S, res, "cpu", dsize, sl)
    if be == "cpu":
        P = torch.empty(n, dtype=torch.float64)
        Q = torch.empty(n, dtype=torch.float64)
        S = torch.empty(n, dtype=torch.float64) if want_samples else None
        _native.check(lib.bq_cpu_cumulative(C.byref(job), threads or max(1, os.cpu_count() or 1),
                                            C.c_void_p(P.data_ptr()), C.c_void_p(Q.data_ptr()),
                                            C.c_void_p(S.data_ptr() if S is not None else 0),
                                            C.byref(res)), "bq_cpu_cumulative")
        return CumulativeResult(P, Q, S, res, "cpu", 1, (0, n))

    if dist is not None and dsize > 1 and (world is None or int(world) == dsize):
        dw = parallel.dist_world()
        ctx = dw.ctx
        dev = torch.device("cuda", ctx.device)
        tiles = (n + 2047) // 2048
        lo = (dw.rank * (8 // dsize) * tiles // 8) * 2048
        hi = min(n, (((dw.rank + 1) * (8 // dsize)) * tiles // 8) * 2048)
        if replicate:
            P = torch.empty(n, dtype=torch.float64, device=dev)
            Q = torch.empty(n, dtype=torch.float64, device=dev)
            opened = dw.share_tables(P, Q)
        else:
            # a rank whose segments hold no tile (tiny n, many ranks) still takes part in the call:
            # it gets a valid, empty slice
            # (torch reports a null data_ptr for an empty view: the kernel gets the base buffer's address)
            m = max(hi - lo, 0)
            P = torch.empty(max(m, 4), dtype=torch.float64, device=dev)
            Q = torch.empty(max(m, 4), dtype=torch.float64, device=dev)
            opened = None
        S = None
        if want_samples:
            S = torch.empty(max(n if replicate else hi - lo, 4), dtype=torch.float64, device=dev)
        torch.cuda.synchronize(dev)
        res = ctx.cumulative(job, P.data_ptr(), Q.data_ptr(), S.data_ptr() if S is not None else 0)
        if opened is not None:
            dw.release_tables(opened)
        if not replicate:
            m = max(hi - lo, 0)
            P, Q, S = P[:m], Q[:m], (S[:m] if S is not None else None)
        return CumulativeResult(P, Q, S, res, "cuda", dsize, (0, n) if replicate else (lo, hi))

    w = 1 if world is None else int(world)
    if w == 1:
        ctx = parallel.context()
        dev = torch.device("cuda", ctx.device)
        P = torch.empty(n, dtype=torch.float64, device=dev)
        Q = torch.empty(n, dtype=torch.float64, device=dev)
        S = torch.empty(n, dtype=torch.float64, device=dev) if want_samples else None
        torch.cuda.synchronize(dev)
        res = ctx.cumulative(job, P.data_ptr(), Q.data_ptr(), S.data_ptr() if S is not None else 0)
        return CumulativeResult(P, Q, S, res, "cuda", 1, (0, n))

    lw = parallel.local_world(w)
    tiles = (n + 2047) // 2048
    Ps, Qs, slices = [], [], []
    for r in range(w):
        dev = torch.device("cuda", lw.devices[r])
        lo = (r * (8 // w) * tiles // 8) * 2048
        hi = min(n, (((r + 1) * (8 // w)) * tiles // 8) * 2048)
        m = n if replicate else max(hi - lo, 0)      # an empty slice is still a valid buffer (tiny n, many ranks)
        # (torch reports a null data_ptr for an empty view: the kernel gets the base buffer's address)
        Ps.append(torch.empty(max(m, 4), dtype=torch.float64, device=dev))
        Qs.append(torch.empty(max(m, 4), dtype=torch.float64, device=dev))
        slices.append((0, n) if replicate else (lo, hi))
    for r in range(w):
        torch.cuda.synchronize(torch.device("cuda", lw.devices[r]))
    out = lw.cumulative(job, [p.data_ptr() for p in Ps], [q.data_ptr() for q in Qs])
    Ps = [p[:hi_ - lo_] for p, (lo_, hi_) in zip(Ps, slices)]
    Qs = [q[:hi_ - lo_] for q, (lo_, hi_) in zip(Qs, slices)]
    result = CumulativeResult(Ps[0], Qs[0], None, out[w - 1], "cuda", w, slices[0])
    result.all_P, result.all_Q, result.slices, result.per_rank = Ps, Qs, slices, out
    result.device_ms = max(float(o.device_ms) for o in out)
    return result
