"""Worlds: who takes part in a call and how their partial results meet.

The reference spreads work over MPI ranks and combines with MPI_Send/Recv,
MPI_Reduce, MPI_Barrier and MPI_Bcast (riemann.cpp:62-85, 4main.c:69-221).
Here a *world* is a set of GPUs whose fused kernels publish partial results
into each other's memory over NVLink and meet at an in-kernel barrier:

  Context      one GPU of this process (world of 1)
  LocalWorld   several GPUs driven by this one process (peer access through UVA)
  DistWorld    one GPU per process under torchrun; torch.distributed (NCCL) is
               only the bootstrap that carries the CUDA-IPC handles around
  CpuWorld     torch.distributed (gloo) ranks on the CPU backend: each computes
               its segments of the fixed tree, an all_gather moves 8 pairs
"""
import ctypes as C
import os
import threading

from .. import _native
from .._native import Job, Result, ScanJob, ScanResult, NativeError, check

_lock = threading.Lock()
_contexts = {}
_local_worlds = {}
_dist_world = None


class Context:
    """Owns every device buffer the fused kernels need on one GPU."""

    def __init__(self, device=0):
        lib, cuda = _native.load()
        if not cuda or lib.bq_cuda_device_count() == 0:
            raise NativeError("no CUDA device available for a b200quad Context")
        self.lib = lib
        self.device = int(device)
        h = C.c_void_p()
        check(lib.bq_ctx_create(self.device, C.byref(h)), "bq_ctx_create")
        self.handle = h
        self.rank, self.world = 0, 1
        self._owned = True

    @classmethod
    def _wrap(cls, lib, handle, device, rank, world):
        self = cls.__new__(cls)
        self.lib, self.handle, self.device = lib, handle, device
        self.rank, self.world = rank, world
        self._owned = False
        return self

    # -- definite integral --------------------------------------------------
    def integrate(self, job):
        res = Result()
        rc = self.lib.bq_integrate(self.handle, C.byref(job), C.byref(res))
        check(rc, "bq_integrate")
        return res

    def integrate_value(self, job):
        """The integral as a float, lowest latency: one launch, no events, the host watches the
        completion word the kernel writes last into the pinned result block."""
        val = C.c_double()
        rc = self.lib.bq_integrate_value(self.handle, C.byref(job), C.byref(val), None)
        if rc:
            check(rc, "bq_integrate_value")
        return val.value

    def integrate_async(self, job, stream=None):
        check(self.lib.bq_integrate_async(self.handle, C.byref(job), C.c_void_p(stream or 0)),
              "bq_integrate_async")

    def fetch(self):
        res = Result()
        check(self.lib.bq_result_fetch(self.handle, C.byref(res)), "bq_result_fetch")
        return res

    def integrate_device_async(self, job, d_value, stream=None):
        """One kernel; the integral is stored at the device address d_value, nothing goes to the host."""
        check(self.lib.bq_integrate_device_async(self.handle, C.byref(job), C.c_void_p(d_value),
                                                 C.c_void_p(stream or 0)), "bq_integrate_device_async")

    def fetch_device(self):
        res = Result()
        check(self.lib.bq_device_result_fetch(self.handle, C.byref(res)), "bq_device_result_fetch")
        return res

    # -- cumulative -----------------------------------------------------------
    def cumulative(self, job, dP, dQ, dS=0):
        res = ScanResult()
        rc = self.lib.bq_cumulative(self.handle, C.byref(job), C.c_void_p(dP), C.c_void_p(dQ),
                                    C.c_void_p(dS or 0), C.byref(res))
        check(rc, "bq_cumulative")
        return res

    # -- knobs ------------------------------------------------------------------
    def configure(self, ctas_per_sm=0, cluster=0, rows_per_tile=0):
        check(self.lib.bq_ctx_configure(self.handle, ctas_per_sm, cluster, rows_per_tile),
              "bq_ctx_configure")

    def set_timeout_ms(self, ms):
        check(self.lib.bq_ctx_set_timeout_ms(self.handle, int(ms)), "bq_ctx_set_timeout_ms")

    def set_profile(self, table):
        arr = (C.c_double * len(table))(*table)
        check(self.lib.bq_ctx_set_profile(self.handle, arr, len(table)), "bq_ctx_set_profile")

    def inject_fault(self, kind=1):
        check(self.lib.bq_ctx_inject_fault(self.handle, int(kind)), "bq_ctx_inject_fault")

    def trace(self, enable=True):
        """Per-CTA timeline of the quad kernel (start ns, end ns, SM id, tiles)."""
        check(self.lib.bq_ctx_trace(self.handle, 1 if enable else 0), "bq_ctx_trace")

    def read_trace(self):
        buf = (C.c_uint64 * (4 * 4096))()
        n = self.lib.bq_ctx_trace_read(self.handle, buf, 4096)
        return [tuple(buf[4 * i + k] for k in range(4)) for i in range(n)]

    def reset_comm(self):
        check(self.lib.bq_ctx_reset_comm(self.handle), "bq_ctx_reset_comm")

    @property
    def stream(self):
        return self.lib.bq_ctx_stream(self.handle)

    def close(self):
        if getattr(self, "_owned", False) and self.handle:
            self.lib.bq_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


def context(device=None):
    """The cached Context of a device (default: torch's current CUDA device, else 0)."""
    if device is None:
        device = 0
        try:
            import torch
            if torch.cuda.is_available():
                device = torch.cuda.current_device()
        except Exception:
            pass
    with _lock:
        if device not in _contexts:
            # the process-wide context of the device (shared with the torch ops), not owned here
            lib, cuda = _native.load()
            if not cuda or lib.bq_cuda_device_count() == 0:
                raise NativeError("no CUDA device available for a b200quad Context")
            h = lib.bq_default_ctx(int(device))
            if not h:
                raise NativeError("bq_default_ctx(%d) failed: %s" % (device, _native.last_error()))
            _contexts[device] = Context._wrap(lib, C.c_void_p(h), int(device), 0, 1)
        return _contexts[device]


class LocalWorld:
    """`size` GPUs of this process; one launch per GPU, one host thread."""

    def __init__(self, size, devices=None, _shared=False):
        lib, cuda = _native.load()
        if not cuda:
            raise NativeError("LocalWorld needs the CUDA library")
        self.lib = lib
        self.size = int(size)
        devs = list(devices) if devices is not None else list(range(self.size))
        self._owned = not _shared
        if _shared:
            # the process-wide world of this size (devices 0..size-1), shared with the torch ops
            h = lib.bq_default_world(self.size)
            if not h:
                raise NativeError("bq_default_world(%d) failed: %s" % (self.size, _native.last_error()))
            h = C.c_void_p(h)
        else:
            arr = (C.c_int * self.size)(*devs)
            h = C.c_void_p()
            check(lib.bq_world_create(self.size, arr, C.byref(h)), "bq_world_create")
        self.handle = h
        self.devices = devs
        self.contexts = [Context._wrap(lib, C.c_void_p(lib.bq_world_ctx(h, r)), devs[r], r, self.size)
                         for r in range(self.size)]

    def integrate(self, job):
        out = (Result * self.size)()
        check(self.lib.bq_world_integrate(self.handle, C.byref(job), out), "bq_world_integrate")
        return list(out)

    def cumulative(self, job, dP, dQ):
        out = (ScanResult * self.size)()
        aP = (C.c_void_p * self.size)(*dP)
        aQ = (C.c_void_p * self.size)(*dQ)
        check(self.lib.bq_world_cumulative(self.handle, C.byref(job), aP, aQ, out),
              "bq_world_cumulative")
        return list(out)

    # -- NVLS (NVSwitch multicast) ---------------------------------------------------
    @property
    def has_multicast(self):
        return bool(self.lib.bq_world_has_multicast(self.handle))

    def set_nvls(self, on=True):
        """Publish through the multicast address (one multimem.st per word / per 16 bytes of a
        replicated table) instead of one store per peer.  Same results either way."""
        check(self.lib.bq_world_set_nvls(self.handle, 1 if on else 0), "bq_world_set_nvls")

    def alloc_tables(self, n):
        """P and Q tables of n doubles on every rank, in symmetric multicast-bound memory; returns
        two lists of device addresses (rank order).  Freed by free_tables() or with the world."""
        aP = (C.c_void_p * self.size)()
        aQ = (C.c_void_p * self.size)()
        check(self.lib.bq_world_alloc_tables(self.handle, int(n), aP, aQ), "bq_world_alloc_tables")
        return [int(p) for p in aP], [int(q) for q in aQ]

    def free_tables(self):
        self.lib.bq_world_free_tables(self.handle)

    def close(self):
        if self.handle and self._owned:
            self.lib.bq_world_destroy(self.handle)
        self.handle = None


def device_tensor(ptr, n, device):
    """A float64 torch tensor over `n` doubles at device address `ptr` (memory owned by the library)."""
    import torch

    class _Raw:
        __cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False), "version": 3}
    return torch.as_tensor(_Raw(), device=torch.device("cuda", device))


def local_world(size):
    with _lock:
        if size not in _local_worlds:
            _local_worlds[size] = LocalWorld(size, _shared=True)
        return _local_worlds[size]


class DistWorld:
    """One GPU per process (torchrun).  torch.distributed moves only the IPC handles."""

    def __init__(self, group=None, device=None):
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            raise NativeError("torch.distributed is not initialised")
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.size = dist.get_world_size(group)
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", self.rank % max(torch.cuda.device_count(), 1)))
        torch.cuda.set_device(device)
        # A context of its own: attaching peers turns a context into rank r of a world of N for good
        # (its kernels then publish into the peers' comm blocks), so the process-wide world-of-1
        # context that single-GPU calls use (parallel.context()) must stay untouched.
        self.ctx = Context(device)
        lib = self.ctx.lib
        self.transport = None
        self._symm = []                       # keeps the symmetric tensors (and their handles) alive
        if os.environ.get("BQ_SYMM", "0") == "1":     # opt-in: CUDA IPC is the default wiring
            self._wire_symmetric(device)
        if self.transport is None:
            mine = C.create_string_buffer(_native.IPC_HANDLE_BYTES)
            check(lib.bq_ctx_export_comm(self.ctx.handle, mine), "bq_ctx_export_comm")
            gathered = [None] * self.size
            dist.all_gather_object(gathered, bytes(mine.raw), group=group)
            blob = b"".join(gathered)
            check(lib.bq_ctx_attach_peers(self.ctx.handle, self.rank, self.size, blob),
                  "bq_ctx_attach_peers")
            self.transport = "cuda-ipc"
        self.ctx.rank, self.ctx.world = self.rank, self.size
        lib.bq_set_dist_ctx(self.ctx.handle)      # the torch ops' world > 1 path uses this context
        dist.barrier(group=group)

    def _symm_alloc(self, nbytes, device):
        """(tensor, handle) of a torch symmetric-memory allocation, rendezvoused over the group."""
        import torch
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(int(nbytes), dtype=torch.uint8, device=torch.device("cuda", device))
        group = self.group if self.group is not None else self.dist.group.WORLD
        hdl = symm_mem.rendezvous(t, group)
        return t, hdl

    def _wire_symmetric(self, device):
        """Comm blocks in torch symmetric memory (cuMem allocations exchanged by the process group,
        bound to an NVLS multicast object when the fabric has one): peers are plain mapped pointers
        and, with multicast, one multimem.st reaches every rank.  Falls back to CUDA IPC when the
        build or the box cannot do it; every rank takes the same branch (the decision is agreed)."""
        import torch
        ok, t, hdl = 1, None, None
        try:
            t, hdl = self._symm_alloc(4096, device)
            t.zero_()
            torch.cuda.synchronize(device)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            if len(ptrs) != self.size or any(p == 0 for p in ptrs):
                ok = 0
        except Exception:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=torch.device("cuda", device))
        self.dist.all_reduce(flag, op=self.dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) != 1:
            return
        arr = (C.c_void_p * _native.MAX_WORLD)(*ptrs)
        check(self.ctx.lib.bq_ctx_adopt_comm(self.ctx.handle, self.rank, self.size, arr, C.c_void_p(mc or None), 4096),
              "bq_ctx_adopt_comm")
        self._symm.append((t, hdl))
        self.has_multicast = mc != 0
        self.transport = "symmetric-memory" + ("+multicast" if mc else "")

    has_multicast = False

    def set_nvls(self, on=True):
        check(self.ctx.lib.bq_ctx_set_nvls(self.ctx.handle, 1 if on else 0), "bq_ctx_set_nvls")

    def symmetric_tables(self, n):
        """Full-size P and Q tables in symmetric memory: returns (P, Q) float64 tensors; every
        rank's copies are wired as scan peers (and as multicast targets when available)."""
        import torch
        device = self.ctx.device
        out = []
        peers, mcs = [], []
        for _ in range(2):
            t, hdl = self._symm_alloc(int(n) * 8, device)
            self._symm.append((t, hdl))
            out.append(t.view(torch.float64))
            peers.append([int(p) for p in hdl.buffer_ptrs])
            mcs.append(int(getattr(hdl, "multicast_ptr", 0) or 0))
        aP = (C.c_void_p * _native.MAX_WORLD)(*peers[0])
        aQ = (C.c_void_p * _native.MAX_WORLD)(*peers[1])
        check(self.ctx.lib.bq_ctx_set_scan_peers(self.ctx.handle, aP, aQ), "bq_ctx_set_scan_peers")
        check(self.ctx.lib.bq_ctx_set_scan_multicast(self.ctx.handle, C.c_void_p(mcs[0] or None),
                                                     C.c_void_p(mcs[1] or None)), "bq_ctx_set_scan_multicast")
        self.dist.barrier(group=self.group)
        return out[0], out[1]

    def integrate(self, job):
        return self.ctx.integrate(job)

    def reset(self):
        self.ctx.reset_comm()
        self.dist.barrier(group=self.group)

    def share_tables(self, tP, tQ):
        """Map every rank's full-size P/Q tables here (replicating scans)."""
        lib = self.ctx.lib
        hp = C.create_string_buffer(_native.IPC_HANDLE_BYTES)
        hq = C.create_string_buffer(_native.IPC_HANDLE_BYTES)
        # A cudaIpc handle names the whole cudaMalloc segment; torch's caching allocator hands out
        # tensors INSIDE segments (two small tables usually share one), so each handle travels with
        # the tensor's offset from its segment base (cuMemGetAddressRange in bq_ipc_export_at).
        op_, oq_ = C.c_uint64(), C.c_uint64()
        check(lib.bq_ipc_export_at(C.c_void_p(tP.data_ptr()), hp, C.byref(op_)), "bq_ipc_export_at")
        check(lib.bq_ipc_export_at(C.c_void_p(tQ.data_ptr()), hq, C.byref(oq_)), "bq_ipc_export_at")
        gathered = [None] * self.size
        self.dist.all_gather_object(gathered, (bytes(hp.raw), bytes(hq.raw), int(op_.value), int(oq_.value)),
                                    group=self.group)
        aP = (C.c_void_p * _native.MAX_WORLD)()
        aQ = (C.c_void_p * _native.MAX_WORLD)()
        opened = []
        for r, (bp, bq_, op, oq) in enumerate(gathered):
            if r == self.rank:
                aP[r], aQ[r] = tP.data_ptr(), tQ.data_ptr()
                continue
            pp, pq = C.c_void_p(), C.c_void_p()
            check(lib.bq_ipc_open(bp, C.byref(pp)), "bq_ipc_open")
            opened.append(pp.value)
            if bq_ == bp:
                pq = pp
            else:
                check(lib.bq_ipc_open(bq_, C.byref(pq)), "bq_ipc_open")
                opened.append(pq.value)
            aP[r], aQ[r] = pp.value + op, pq.value + oq
        check(lib.bq_ctx_set_scan_peers(self.ctx.handle, aP, aQ), "bq_ctx_set_scan_peers")
        self.dist.barrier(group=self.group)
        return opened

    def release_tables(self, opened):
        self.dist.barrier(group=self.group)
        for p in opened:
            self.ctx.lib.bq_ipc_close(C.c_void_p(p))
        self.ctx.lib.bq_ctx_set_scan_peers(self.ctx.handle, None, None)


def dist_world():
    """The DistWorld over the default process group (created on first use)."""
    global _dist_world
    if _dist_world is None:
        _dist_world = DistWorld()
    return _dist_world


class CpuWorld:
    """gloo ranks (or a single process) on the CPU backend."""

    def __init__(self, group=None, threads=None):
        self.lib, _ = _native.load()
        self.group = group
        self.threads = threads or max(1, (os.cpu_count() or 1))
        try:
            import torch.distributed as dist
            self.dist = dist if dist.is_initialized() else None
        except Exception:
            self.dist = None
        self.rank = self.dist.get_rank(group) if self.dist else 0
        self.size = self.dist.get_world_size(group) if self.dist else 1

    def integrate(self, job):
        res = Result()
        if self.size == 1:
            check(self.lib.bq_cpu_integrate(C.byref(job), self.threads, C.byref(res)),
                  "bq_cpu_integrate")
            return res
        import torch
        check(self.lib.bq_cpu_integrate_segments(C.byref(job), self.rank, self.size,
                                                 max(1, self.threads // self.size), C.byref(res)),
              "bq_cpu_integrate_segments")
        mine = torch.tensor([list(res.seg_hi), list(res.seg_lo)], dtype=torch.float64)
        parts = [torch.zeros_like(mine) for _ in range(self.size)]
        self.dist.all_gather(parts, mine, group=self.group)
        segs = _native.SEGMENTS // self.size
        hi = (C.c_double * _native.SEGMENTS)()
        lo = (C.c_double * _native.SEGMENTS)()
        for r, t in enumerate(parts):          # rank r owns segments [r*segs, (r+1)*segs)
            for s in range(r * segs, (r + 1) * segs):
                hi[s], lo[s] = float(t[0, s]), float(t[1, s])
        out = Result()
        check(self.lib.bq_fold_segments(C.byref(job), hi, lo, C.byref(out)), "bq_fold_segments")
        out.world, out.rank = self.size, self.rank
        return out

    def cumulative(self, job, want_samples=False, replicate=True):
        """Running sums over the gloo ranks: every rank scans its own segments, ONE all_gather moves
        the 8 segment totals (32 doubles), every rank finishes its slice with the right bases.  The
        reference gathers whole slices to rank 0, fixes the carries there serially and broadcasts
        144 MB (4main.c:141-157); here slices only travel when replicate=True asks for whole tables.
        Returns (P, Q, samples or None, ScanResult, (lo, hi))."""
        import torch
        n = int(job.n)
        res = ScanResult()
        if self.size == 1:
            P = torch.empty(n, dtype=torch.float64)
            Q = torch.empty(n, dtype=torch.float64)
            S = torch.empty(n, dtype=torch.float64) if want_samples else None
            check(self.lib.bq_cpu_cumulative(C.byref(job), self.threads, C.c_void_p(P.data_ptr()),
                                             C.c_void_p(Q.data_ptr()), C.c_void_p(S.data_ptr() if S is not None else 0),
                                             C.byref(res)), "bq_cpu_cumulative")
            return P, Q, S, res, (0, n)
        threads = max(1, self.threads // self.size)
        mine = torch.zeros(32, dtype=torch.float64)
        buf = (C.c_double * 32)()
        check(self.lib.bq_cpu_scan_totals(C.byref(job), self.rank, self.size, threads, buf), "bq_cpu_scan_totals")
        mine[:] = torch.tensor(list(buf), dtype=torch.float64)
        parts = [torch.zeros_like(mine) for _ in range(self.size)]
        self.dist.all_gather(parts, mine, group=self.group)
        segs = _native.SEGMENTS // self.size
        allt = (C.c_double * 32)()
        for r, t in enumerate(parts):
            for k in range(4 * r * segs, 4 * (r + 1) * segs):
                allt[k] = float(t[k])
        tiles = (n + 2047) // 2048
        lo = (self.rank * segs * tiles // 8) * 2048
        hi = min(n, ((self.rank + 1) * segs * tiles // 8) * 2048)
        m = max(hi - lo, 0)
        P = torch.empty(max(m, 1), dtype=torch.float64)
        Q = torch.empty(max(m, 1), dtype=torch.float64)
        S = torch.empty(max(m, 1), dtype=torch.float64) if want_samples else None
        check(self.lib.bq_cpu_scan_finish(C.byref(job), self.rank, self.size, threads, allt, C.c_void_p(P.data_ptr()),
                                          C.c_void_p(Q.data_ptr()), C.c_void_p(S.data_ptr() if S is not None else 0),
                                          C.byref(res)), "bq_cpu_scan_finish")
        P, Q = P[:m], Q[:m]
        S = S[:m] if S is not None else None
        # the three scalars of the result live on the rank that holds the end of the grid: share them
        fin = torch.tensor([res.p_last, res.q_last, res.p_second_last, 1.0 if hi == n and m > 0 else 0.0],
                           dtype=torch.float64)
        fins = [torch.zeros_like(fin) for _ in range(self.size)]
        self.dist.all_gather(fins, fin, group=self.group)
        owner = max(r for r, f in enumerate(fins) if float(f[3]) == 1.0)
        res.p_last, res.q_last = float(fins[owner][0]), float(fins[owner][1])
        res.p_second_last = float(fins[owner][2])
        if n >= 2 and float(fins[owner][2]) == 0.0 and owner > 0:      # P[n-2] sits at the end of the previous slice
            prev_last = torch.tensor([float(P[-1]) if m > 0 else 0.0], dtype=torch.float64)
            lasts = [torch.zeros_like(prev_last) for _ in range(self.size)]
            self.dist.all_gather(lasts, prev_last, group=self.group)
            res.p_second_last = float(lasts[owner - 1][0])
        if not replicate:
            return P, Q, S, res, (lo, hi)
        # whole tables everywhere (the MPI_Bcast semantics, 4main.c:157,221): rank by rank broadcast
        fullP = torch.empty(n, dtype=torch.float64)
        fullQ = torch.empty(n, dtype=torch.float64)
        fullS = torch.empty(n, dtype=torch.float64) if want_samples else None
        for r in range(self.size):
            rlo = (r * segs * tiles // 8) * 2048
            rhi = min(n, ((r + 1) * segs * tiles // 8) * 2048)
            if rhi <= rlo:
                continue
            for full, part in ((fullP, P), (fullQ, Q), (fullS, S)):
                if full is None:
                    continue
                if r == self.rank:
                    full[rlo:rhi] = part
                self.dist.broadcast(full[rlo:rhi], src=r if self.group is None else self.dist.get_global_rank(self.group, r),
                                    group=self.group)
        return fullP, fullQ, fullS, res, (0, n)
