"""ctypes view of the C ABI declared in include/b200quad.h.

The CUDA build (libb200quad.so) is used whenever it loads and a GPU is present;
on GPU-less boxes the CPU-only build (libb200quad_cpu.so) provides the same
host entry points.  On a machine that has a GPU the CUDA library MUST load:
a silent CPU fallback there would hide a broken build.
"""
import ctypes as C
import os

_LIBDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")

SEGMENTS = 8
MAX_WORLD = 8
IPC_HANDLE_BYTES = 64

INTEGRANDS = {"sin": 0, "acc": 1, "vel": 2, "dis": 3, "profile": 4, "custom": 5}
RULES = {"left": 0, "right": 1, "mid": 2, "midpoint": 2, "trap": 3, "trapezoid": 3}
DTYPES = {"fp64": 0, "f64": 0, "float64": 0, "fp32": 1, "f32": 1, "fp32c": 1, "float32": 1}
EVALS = {"auto": 0, "direct": 1, "rotate": 2}
EVAL_NAMES = {0: "auto", 1: "direct", 2: "rotate"}
STATUS = {0: "ok", 1: "invalid argument", 2: "CUDA error", 3: "peer time-out", 4: "unsupported",
          5: "outside integrand range"}


class Job(C.Structure):
    _fields_ = [("integrand", C.c_int32), ("rule", C.c_int32), ("dtype", C.c_int32),
                ("eval", C.c_int32), ("a", C.c_double), ("b", C.c_double), ("n", C.c_uint64),
                ("p", C.c_double * 4)]


class Result(C.Structure):
    _fields_ = [("value", C.c_double), ("value_lo", C.c_double), ("raw_sum_hi", C.c_double),
                ("raw_sum_lo", C.c_double), ("seg_hi", C.c_double * SEGMENTS),
                ("seg_lo", C.c_double * SEGMENTS), ("t_start_ns", C.c_uint64),
                ("t_end_ns", C.c_uint64), ("n", C.c_uint64), ("status", C.c_int32),
                ("eval_used", C.c_int32), ("world", C.c_int32), ("rank", C.c_int32),
                ("device_ms", C.c_float), ("launches", C.c_int32), ("seq", C.c_uint64),
                ("reserved", C.c_uint64), ("stamp_ns", C.c_uint32 * 8)]


STAMPS = ("prologue", "tiles", "ticket", "fold", "published", "barrier", "result", "spins")


class ScanJob(C.Structure):
    _fields_ = [("integrand", C.c_int32), ("replicate", C.c_int32), ("want_samples", C.c_int32),
                ("reserved", C.c_int32), ("x0", C.c_double), ("dx", C.c_double), ("n", C.c_uint64),
                ("p", C.c_double * 4)]


class ScanResult(C.Structure):
    _fields_ = [("p_last", C.c_double), ("q_last", C.c_double), ("p_second_last", C.c_double),
                ("slice_begin", C.c_uint64), ("slice_end", C.c_uint64), ("t_start_ns", C.c_uint64),
                ("t_end_ns", C.c_uint64), ("status", C.c_int32), ("world", C.c_int32),
                ("rank", C.c_int32), ("device_ms", C.c_float), ("reserved", C.c_uint64 * 2)]


def _gpu_present():
    if os.environ.get("B200QUAD_FORCE_CPU") == "1":
        return False
    return os.path.exists("/dev/nvidiactl") or os.path.exists("/dev/nvidia0")


def _declare(lib, cuda):
    vp, i32, u64, dbl = C.c_void_p, C.c_int, C.c_uint64, C.c_double
    P = C.POINTER

    def sig(name, res, *args):
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = list(args)

    sig("bq_version", C.c_char_p)
    sig("bq_last_error", C.c_char_p)
    sig("bq_default_profile", P(dbl), P(i32))
    sig("bq_oracle_sin_sum", dbl, dbl, dbl, u64, i32)
    sig("bq_sizeof_params", C.c_size_t)
    sig("bq_sin_bounded", dbl, dbl)
    sig("bq_cos_bounded", dbl, dbl)
    sig("bq_set_custom_integrand", i32, vp, vp)
    sig("bq_register_custom_integrand", i32, vp, vp)
    sig("bq_cpu_integrate", i32, P(Job), i32, P(Result))
    sig("bq_cpu_integrate_segments", i32, P(Job), i32, i32, i32, P(Result))
    sig("bq_fold_segments", i32, P(Job), P(dbl), P(dbl), P(Result))
    sig("bq_cpu_cumulative", i32, P(ScanJob), i32, vp, vp, vp, P(ScanResult))
    sig("bq_cpu_scan_totals", i32, P(ScanJob), i32, i32, i32, P(dbl))
    sig("bq_cpu_scan_finish", i32, P(ScanJob), i32, i32, i32, P(dbl), vp, vp, vp, P(ScanResult))
    if not cuda:
        return
    sig("bq_cuda_device_count", i32)
    sig("bq_ctx_create", i32, i32, P(vp))
    sig("bq_ctx_destroy", None, vp)
    sig("bq_ctx_device", i32, vp)
    sig("bq_ctx_stream", vp, vp)
    sig("bq_ctx_set_profile", i32, vp, P(dbl), i32)
    sig("bq_ctx_configure", i32, vp, i32, i32, i32)
    sig("bq_ctx_set_timeout_ms", i32, vp, i32)
    sig("bq_integrate", i32, vp, P(Job), P(Result))
    sig("bq_integrate_value", i32, vp, P(Job), P(dbl), P(Result))
    sig("bq_integrate_async", i32, vp, P(Job), vp)
    sig("bq_integrate_device_async", i32, vp, P(Job), vp, vp)
    sig("bq_device_result_fetch", i32, vp, P(Result))
    sig("bq_result_fetch", i32, vp, P(Result))
    sig("bq_ctx_device_result", vp, vp)
    sig("bq_cumulative", i32, vp, P(ScanJob), vp, vp, vp, P(ScanResult))
    sig("bq_cumulative_async", i32, vp, P(ScanJob), vp, vp, vp, vp)
    sig("bq_scan_result_fetch", i32, vp, P(ScanResult))
    sig("bq_ctx_reserve_scan", i32, vp, u64)
    sig("bq_noalloc_max_samples", u64)
    sig("bq_cumulative_noalloc", i32, i32, P(ScanJob), vp, vp, vp, vp)
    sig("bq_ctx_export_comm", i32, vp, C.c_char_p)
    sig("bq_ctx_attach_peers", i32, vp, i32, i32, C.c_char_p)
    sig("bq_ctx_reset_comm", i32, vp)
    sig("bq_ctx_adopt_comm", i32, vp, i32, i32, P(vp), vp, C.c_size_t)
    sig("bq_ctx_set_nvls", i32, vp, i32)
    sig("bq_ctx_set_scan_multicast", i32, vp, vp, vp)
    sig("bq_default_ctx", vp, i32)
    sig("bq_default_world", vp, i32)
    sig("bq_set_dist_ctx", i32, vp)
    sig("bq_dist_ctx", vp)
    sig("bq_ctx_world", i32, vp)
    sig("bq_ctx_rank", i32, vp)
    sig("bq_world_has_multicast", i32, vp)
    sig("bq_world_set_nvls", i32, vp, i32)
    sig("bq_world_alloc_tables", i32, vp, u64, P(vp), P(vp))
    sig("bq_world_free_tables", None, vp)
    sig("bq_ipc_export", i32, vp, C.c_char_p)
    sig("bq_ipc_export_at", i32, vp, C.c_char_p, P(u64))
    sig("bq_ipc_open", i32, C.c_char_p, P(vp))
    sig("bq_ipc_close", i32, vp)
    sig("bq_ctx_set_scan_peers", i32, vp, P(vp), P(vp))
    sig("bq_world_create", i32, i32, P(i32), P(vp))
    sig("bq_world_destroy", None, vp)
    sig("bq_world_size", i32, vp)
    sig("bq_world_ctx", vp, vp, i32)
    sig("bq_world_integrate", i32, vp, P(Job), P(Result))
    sig("bq_world_cumulative", i32, vp, P(ScanJob), P(vp), P(vp), P(ScanResult))
    sig("bq_integrate_local_async", i32, vp, P(Job), i32, i32, vp, vp)
    sig("bq_ctx_inject_fault", i32, vp, i32)
    sig("bq_ctx_trace", i32, vp, i32)
    sig("bq_ctx_trace_read", i32, vp, P(u64), i32)


class NativeError(RuntimeError):
    pass


_lib = None
_is_cuda = False
_load_note = ""


def load():
    """Returns (lib, is_cuda)."""
    global _lib, _is_cuda, _load_note
    if _lib is not None:
        return _lib, _is_cuda
    # B200QUAD_LIB: load another build of the core library (tuning runs compare variants)
    cuda_path = os.environ.get("B200QUAD_LIB") or os.path.join(_LIBDIR, "libb200quad.so")
    # B200QUAD_CPU_LIB: another build of the CPU library (the sanitizer tier loads its ASan/UBSan build)
    cpu_path = os.environ.get("B200QUAD_CPU_LIB") or os.path.join(_LIBDIR, "libb200quad_cpu.so")
    if not os.path.exists(cuda_path) and not os.path.exists(cpu_path):
        raise NativeError("native libraries are not built: run `python build.py` "
                          "(or __graft_entry__.build()) in the repository root")
    err = None
    if os.environ.get("B200QUAD_FORCE_CPU") != "1":
        try:
            lib = C.CDLL(cuda_path, mode=C.RTLD_GLOBAL)
            _declare(lib, True)
            _lib, _is_cuda = lib, True
            return _lib, _is_cuda
        except OSError as e:   # e.g. libcudart missing on a CPU-only host
            err = e
    if _gpu_present():
        raise NativeError("a GPU is present but %s failed to load: %s" % (cuda_path, err))
    lib = C.CDLL(cpu_path, mode=C.RTLD_GLOBAL)
    _declare(lib, False)
    _lib, _is_cuda = lib, False
    _load_note = str(err) if err else "forced"
    return _lib, _is_cuda


def last_error():
    lib, _ = load()
    return lib.bq_last_error().decode()


def check(rc, what):
    if rc != 0:
        raise NativeError("%s failed: %s (%s)" % (what, STATUS.get(rc, rc), last_error()))


def cuda_device_count():
    lib, cuda = load()
    if not cuda:
        return 0
    return int(lib.bq_cuda_device_count())
