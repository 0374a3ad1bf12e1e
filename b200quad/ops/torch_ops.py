"""torch.ops.b200quad.* — the ops registered by csrc/torch_ops.cpp."""
import os

import torch

from .. import _native, models

_loaded = False


def load():
    global _loaded
    if not _loaded:
        _native.load()   # core library first (RTLD_GLOBAL) so the op library resolves against it
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib",
                            "libb200quad_torch.so")
        if not os.path.exists(path):
            raise _native.NativeError("libb200quad_torch.so is not built: run `python build.py`")
        torch.ops.load_library(path)
        _loaded = True
    return torch.ops.b200quad


def _native_id(integ):
    if getattr(integ, "expr", None) is not None:
        from ..models import jit as _custom
        return _custom.integrand_code(integ.expr)
    return integ.native_id


def integrate(f, a, b, n, rule="left", dtype="fp64", eval="auto", device=None, world=1, params=None):
    """0-dim float64 tensor; on CUDA the work is enqueued on torch's current stream.
    world > 1: this rank's part of the collective call under torch.distributed (create
    b200quad.parallel.dist_world() first), or the process's world of that many GPUs."""
    ops = load()
    integ, _ = models.get(f)
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else -1
    if params is None:
        params = integ.params if getattr(integ, "expr", None) is not None else ()
    return ops.integrate(_native_id(integ), float(a), float(b), int(n), _native.RULES[rule],
                         _native.DTYPES[dtype], _native.EVALS[eval], int(device), int(world),
                         [float(p) for p in params])


def cumulative(f, x0, dx, n, device=None, world=1, params=None):
    ops = load()
    integ, _ = models.get(f)
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else -1
    return ops.cumulative(integ.native_id, float(x0), float(dx), int(n), int(device), int(world),
                          [float(p) for p in (params or ())])
