"""Integrand library ("model families" of this framework).

Mirrors what the reference integrates:
  sin      riemann.cpp:37, cintegrate.cu:68
  acc/vel/dis  the analytic train model, riemann.cpp:103-116 (declared, never called there)
  profile  the 1801-sample velocity table of ex4vel.h, lerped as faccel() does
           (4main.c:262-268, cintegrate.cu:36-44)

Each entry knows its native enum id (fused-kernel path), a NumPy/Torch
evaluator (slow path, oracles) and, where one exists, an antiderivative.
"""
import ctypes as C
import math

from .. import _native

TSCALE = 286.4788975     # riemann.cpp:7
ASCALE = 0.2365890       # riemann.cpp:8
VSCALE = 67.7777777      # riemann.cpp:9


class Integrand:
    """A named integrand of the native library."""

    def __init__(self, name, native_id, params=(0.0, 0.0, 0.0, 0.0), fn=None, antiderivative=None,
                 domain=None, doc=""):
        self.name = name
        self.native_id = native_id
        self.params = tuple(float(p) for p in params) + (0.0,) * (4 - len(params))
        self.fn = fn
        self.antiderivative = antiderivative
        self.domain = domain
        self.__doc__ = doc

    def __call__(self, x):
        return self.fn(x)

    def exact(self, a, b):
        if self.antiderivative is None:
            raise ValueError("integrand %r has no closed-form antiderivative" % self.name)
        return self.antiderivative(b) - self.antiderivative(a)

    def __repr__(self):
        return "Integrand(%r)" % self.name


def _xp(x):
    try:
        import torch
        if isinstance(x, torch.Tensor):
            return torch
    except Exception:  # pragma: no cover
        pass
    import numpy as np
    return np


def default_profile():
    """The 1801-entry velocity table (t = 0..1800 s, 1 s spacing) as a list of floats."""
    lib, _ = _native.load()
    n = C.c_int(0)
    ptr = lib.bq_default_profile(C.byref(n))
    return [ptr[i] for i in range(n.value)]


def _profile_fn(table=None, spacing=1.0):
    def fn(x):
        xp = _xp(x)
        tab = table if table is not None else default_profile()
        if xp.__name__ == "torch":
            t = xp.as_tensor(tab, dtype=x.dtype, device=x.device)
            u = x / spacing
            i = xp.clamp(u.floor().long(), 0, len(tab) - 2)
            frac = u - i.to(x.dtype)
            return t[i] + (t[i + 1] - t[i]) * frac
        import numpy as np
        t = np.asarray(tab, dtype=np.float64)
        u = np.asarray(x, dtype=np.float64) / spacing
        i = np.clip(np.floor(u).astype(np.int64), 0, len(tab) - 2)
        frac = u - i
        return t[i] + (t[i + 1] - t[i]) * frac
    return fn


def _mk():
    reg = {}
    reg["sin"] = Integrand("sin", 0, fn=lambda x: _xp(x).sin(x),
                           antiderivative=lambda x: -math.cos(x), doc="sin(x)")
    reg["acc"] = Integrand("acc", 1, (TSCALE, ASCALE, VSCALE),
                           fn=lambda x: -(_xp(x).sin(x / TSCALE) * ASCALE),
                           antiderivative=lambda x: ASCALE * TSCALE * math.cos(x / TSCALE),
                           doc="train acceleration, riemann.cpp:103-106")
    reg["vel"] = Integrand("vel", 2, (TSCALE, ASCALE, VSCALE),
                           fn=lambda x: (-_xp(x).cos(x / TSCALE) + 1.0) * VSCALE,
                           antiderivative=lambda x: VSCALE * (x - TSCALE * math.sin(x / TSCALE)),
                           doc="train velocity, riemann.cpp:108-111")
    reg["dis"] = Integrand("dis", 3, (TSCALE, ASCALE, VSCALE),
                           fn=lambda x: VSCALE * (x - TSCALE * _xp(x).sin(x / TSCALE)),
                           antiderivative=lambda x: VSCALE * (0.5 * x * x + TSCALE * TSCALE * math.cos(x / TSCALE)),
                           doc="train position, riemann.cpp:113-116")
    reg["profile"] = Integrand("profile", 4, (1.0,), fn=_profile_fn(), domain=(0.0, 1800.0),
                               doc="lerped ex4vel.h velocity table, 4main.c:262-268")
    return reg


REGISTRY = _mk()


def get(f):
    """Resolve a name / Integrand / callable to (Integrand or None, callable or None)."""
    if isinstance(f, Integrand):
        return f, None
    if isinstance(f, str):
        key = f.lower()
        if key not in REGISTRY:
            raise ValueError("unknown integrand %r (have: %s)" % (f, ", ".join(sorted(REGISTRY))))
        return REGISTRY[key], None
    if callable(f):
        return None, f
    raise TypeError("f must be an integrand name, an Integrand or a callable")


def names():
    return sorted(REGISTRY)


def custom(expr, params=(0.0, 0.0, 0.0, 0.0), name=None):
    """An integrand from a C++ expression in `x` and `p0..p3`, compiled into the fused kernel
    (see models/jit.py).  The build happens on first use and is cached on disk."""
    from . import jit as _custom
    expr = str(expr).strip()
    integ = Integrand(name or "custom:" + expr, _native.INTEGRANDS["custom"], params,
                      doc="custom integrand: " + expr)
    integ.expr = expr
    ev = None

    def fn(x):
        # slow path / oracles: element-wise through the compiled host evaluator
        nonlocal ev
        if ev is None:
            ev = _custom.evaluator(expr, integ.params)
        xp = _xp(x)
        if xp.__name__ == "torch":
            flat = [ev(v) for v in x.detach().cpu().double().reshape(-1).tolist()]
            return xp.tensor(flat, dtype=x.dtype, device=x.device).reshape(x.shape)
        import numpy as np
        return np.vectorize(ev)(np.asarray(x, dtype=np.float64))
    integ.fn = fn
    return integ
