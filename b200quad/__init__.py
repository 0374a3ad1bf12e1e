"""b200quad — Blackwell-native numerical integration.

Same capabilities as Excalibur1224/Cuda-v-MPI (definite integrals by Riemann
sums, cumulative integration of the train velocity profile, on many workers),
re-designed for B200: one fused sm_100a kernel per GPU per call, the cross-GPU
combine done inside that kernel over NVLink peer memory.

    import b200quad as bq
    bq.integrate("sin", 0.0, math.pi, 2**32)              # left rule, this GPU
    bq.integrate("sin", 0.0, math.pi, 2**32, world=8)     # 8 GPUs, one process
    bq.integrate("profile", 0, 1800, 18_000_000, "trap")  # the ex4vel.h table
    r = bq.cumulative_integrate()                         # 4main.c's two running sums
"""
from . import _native, models, parallel, utils
from .ops import (integrate, integrate_result, cumulative_integrate, IntegrateResult, backend,
                  make_job)
from ._native import NativeError

__version__ = "0.1.0"
__all__ = ["integrate", "integrate_result", "cumulative_integrate", "IntegrateResult", "backend",
           "make_job", "models", "parallel", "utils", "NativeError"]
