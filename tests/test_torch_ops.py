import math

import torch

from b200quad.ops import torch_ops
from b200quad.utils import oracle_sin


def test_integrate_op_cpu():
    t = torch_ops.integrate("sin", 0.0, math.pi, 100_000, "mid", device=-1)
    assert t.dtype == torch.float64 and t.dim() == 0
    assert abs(float(t) - oracle_sin(0.0, math.pi, 100_000, "mid")) < 1e-15


def test_cumulative_op_cpu():
    P, Q = torch_ops.cumulative("profile", 0.0, 0.01, 180_000, device=-1)
    assert P.shape == (180_000,) and abs(float(P[-1]) * 0.01 - 122000.004) < 1e-6
    assert float(Q[-1]) > float(P[-1])


def test_op_schema_registered():
    ops = torch_ops.load()
    assert "integrate" in dir(ops) or hasattr(ops, "integrate")
