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


def test_integrate_op_params_cpu():
    # vel = (1 - cos(x / tscale)) * vscale with tscale = 2, vscale = 3 passed through the op
    t = torch_ops.integrate("vel", 0.0, 4.0, 200_000, "mid", device=-1, params=(2.0, 0.0, 3.0))
    exact = 3.0 * (4.0 - 2.0 * math.sin(2.0))
    assert abs(float(t) - exact) < 1e-9


import pytest  # noqa: E402


@pytest.mark.gpu
def test_ops_share_the_process_context_and_take_world():
    """The op resolves to the same per-device context as the Python layer (a table set there is the
    table the op integrates) and accepts the world size as an input (riemann.cpp:63-65)."""
    import b200quad as bq
    from b200quad import parallel
    from conftest import gpu_count
    ctx = parallel.context(0)
    tab = [float(i * i) for i in range(101)]
    ctx.set_profile(tab)
    try:
        t = torch_ops.integrate("profile", 0.0, 100.0, 100_000, "trap", device=0)
        exact = sum((tab[i] + tab[i + 1]) / 2 for i in range(100))
        assert abs(float(t) - exact) / exact < 1e-13
    finally:
        ctx.set_profile(bq.models.default_profile())
    one = float(torch_ops.integrate("sin", 0.0, math.pi, (1 << 24) + 3, device=0))
    if gpu_count() >= 2:
        two = torch_ops.integrate("sin", 0.0, math.pi, (1 << 24) + 3, device=0, world=2)
        assert two.is_cuda and float(two) == one
    v = torch_ops.integrate("vel", 0.0, 4.0, 1 << 22, "mid", device=0, params=(2.0, 0.0, 3.0))
    assert abs(float(v) - 3.0 * (4.0 - 2.0 * math.sin(2.0))) < 1e-11
