"""CPU backend of cumulative_integrate() (capability of 4main.c:76-225)."""
import math

import pytest
import torch

import b200quad as bq


from oracle_prefix import exact_prefix, max_rel_error


def _oracle(name, x0, dx, n):
    """Samples from the plain PyTorch fp64 model; P and Q = exact running sums of the samples the
    library stored (oracle_prefix.py), so the comparison judges the summation alone."""
    x = x0 + torch.arange(n, dtype=torch.float64) * dx
    return bq.models.REGISTRY[name](x)


@pytest.mark.parametrize("n", [1, 2, 5, 2047, 2048, 2049, 100_003])
@pytest.mark.parametrize("name", ["profile", "vel", "sin"])
def test_tables_match_cumsum(name, n):
    dx = 1800.0 / max(n, 2) if name != "sin" else math.pi / max(n, 2)
    r = bq.cumulative_integrate(name, 0.0, dx, n, want_samples=True, backend="cpu")
    v = _oracle(name, 0.0, dx, n)
    assert float((r.samples - v).abs().max()) <= 1e-13 * max(float(v.abs().max()), 1.0)
    P, Q = exact_prefix(r.samples.numpy())
    assert max_rel_error(r.P.numpy(), P, floor=1.0) <= 1e-15
    assert max_rel_error(r.Q.numpy(), Q, floor=1.0) <= 1e-15
    assert r.p_last == float(r.P[-1]) and r.q_last == float(r.Q[-1])
    if n >= 2:
        assert r.p_second_last == float(r.P[-2])


def test_reference_train_problem():
    # 4main.c: 18e6 samples, prints default_sum[N-2]/1e4 = 122000.004030 (naive sum, P=1).
    r = bq.cumulative_integrate(backend="cpu")
    assert abs(r.p_second_last / 1e4 - 122000.004) < 1e-6
    assert abs(r.p_last / 1e4 - 122000.00400000015) < 1e-9
    assert abs(r.q_last / 1e8 - 1.0986e8) / 1.0986e8 < 1e-4
    # position is monotone (velocity >= 0)
    assert bool((r.P[1:] >= r.P[:-1]).all())
