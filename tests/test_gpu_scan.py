"""The fused cumulative-integration kernel vs a plain PyTorch cumsum reference and the CPU tree."""
import math

import pytest
import torch

import b200quad as bq
from b200quad import parallel
from conftest import gpu_count

pytestmark = pytest.mark.gpu


from oracle_prefix import exact_prefix, max_rel_error


def _oracle(name, x0, dx, n):
    """Plain PyTorch fp64 reference of the sampled integrand."""
    x = x0 + torch.arange(n, dtype=torch.float64, device="cuda") * dx
    return bq.models.REGISTRY[name](x)


@pytest.mark.parametrize("n", [1, 2, 9, 2047, 2048, 2049, 300_001, 18_000_000])
@pytest.mark.parametrize("name", ["profile", "vel"])
def test_tables_match_exact_running_sums(name, n):
    dx = 1800.0 / max(n, 2)
    r = bq.cumulative_integrate(name, 0.0, dx, n, want_samples=True, backend="cuda")
    v = _oracle(name, 0.0, dx, n)
    assert r.status == 0
    assert float((r.samples - v).abs().max()) <= 1e-12 * max(float(v.abs().max()), 1.0)
    # P and Q against the EXACT running sums of the samples the kernel evaluated (a naive fp64
    # cumsum would be off by 1e-10 at 18e6 terms; the kernel is not)
    P, Q = exact_prefix(r.samples.cpu().numpy())
    assert max_rel_error(r.P.cpu().numpy(), P, floor=1.0) <= 1e-15
    assert max_rel_error(r.Q.cpu().numpy(), Q, floor=1.0) <= 1e-15
    assert r.p_last == float(r.P[-1]) and r.q_last == float(r.Q[-1])
    if n >= 2:
        assert r.p_second_last == float(r.P[-2])


def test_reference_train_problem_bitwise_vs_cpu_tree():
    g = bq.cumulative_integrate(backend="cuda")            # 18e6 samples of the profile
    c = bq.cumulative_integrate(backend="cpu")
    assert abs(g.p_second_last / 1e4 - 122000.004) < 1e-6
    assert abs(g.p_last / 1e4 - 122000.00400000015) < 1e-9
    assert torch.equal(g.P.cpu(), c.P) and torch.equal(g.Q.cpu(), c.Q)   # same tree, no libm


def test_repeated_calls_reuse_state():
    a = bq.cumulative_integrate("profile", 0.0, 1e-3, 1_800_000, backend="cuda")
    for _ in range(5):
        b = bq.cumulative_integrate("profile", 0.0, 1e-3, 1_800_000, backend="cuda")
        assert torch.equal(a.P, b.P) and torch.equal(a.Q, b.Q)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("replicate", [True, False])
def test_local_world_scan(world, replicate):
    if gpu_count() < world:
        pytest.skip("needs %d GPUs" % world)
    n = 18_000_000
    one = bq.cumulative_integrate(n=n, backend="cuda")
    r = bq.cumulative_integrate(n=n, world=world, replicate=replicate, backend="cuda")
    for rank in range(world):
        lo, hi = r.slices[rank]
        assert r.per_rank[rank].status == 0
        assert torch.equal(r.all_P[rank].cpu(), one.P[lo:hi].cpu())
        assert torch.equal(r.all_Q[rank].cpu(), one.Q[lo:hi].cpu())
        assert r.per_rank[rank].p_last == one.p_last
        assert r.per_rank[rank].p_second_last == one.p_second_last
