"""Custom integrands (b200quad.models.custom): a C++ expression compiled into the fused kernel.
CPU tier here (the unit is built with g++ and feeds the CPU backend); the GPU tier checks the
nvcc-built kernel against it."""
import math

import pytest

import b200quad as bq
from b200quad.utils import oracle_fsum


def test_polynomial_cpu_backend_matches_exact_sum():
    f = bq.models.custom("x*x*x - 2.0*x + p0", params=(0.25,))
    n = 20_000
    r = bq.integrate_result(f, 0.0, 2.0, n, "mid", backend="cpu")
    ref = oracle_fsum(lambda x: x * x * x - 2.0 * x + 0.25, 0.0, 2.0, n, "mid")
    assert r.status == 0
    assert abs(r.value - ref) <= 2e-15 * max(1.0, abs(ref))
    assert abs(r.value - 0.5) < 1e-7          # the integral itself is 4 - 4 + 0.5


def test_two_expressions_can_alternate():
    f = bq.models.custom("exp(-x*x)")
    g = bq.models.custom("1.0/(1.0+x*x)")
    for _ in range(2):
        a = bq.integrate(f, 0.0, 3.0, 50_000, "mid", backend="cpu")
        b = bq.integrate(g, 0.0, 1.0, 50_000, "trap", backend="cpu")
        assert abs(a - 0.5 * math.sqrt(math.pi) * math.erf(3.0)) < 1e-9
        assert abs(b - math.pi / 4) < 1e-9


def test_bad_expression_reports_compiler_output():
    from b200quad.models.jit import CustomBuildError
    with pytest.raises(CustomBuildError):
        bq.integrate(bq.models.custom("this is not C++"), 0.0, 1.0, 10, backend="cpu")


def test_cumulative_rejects_custom():
    with pytest.raises(Exception):
        bq.cumulative_integrate(bq.models.custom("x"), 0.0, 1e-3, 1000, backend="cpu")


@pytest.mark.gpu
def test_gpu_kernel_matches_cpu_tree_bitwise_for_a_polynomial():
    f = bq.models.custom("x*x*x - 2.0*x + p0", params=(0.25,))
    for n in (1000, (1 << 20) + 17, 1 << 26):
        for rule in ("left", "mid", "trap"):
            g = bq.integrate_result(f, 0.0, 2.0, n, rule, backend="cuda")
            c = bq.integrate_result(f, 0.0, 2.0, n, rule, backend="cpu")
            assert g.status == 0 and g.value == c.value and g.segments == c.segments, (n, rule)


@pytest.mark.gpu
def test_gpu_transcendental_and_fp32():
    f = bq.models.custom("exp(-x*x)")
    exact = 0.5 * math.sqrt(math.pi) * math.erf(3.0)
    assert abs(bq.integrate(f, 0.0, 3.0, 1 << 28, "mid") - exact) < 1e-12
    assert abs(bq.integrate(f, 0.0, 3.0, 1 << 24, "mid", dtype="fp32") - exact) < 2e-5   # x itself is rounded to fp32


@pytest.mark.gpu
def test_gpu_custom_on_several_gpus_same_bits():
    from conftest import gpu_count
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    f = bq.models.custom("sqrt(x) * p0", params=(3.0,))
    one = bq.integrate_result(f, 0.0, 4.0, (1 << 24) + 5, "mid", backend="cuda")
    two = bq.integrate_result(f, 0.0, 4.0, (1 << 24) + 5, "mid", world=2, backend="cuda")
    assert one.value == two.value and abs(one.value - 16.0) < 1e-9
