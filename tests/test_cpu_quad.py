"""CPU backend of integrate(): same summation tree as the kernel, checked against
closed forms and math.fsum (oracles 1-2, reference-behaviour 3)."""
import ctypes as C
import math

import pytest

import b200quad as bq
from b200quad import _native
from b200quad.utils import oracle_sin, oracle_profile, oracle_fsum

RULES = ["left", "right", "mid", "trap"]


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.mark.parametrize("rule", RULES)
@pytest.mark.parametrize("ev", ["direct", "rotate"])
@pytest.mark.parametrize("n", [1, 2, 7, 1000, 1023, 1024, 1025, 65536 + 3, 10**6, (1 << 23) + 11])
def test_sin_closed_form(rule, ev, n):
    v = bq.integrate("sin", 0.0, math.pi, n, rule, eval=ev, backend="cpu")
    ref = oracle_sin(0.0, math.pi, n, rule)
    assert abs(v - ref) <= 4e-16 * max(abs(ref), 1.0), (v, ref)


@pytest.mark.parametrize("n,rule", [((1 << 29) + 12345, "trap"), ((1 << 30) + 999, "mid")])
def test_sin_closed_form_with_several_tiles_per_part(n, rule):
    """Full 32-row tiles, three to seven per part: whole and partial flush groups of the fused fp64 loop
    (column-side rotation once per group, row constants shared by all groups: PartWorker::run_uniform
    and its mirror in cpu_backend.cpp), and the group with the ragged last tile on the general path."""
    r = bq.integrate_result("sin", 0.0, math.pi, n, rule, eval="rotate", backend="cpu")
    ref = oracle_sin(0.0, math.pi, n, rule)
    assert r.status == 0 and abs(r.value - ref) <= 1e-15 * ref


def test_row_constants_follow_the_job():
    """The row constants of the fused fp64 loop are cached per thread for the last (step, time scale)
    (host_common.cpp: fill_row_consts): calls that alternate between grids and time scales must each
    get their own."""
    import ctypes as C
    lib, _ = _native.load()

    def vel(tscale, n):
        job = bq.make_job("vel", 0.0, 900.0, n, "mid", "fp64", "rotate", params=(tscale, 0.0, 67.7777777, 0.0))
        res = _native.Result()
        assert lib.bq_cpu_integrate(C.byref(job), 1, C.byref(res)) == 0    # one thread: one cache
        exact = 67.7777777 * (900.0 - tscale * math.sin(900.0 / tscale))   # integral of vscale * (1 - cos(x / tscale))
        return res.value, exact

    n1, n2 = (1 << 24) + 3, (1 << 25) + 5          # two grids (4 and 7 rows per tile), same constants' cache
    seq = [(286.4788975, n1), (100.0, n1), (286.4788975, n1), (100.0, n2), (286.4788975, n2)]
    for tscale, n in seq:
        got, exact = vel(tscale, n)
        assert abs(got - exact) <= 1e-11 * abs(exact), (tscale, n, got, exact)   # midpoint rule error ~ h^2
    s = bq.integrate_result("sin", 0.0, math.pi, n1, "mid", eval="rotate", backend="cpu")
    assert abs(s.value - oracle_sin(0.0, math.pi, n1, "mid")) <= 2e-15


def test_reference_riemann_value():
    # riemann.cpp prints 2.00000000000222 at n=1e9 with one worker (rel. err 1.1e-12);
    # the true left sum is 2 - 1.6e-18.  We must be far closer than the reference.
    v = bq.integrate("sin", 0.0, math.pi, 10**9, "left", backend="cpu")
    assert abs(v - 2.0) < 1e-15


def test_other_interval_and_sign():
    v = bq.integrate("sin", 1.0, 2.5, 12345, "mid", backend="cpu")
    assert rel(v, oracle_sin(1.0, 2.5, 12345, "mid")) < 1e-15
    v = bq.integrate("sin", 2.5, 1.0, 12345, "mid", backend="cpu")      # reversed: negative h
    assert rel(v, oracle_sin(2.5, 1.0, 12345, "mid")) < 1e-15


@pytest.mark.parametrize("rule", RULES)
def test_profile_matches_fsum(rule):
    n = 180_000
    v = bq.integrate("profile", 0.0, 1800.0, n, rule, backend="cpu")
    assert rel(v, oracle_profile(n, rule=rule)) < 1e-15


def test_profile_reference_numbers():
    # 4main.c prints 122000.004030 (P=1) .. 122000.003997 (P=8); exactly rounded: 122000.00400000015
    left = bq.integrate("profile", 0.0, 1800.0, 18_000_000, "left", backend="cpu")
    trap = bq.integrate("profile", 0.0, 1800.0, 18_000_000, "trap", backend="cpu")
    assert abs(trap - 122000.00400000015) < 1e-9      # piecewise linear: trapezoid is exact
    assert abs(left - 122000.004) < 1e-6


@pytest.mark.parametrize("name", ["acc", "vel", "dis"])
@pytest.mark.parametrize("ev", ["direct", "rotate"])
def test_train_model_against_antiderivative(name, ev):
    f = bq.models.REGISTRY[name]
    n = 2_000_000
    v = bq.integrate(name, 0.0, 900.0, n, "mid", eval=ev, backend="cpu")
    assert rel(v, f.exact(0.0, 900.0)) < 1e-9         # midpoint rule truncation error ~ h^2
    d = bq.integrate(name, 0.0, 900.0, n, "mid", eval="direct", backend="cpu")
    assert rel(v, d) < 1e-14                          # rotate == direct to rounding


def test_vel_integral_is_distance():
    # dis(1800) ~ 122000: the analytic twin of the ex4vel.h table (riemann.cpp:113-116)
    v = bq.integrate("vel", 0.0, 1800.0, 1_000_000, "trap", backend="cpu")
    assert abs(v - 122000.0) < 1.0


@pytest.mark.parametrize("ev", ["direct", "rotate"])
def test_fp32_compensated(ev):
    n = 1 << 22
    v = bq.integrate("sin", 0.0, math.pi, n, "mid", dtype="fp32", eval=ev, backend="cpu")
    assert abs(v - 2.0) < 5e-8 * 2.0          # fp32 evaluation, compensated accumulation


@pytest.mark.parametrize("a,b,n,rule", [
    (0.0, math.pi, 1 << 26, "left"),               # the headline interval: the old bias cancelled by symmetry here
    (0.0, math.pi / 2, (1 << 26) + 5, "mid"),      # ... and not here (6.9e-8 at 2^32 before the error diffusion)
    (1.0, 1.5, 1 << 26, "right"),
    (0.2, 2.9, (1 << 25) + 12345, "trap"),         # ragged last tile
    (-3.0, 2.0, 1 << 26, "mid"),
    (0.0, math.pi / 2, 1 << 22, "mid"),            # 4 rows per tile
])
def test_fp32_rotate_on_a_fine_grid_is_accurate_to_fp64_level(a, b, n, rule):
    """fp32 FMA runs over error-diffused row values (quad_core.cuh: RowDiffusion, fused_hi,
    diffusion_owed): on a fine grid every partial sum of a run is exact in fp32 and what the
    roundings of the row values and of the column constants leave is added once per part in fp64."""
    r = bq.integrate_result("sin", a, b, n, rule, dtype="fp32", eval="rotate", backend="cpu")
    ref = oracle_sin(a, b, n, rule)
    assert abs(r.value - ref) <= 1e-12 * abs(ref)


def test_dis_linear_part_in_closed_form_matches_per_sample_evaluation():
    """ROTATE sums the linear part of dis as an arithmetic series (quad_core.cuh: grid_sum_x) and only the
    sine part sample by sample; DIRECT forms every sample.  Same tree otherwise: the two agree to an ulp."""
    for a, b, n, rule in ((0.0, 1800.0, (1 << 24) + 5, "mid"), (100.0, 1500.0, 1 << 25, "trap"),
                          (-50.0, 900.0, 12_345_678, "right"), (3.0, 4.0, 10_000_019, "left")):
        r = bq.integrate_result("dis", a, b, n, rule, eval="rotate", backend="cpu")
        d = bq.integrate_result("dis", a, b, n, rule, eval="direct", backend="cpu")
        assert r.eval == "rotate" and d.eval == "direct"
        # x and tscale * sin(x / tscale) cancel for small x (to x^3 / (6 tscale^2)): the bound is an ulp of
        # the terms that cancel, vscale * |x| per sample, not of their difference
        cancel = 67.7777777 * max(abs(a), abs(b)) * (b - a)
        assert abs(r.value - d.value) <= 4e-16 * (abs(d.value) + cancel), (r.value, d.value)


@pytest.mark.parametrize("name,a,b", [("vel", 0.0, 1800.0), ("acc", 5.0, 900.0), ("vel", 100.0, 700.0),
                                      ("dis", 0.0, 1800.0)])
def test_fp32_rotate_train_models_match_fp64(name, a, b):
    """-ascale * sin and vscale - vscale * cos run the fused fp32 loop on the integrand divided by its
    amplitude (quad_core.cuh: fused_norm; the cos type with its two streams swapped), so the unit
    column constant and the exact partial sums carry over."""
    for n, rule in ((1 << 24, "mid"), ((1 << 25) + 777, "left")):
        r = bq.integrate_result(name, a, b, n, rule, dtype="fp32", eval="rotate", backend="cpu")
        d = bq.integrate_result(name, a, b, n, rule, dtype="fp64", eval="rotate", backend="cpu")
        assert abs(r.value - d.value) <= 1e-11 * abs(d.value)      # the ragged last row is outside the scheme


def test_fp32_rotate_coarse_grid_keeps_the_old_bound():
    # 2048 columns span more than 2^-10 rad: the two products meet in an fp32 add at every run
    for n in (1 << 16, 1 << 20):
        r = bq.integrate_result("sin", 0.0, math.pi, n, "left", dtype="fp32", eval="rotate", backend="cpu")
        assert abs(r.value - oracle_sin(0.0, math.pi, n, "left")) <= 5e-9 * 2.0


def _segments(job, rank, world):
    lib, _ = _native.load()
    res = _native.Result()
    assert lib.bq_cpu_integrate_segments(C.byref(job), rank, world, 2, C.byref(res)) == 0
    return list(res.seg_hi), list(res.seg_lo)


@pytest.mark.parametrize("n", [1000, 10**6 + 7])
def test_world_size_does_not_change_bits(n):
    job = bq.make_job("sin", 0.0, math.pi, n, "left")
    whole = bq.integrate_result("sin", 0.0, math.pi, n, "left", backend="cpu")
    lib, _ = _native.load()
    for world in (1, 2, 4, 8):
        hi = [0.0] * 8
        lo = [0.0] * 8
        for r in range(world):
            h, l = _segments(job, r, world)
            per = 8 // world
            for s in range(r * per, (r + 1) * per):
                hi[s], lo[s] = h[s], l[s]
        out = _native.Result()
        assert lib.bq_fold_segments(C.byref(job), (C.c_double * 8)(*hi), (C.c_double * 8)(*lo),
                                    C.byref(out)) == 0
        assert out.value == whole.value and out.value_lo == whole.value_lo


def test_threads_do_not_change_bits():
    a = bq.integrate("sin", 0.0, math.pi, 3_000_001, backend="cpu", threads=1)
    b = bq.integrate("sin", 0.0, math.pi, 3_000_001, backend="cpu", threads=7)
    assert a == b


def test_bad_arguments():
    with pytest.raises(_native.NativeError):
        bq.integrate("sin", 0.0, 1.0, 0, backend="cpu")
    with pytest.raises(_native.NativeError):
        bq.integrate("profile", 0.0, 2000.0, 100, backend="cpu")      # outside the table
    with pytest.raises(ValueError):
        bq.integrate("nope", 0.0, 1.0, 10, backend="cpu")
    with pytest.raises(ValueError):
        bq.integrate("sin", 0.0, 1.0, 10, "simpson", backend="cpu")


def test_python_callable_slow_path():
    v = bq.integrate(lambda x: x.sin(), 0.0, math.pi, 100_000, "trap")
    assert rel(v, oracle_sin(0.0, math.pi, 100_000, "trap")) < 1e-13


def test_fsum_oracle_agrees_with_closed_form():
    assert rel(oracle_fsum(math.sin, 0.0, math.pi, 5000, "mid"), oracle_sin(0.0, math.pi, 5000, "mid")) < 1e-15


def test_default_profile_table():
    t = bq.models.default_profile()
    assert len(t) == 1801 and t[0] == 0.0 and t[1] == 0.002904762
    assert abs(math.fsum(t) - 122000.004) < 1e-9
    assert abs(max(t) - 87.14286) < 1e-4


def test_bounded_sincos_accuracy():
    """The sin/cos of the DIRECT fast path (bq_sin_bounded / bq_cos_bounded = the host build of the
    code the kernel runs) against 80-bit sinl/cosl: absolute error below one ulp of 1 everywhere on
    |x| <= 1e6 (what a sum of samples needs), and < 1.5 ulp relative away from the zeros."""
    import numpy as np
    from b200quad import _native
    lib, _ = _native.load()
    rng = np.random.default_rng(7)
    xs = np.concatenate([rng.uniform(-1e6, 1e6, 60000), rng.uniform(-10, 10, 30000),
                         np.arange(-2000, 2000) * (math.pi / 2) + rng.uniform(-1e-6, 1e-6, 4000)])
    ld = np.longdouble
    worst_abs, worst_rel = 0.0, 0.0
    for x in xs:
        for fn, ref in ((lib.bq_sin_bounded, np.sin(ld(x))), (lib.bq_cos_bounded, np.cos(ld(x)))):
            err = float(abs(ld(fn(float(x))) - ref))
            worst_abs = max(worst_abs, err)
            if abs(float(ref)) >= 1e-3:
                worst_rel = max(worst_rel, err / float(np.spacing(np.float64(abs(float(ref))))))
    assert worst_abs <= 1.25 * 2.0 ** -53, worst_abs      # measured 1.11 * 2^-53
    assert worst_rel <= 1.5, worst_rel                   # ulps; measured 1.39
