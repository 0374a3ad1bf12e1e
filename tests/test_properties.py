"""Property tests on the CPU backend (it walks the same summation tree as the sm_100a kernel, and
the GPU tier asserts GPU == CPU bit for bit, so what holds here holds for the kernel's arithmetic)."""
import ctypes as C
import math

from hypothesis import given, settings, strategies as st

import b200quad as bq
from b200quad import _native
from b200quad.utils import oracle_sin

RULES = ["left", "right", "mid", "trap"]


@settings(max_examples=60, deadline=None, derandomize=True)
@given(n=st.integers(min_value=1, max_value=300_000), rule=st.sampled_from(RULES),
       ev=st.sampled_from(["direct", "rotate"]),
       a=st.floats(min_value=-3.0, max_value=3.0), width=st.floats(min_value=1e-3, max_value=6.0))
def test_sin_any_interval_matches_closed_form(n, rule, ev, a, width):
    b = a + width
    r = bq.integrate_result("sin", a, b, n, rule, eval=ev, backend="cpu")
    ref = oracle_sin(a, b, n, rule)
    # the Riemann sum may cancel (e.g. around a zero of the integral): bound against the sum of |terms|
    scale = max(abs(ref), width * 0.05, 1e-300)
    assert r.status == 0
    assert abs(r.value - ref) <= 6e-15 * scale + 1e-18, (r.value, ref)


@settings(max_examples=40, deadline=None, derandomize=True)
@given(n=st.integers(min_value=1, max_value=6_000_000), rule=st.sampled_from(RULES),
       name=st.sampled_from(["sin", "vel", "acc"]), a=st.floats(min_value=-3.0, max_value=3.0),
       width=st.floats(min_value=1e-3, max_value=6.0))
def test_fp32_rotate_tracks_fp64_on_any_interval(n, rule, name, a, width):
    """The fused fp32 loop (error-diffused row values, amplitude-normalised integrand, swapped streams
    for the cos type, ragged last row, closing term per part) against the fp64 loop over the same tree.
    Coarse grids round every product (a few 1e-9 of the sum of |terms|); fine ones are exact."""
    scale_t = 1.0 if name == "sin" else 300.0          # the train models take x in seconds
    a, b = a * scale_t, (a + width) * scale_t
    r = bq.integrate_result(name, a, b, n, rule, dtype="fp32", eval="rotate", backend="cpu")
    d = bq.integrate_result(name, a, b, n, rule, dtype="fp64", eval="rotate", backend="cpu")
    peak = {"sin": 1.0, "vel": 2.0 * 122000.004 / 1800.0, "acc": 1.0}[name]    # bound on |f| (acc: ~0.71)
    assert r.status == 0 and d.status == 0
    # a few rows of samples: all columns of a row share the row value's fp32 rounding (6e-8); many rows: it averages out
    assert abs(r.value - d.value) <= (3e-7 if n < 100_000 else 1e-8) * peak * (b - a), (r.value, d.value)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(n=st.integers(min_value=1, max_value=200_000), rule=st.sampled_from(RULES),
       name=st.sampled_from(["sin", "vel", "dis", "profile"]), world=st.sampled_from([2, 4, 8]))
def test_any_world_size_reassembles_the_same_bits(n, rule, name, world):
    """Segments computed rank by rank (as the ranks of a world would) fold to the single-rank bits."""
    lib, _ = _native.load()
    a, b = (0.0, math.pi) if name == "sin" else (0.0, 1800.0)
    job = bq.make_job(name, a, b, n, rule)
    whole = _native.Result()
    assert lib.bq_cpu_integrate(C.byref(job), 2, C.byref(whole)) == 0
    hi = (C.c_double * 8)()
    lo = (C.c_double * 8)()
    for rank in range(world):
        part = _native.Result()
        assert lib.bq_cpu_integrate_segments(C.byref(job), rank, world, 1, C.byref(part)) == 0
        for s in range(rank * (8 // world), (rank + 1) * (8 // world)):
            hi[s], lo[s] = part.seg_hi[s], part.seg_lo[s]
    folded = _native.Result()
    assert lib.bq_fold_segments(C.byref(job), hi, lo, C.byref(folded)) == 0
    assert folded.value == whole.value and folded.value_lo == whole.value_lo


@settings(max_examples=20, deadline=None, derandomize=True)
@given(n=st.integers(min_value=1, max_value=60_000), x0=st.floats(min_value=0.0, max_value=100.0),
       dx=st.floats(min_value=1e-4, max_value=2e-2))
def test_cumulative_tables_are_running_sums(n, x0, dx):
    from oracle_prefix import exact_prefix, max_rel_error
    c = bq.cumulative_integrate("profile", x0, dx, n, want_samples=True, backend="cpu")
    # exact running sums of the stored samples (a naive fp64 cumsum is the weaker party here:
    # it drifts by ~1e-13 where the library stays within a few ulp)
    P, Q = exact_prefix(c.samples.numpy())
    assert max_rel_error(c.P.numpy(), P) <= 1e-15
    assert max_rel_error(c.Q.numpy(), Q) <= 1e-15
    assert c.p_last == float(c.P[-1]) and c.q_last == float(c.Q[-1])
