"""The fused sm_100a definite-integral kernel against a plain fp32/fp64 PyTorch
reference of the same op, closed forms, and the CPU backend (same tree)."""
import math

import pytest
import torch

import b200quad as bq
from b200quad import _native, parallel
from b200quad.utils import oracle_sin, oracle_profile
from conftest import gpu_count

pytestmark = pytest.mark.gpu

RULES = ["left", "right", "mid", "trap"]


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


def torch_reference(name, a, b, n, rule, dtype=torch.float64):
    """Plain PyTorch evaluation of the same Riemann sum (chunked), fp64 accumulation."""
    f = bq.models.REGISTRY[name]
    h = (b - a) / n
    off = {"left": 0.0, "right": 1.0, "mid": 0.5, "trap": 0.0}[rule]
    total = 0.0
    for s in range(0, n, 1 << 24):
        m = min(1 << 24, n - s)
        i = torch.arange(s, s + m, device="cuda", dtype=torch.float64)
        x = (a + (i + off) * h).to(dtype)
        total += float(f(x).to(torch.float64).sum())
    val = total * h
    if rule == "trap":
        e = f(torch.tensor([a, b], device="cuda", dtype=torch.float64))
        val += 0.5 * h * float(e[1] - e[0])
    return val


def test_loaded_cuda_library():
    lib, cuda = _native.load()
    assert cuda and _native.cuda_device_count() >= 1


@pytest.mark.parametrize("rule", RULES)
@pytest.mark.parametrize("ev", ["direct", "rotate"])
@pytest.mark.parametrize("n", [1, 5, 1000, 1024 * 37 + 5, 1 << 20, (1 << 24) + 12345])
def test_sin_vs_closed_form(rule, ev, n):
    r = bq.integrate_result("sin", 0.0, math.pi, n, rule, eval=ev, backend="cuda")
    ref = oracle_sin(0.0, math.pi, n, rule)
    assert r.status == 0
    assert abs(r.value - ref) <= 1e-15 * max(abs(ref), 1.0), (r.value, ref)


@pytest.mark.parametrize("rule", RULES)
def test_sin_vs_torch_reference(rule):
    n = (1 << 25) + 3
    r = bq.integrate("sin", 0.3, 2.9, n, rule, backend="cuda")
    assert rel(r, torch_reference("sin", 0.3, 2.9, n, rule)) < 1e-12


def test_large_n_both_modes_agree():
    n = 1 << 30
    rot = bq.integrate_result("sin", 0.0, math.pi, n, "left", eval="rotate", backend="cuda")
    dire = bq.integrate_result("sin", 0.0, math.pi, n, "left", eval="direct", backend="cuda")
    ref = oracle_sin(0.0, math.pi, n, "left")
    assert rel(rot.value, ref) < 1e-15 and rel(dire.value, ref) < 1e-15
    assert rot.eval == "rotate" and dire.eval == "direct"


def test_64bit_index_space():
    n = (1 << 33) + 7          # more points than a 32-bit index (the reference caps at 2^31-1)
    r = bq.integrate_result("sin", 0.0, math.pi, n, "mid", backend="cuda")
    assert rel(r.value, oracle_sin(0.0, math.pi, n, "mid")) < 1e-15


@pytest.mark.parametrize("name", ["acc", "vel", "dis"])
@pytest.mark.parametrize("ev", ["direct", "rotate"])
def test_train_model(name, ev):
    n = (1 << 22) + 99
    r = bq.integrate(name, 0.0, 900.0, n, "mid", eval=ev, backend="cuda")
    assert rel(r, torch_reference(name, 0.0, 900.0, n, "mid")) < 1e-12
    assert rel(r, bq.models.REGISTRY[name].exact(0.0, 900.0)) < 1e-9


@pytest.mark.parametrize("rule", RULES)
def test_profile(rule):
    n = 1_800_000
    r = bq.integrate("profile", 0.0, 1800.0, n, rule, backend="cuda")
    assert rel(r, oracle_profile(n, rule=rule)) < 1e-15
    assert rel(r, torch_reference("profile", 0.0, 1800.0, n, rule)) < 1e-12


def test_profile_reference_size():
    r = bq.integrate("profile", 0.0, 1800.0, 18_000_000, "trap", backend="cuda")
    assert abs(r - 122000.00400000015) < 1e-9


def test_custom_profile_table():
    ctx = parallel.context()
    tab = [float(i * i) for i in range(101)]           # x^2 sampled at integers
    ctx.set_profile(tab)
    try:
        r = bq.integrate("profile", 0.0, 100.0, 100_000, "trap", backend="cuda")
        exact = sum((tab[i] + tab[i + 1]) / 2 for i in range(100))
        assert rel(r, exact) < 1e-13
    finally:
        ctx.set_profile(bq.models.default_profile())


@pytest.mark.parametrize("ev", ["direct", "rotate"])
def test_fp32_compensated_vs_fp32_torch(ev):
    n = 1 << 24
    r = bq.integrate("sin", 0.0, math.pi, n, "mid", dtype="fp32", eval=ev, backend="cuda")
    naive = torch_reference("sin", 0.0, math.pi, n, "mid", dtype=torch.float32)
    assert abs(r - 2.0) < 1e-7            # compensated fp32
    assert abs(r - 2.0) <= abs(naive - 2.0) + 1e-7


def test_matches_cpu_backend_tree():
    # Same tree, libm vs libdevice sin/sincos differ by ulps per point only.
    for ev in ("direct", "rotate"):
        n = 5_000_011
        g = bq.integrate("sin", 0.0, math.pi, n, "left", eval=ev, backend="cuda")
        c = bq.integrate("sin", 0.0, math.pi, n, "left", eval=ev, backend="cpu")
        assert rel(g, c) < 1e-15
    g = bq.integrate("profile", 0.0, 1800.0, 1_234_567, "left", backend="cuda")
    c = bq.integrate("profile", 0.0, 1800.0, 1_234_567, "left", backend="cpu")
    assert g == c                          # no libm in this integrand: bit-identical


def test_launch_shape_does_not_change_bits():
    ctx = parallel.context()
    n = (1 << 26) + 17
    base = bq.integrate_result("sin", 0.0, math.pi, n, backend="cuda")
    try:
        for ctas, cluster in [(1, 1), (1, 2), (2, 1), (3, 2), (4, 1)]:
            ctx.configure(ctas, cluster, 0)
            r = bq.integrate_result("sin", 0.0, math.pi, n, backend="cuda")
            assert (r.value, r.value_lo) == (base.value, base.value_lo), (ctas, cluster)
    finally:
        ctx.configure(0, 0, 0)


def test_back_to_back_async_calls():
    ctx = parallel.context()
    job = bq.make_job("sin", 0.0, math.pi, 1 << 22, "mid")
    for _ in range(300):                   # many launches in flight on one stream
        ctx.integrate_async(job)
    r = ctx.fetch()
    assert r.status == 0 and rel(r.value, oracle_sin(0.0, math.pi, 1 << 22, "mid")) < 1e-15


def test_torch_op_on_current_stream():
    from b200quad.ops import torch_ops
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        t = torch_ops.integrate("sin", 0.0, math.pi, 1 << 24, "mid")
        u = t * 2.0
    s.synchronize()
    assert t.is_cuda and abs(float(u) - 4.0) < 1e-12


def test_kernel_timestamps():
    r = bq.integrate_result("sin", 0.0, math.pi, 1 << 28, backend="cuda")
    assert r.kernel_ns > 0 and r.device_ms > 0
    assert r.kernel_ns * 1e-6 <= r.device_ms * 1.5 + 0.05


@pytest.mark.parametrize("world", [2, 4, 8])
def test_local_world_bitwise_equal_to_single_gpu(world):
    if gpu_count() < world:
        pytest.skip("needs %d GPUs" % world)
    n = (1 << 28) + 5
    one = bq.integrate_result("sin", 0.0, math.pi, n, backend="cuda")
    w = parallel.local_world(world)
    job = bq.make_job("sin", 0.0, math.pi, n)
    for _ in range(20):
        res = w.integrate(job)
        for r in res:
            assert r.status == 0
            assert (r.value, r.value_lo) == (one.value, one.value_lo)


def test_peer_timeout_is_reported_not_hung():
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    w = parallel.local_world(2)
    job = bq.make_job("sin", 0.0, math.pi, 1 << 20)
    for c in w.contexts:
        c.set_timeout_ms(200)
    w.contexts[1].inject_fault(1)          # rank 1 never publishes
    with pytest.raises(_native.NativeError):
        w.integrate(job)
    for c in w.contexts:
        c.reset_comm()
        c.set_timeout_ms(2000)
    res = w.integrate(job)
    assert all(r.status == 0 for r in res)


def test_value_fast_path_matches_and_survives_back_to_back_calls():
    """bq.integrate() takes the event-free path (completion word in the pinned result block):
    same bits as the event-timed call, and the word protocol holds over many calls in a row."""
    ctx = parallel.context(0)
    for n in (1, 1000, (1 << 20) + 3, 1 << 26):
        for rule in RULES:
            job = bq.make_job("sin", 0.0, math.pi, n, rule)
            assert ctx.integrate_value(job) == ctx.integrate(job).value
    ref = bq.integrate_result("sin", 0.1, 2.0, 12345, "mid", backend="cuda").value
    for _ in range(2000):
        assert bq.integrate("sin", 0.1, 2.0, 12345, "mid") == ref
    big = bq.integrate_result("vel", 0.0, 1800.0, 1 << 24, backend="cuda").value
    for _ in range(20):     # alternate sizes so a stale block would be noticed
        assert bq.integrate("vel", 0.0, 1800.0, 1 << 24) == big
        assert bq.integrate("sin", 0.1, 2.0, 12345, "mid") == ref


def test_result_block_carries_call_number():
    ctx = parallel.context(0)
    job = bq.make_job("sin", 0.0, math.pi, 4096)
    a = ctx.integrate(job).seq
    b = ctx.integrate(job).seq
    assert b == a + 1
