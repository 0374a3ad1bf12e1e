"""Protocol stress and large-n accuracy of the fused kernels (GPU tier).

The cross-GPU exchange carries no fence and no counter: a partial result is a few 8-byte words that
hold 32 bits of payload and the 32-bit call number (device_state.h), double-buffered by the
call number's parity.  What could go wrong is a word of an OLD call being taken for the current one, so
these tests run 10^4 calls back to back with alternating problem sizes — a stale word would change
the value — in both process layouts (one process driving all GPUs; one process per GPU as under
torchrun).  The reference's counterpart is a blocking MPI_Send/Recv per call (riemann.cpp:76,83).
"""
import math
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

import b200quad as bq
from b200quad import _native, parallel
from b200quad.utils import oracle_sin
from conftest import gpu_count

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CALLS = 10_000


def _world_sizes():
    return [w for w in (2, 4, 8) if w <= gpu_count()]


def test_ten_thousand_integrals_back_to_back_local_world():
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = max(_world_sizes())
    w = parallel.local_world(world)
    jobs = [bq.make_job("sin", 0.0, math.pi, n) for n in (4096, (1 << 16) + 7, 1 << 12, (1 << 20) + 3)]
    want = [bq.integrate_result("sin", 0.0, math.pi, int(j.n), backend="cuda").value for j in jobs]
    for i in range(CALLS):
        k = i % len(jobs)
        res = w.integrate(jobs[k])
        assert all(r.status == 0 and r.value == want[k] for r in res), (i, [r.value for r in res], want[k])


def test_two_thousand_scans_back_to_back_local_world():
    """The scan's descriptors are generation-stamped and never cleared: alternate sizes so that every
    call finds the previous call's descriptors of a different grid in its scratch."""
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = max(_world_sizes())
    sizes = (70_001, 300_000, 4_097, 1_000_003)
    want = {}
    for n in sizes:
        c = bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, backend="cuda")
        want[n] = (c.p_last, c.q_last, c.p_second_last)
    for i in range(2000):
        n = sizes[i % len(sizes)]
        r = bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, world=world, replicate=bool(i & 4), backend="cuda")
        for o in r.per_rank:
            assert o.status == 0 and (o.p_last, o.q_last, o.p_second_last) == want[n], (i, n)


def test_scans_back_to_back_one_gpu_alternating_sizes():
    sizes = (2_049, 5_000_000, 65_537, 18_000_000)
    ref = {n: bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, backend="cpu") for n in sizes[:3]}
    for i in range(300):
        n = sizes[i % len(sizes)]
        g = bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, backend="cuda")
        assert g.status == 0
        if n in ref and i < 12:
            assert torch.equal(g.P.cpu(), ref[n].P) and torch.equal(g.Q.cpu(), ref[n].Q)
        if n in ref:
            assert (g.p_last, g.q_last) == (ref[n].p_last, ref[n].q_last)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import b200quad as bq2
    ns = (4096, (1 << 16) + 7, 1 << 12, (1 << 20) + 3)
    got = {n: set() for n in ns}
    for i in range(CALLS):
        n = ns[i % len(ns)]
        got[n].add(bq2.integrate("sin", 0.0, math.pi, n))
    scans = set()
    for i in range(500):
        n = (70_001, 300_000)[i & 1]
        c = bq2.cumulative_integrate("profile", 0.0, 1800.0 / n, n, replicate=False)
        scans.add((n, c.p_last, c.q_last, c.status))
    q.put((rank, {n: sorted(v) for n, v in got.items()}, sorted(scans)))
    dist.barrier()
    dist.destroy_process_group()


def test_ten_thousand_integrals_back_to_back_torchrun_layout():
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = max(_world_sizes())
    ns = (4096, (1 << 16) + 7, 1 << 12, (1 << 20) + 3)
    want = {n: bq.integrate_result("sin", 0.0, math.pi, n, backend="cuda").value for n in ns}
    want_scan = []
    for n in (70_001, 300_000):
        c = bq.cumulative_integrate("profile", 0.0, 1800.0 / n, n, backend="cuda")
        want_scan.append((n, c.p_last, c.q_last, 0))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, got, scans in out:
        for n in ns:
            assert got[n] == [want[n]], (rank, n, got[n], want[n])      # one value ever seen per size: the right one
        assert scans == sorted(want_scan), (rank, scans)


# ---- accuracy at the benchmark's size -----------------------------------------------------------

@pytest.mark.parametrize("ev", ["rotate", "direct"])
def test_fp32_accuracy_at_2p32(ev):
    """Compensated fp32 at the headline size: fp32 evaluation, every combine in fp64 pairs."""
    n = 1 << 32
    r = bq.integrate_result("sin", 0.0, math.pi, n, "left", dtype="fp32", eval=ev, backend="cuda")
    ref = oracle_sin(0.0, math.pi, n, "left")
    assert r.status == 0
    # ROTATE: fp32 FMA runs of 32 rows over error-diffused row values, promoted to fp64 (measured 1e-16);
    # DIRECT: x rounded to fp32 (ulp(pi) = 2.4e-7)
    bound = 1e-12 if ev == "rotate" else 5e-8
    assert abs(r.value - ref) / ref <= bound, (ev, r.value)


def test_fp32_rotate_accuracy_at_2p32_off_the_symmetric_interval():
    """[0, pi/2]: the bias of plain fp32 runs (6.9e-8 here with 16-row runs) has nothing to cancel against."""
    n = 1 << 32
    r = bq.integrate_result("sin", 0.0, math.pi / 2, n, "mid", dtype="fp32", eval="rotate", backend="cuda")
    ref = oracle_sin(0.0, math.pi / 2, n, "mid")
    assert r.status == 0 and abs(r.value - ref) / ref <= 1e-12, r.value


@pytest.mark.parametrize("name,a,b", [("vel", 0.0, 1800.0), ("acc", 5.0, 900.0)])
def test_fp32_rotate_train_models_match_fp64_at_2p30(name, a, b):
    n = 1 << 30
    r = bq.integrate_result(name, a, b, n, "mid", dtype="fp32", eval="rotate", backend="cuda")
    d = bq.integrate_result(name, a, b, n, "mid", dtype="fp64", eval="rotate", backend="cuda")
    assert r.status == 0 and abs(r.value - d.value) <= 1e-12 * abs(d.value), (r.value, d.value)


@pytest.mark.parametrize("ev", ["rotate", "direct"])
def test_fp64_accuracy_at_2p32(ev):
    n = 1 << 32
    r = bq.integrate_result("sin", 0.0, math.pi, n, "mid", eval=ev, backend="cuda")
    ref = oracle_sin(0.0, math.pi, n, "mid")
    assert r.status == 0 and abs(r.value - ref) / ref <= 1e-15, (ev, r.value, ref)


def test_direct_fast_path_matches_cpu_tree_bitwise():
    """The bounded sin/cos and the one-fma table runs are the same code on host and device."""
    for name, a, b in (("sin", 0.0, math.pi), ("vel", 0.0, 1800.0), ("dis", 0.0, 1800.0), ("acc", 5.0, 900.0),
                       ("profile", 0.0, 1800.0)):
        for n in (1000, 3_000_017):
            for rule in ("left", "trap"):
                g = bq.integrate_result(name, a, b, n, rule, eval="direct", backend="cuda")
                c = bq.integrate_result(name, a, b, n, rule, eval="direct", backend="cpu")
                assert g.status == 0 and (g.value, g.value_lo) == (c.value, c.value_lo), (name, n, rule)
                assert g.segments == c.segments


def test_direct_falls_back_to_libm_outside_the_bounded_range():
    """|x| up to 3e6 > 1e6: the plan must not select the bounded sin there.  At this offset the step
    h = fl((b - a) / n) and the rounding of each x matter at the 1e-7 level, so the oracle is the CPU
    tree (same x, libm sin) and, loosely, a plain fp64 PyTorch sum of the same samples."""
    a, b, n = 2.0e6, 3.0e6, 2_000_003
    g = bq.integrate_result("sin", a, b, n, "mid", eval="direct", backend="cuda")
    c = bq.integrate_result("sin", a, b, n, "mid", eval="direct", backend="cpu")
    assert g.status == 0 and abs(g.value - c.value) <= 1e-11
    h = (b - a) / n
    x = a + (torch.arange(n, dtype=torch.float64, device="cuda") + 0.5) * h
    assert abs(g.value - float(torch.sin(x).sum()) * h) <= 1e-5


# ---- stream ordering of one context -----------------------------------------------------------------

def test_calls_on_two_streams_do_not_overlap_on_one_context():
    """A context owns one set of scratch; a call that arrives on another stream than the previous
    one is ordered after it (ctx.h: order_after_previous)."""
    from b200quad.ops import torch_ops as ops
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    want_a = float(ops.integrate("sin", 0.0, math.pi, (1 << 24) + 1))
    want_b = float(ops.integrate("vel", 0.0, 1800.0, (1 << 22) + 5))
    torch.cuda.synchronize()
    outs = []
    for i in range(200):
        with torch.cuda.stream(s1 if i & 1 else s2):
            outs.append((i & 1, ops.integrate("sin", 0.0, math.pi, (1 << 24) + 1) if i & 1
                         else ops.integrate("vel", 0.0, 1800.0, (1 << 22) + 5)))
    torch.cuda.synchronize()
    for odd, t in outs:
        assert float(t) == (want_a if odd else want_b)


def test_scan_peer_timeout_is_reported_not_hung():
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    w = parallel.local_world(2)
    for c in w.contexts:
        c.set_timeout_ms(200)
    w.contexts[0].inject_fault(1)          # rank 0 never publishes its segment totals
    try:
        with pytest.raises(_native.NativeError):
            bq.cumulative_integrate("profile", 0.0, 1e-3, 1_800_000, world=2, replicate=False, backend="cuda")
    finally:
        for c in w.contexts:
            c.reset_comm()
            c.set_timeout_ms(2000)
    r = bq.cumulative_integrate("profile", 0.0, 1e-3, 1_800_000, world=2, replicate=False, backend="cuda")
    assert all(o.status == 0 for o in r.per_rank)


@pytest.mark.parametrize("n", [1, 1000, 2048, 2049, 4097])
def test_scan_with_ranks_that_own_no_tile(n):
    """Tiny grids on several GPUs: some ranks' segments hold no tile; they still take part in the call
    (empty slice, their segment totals are published as zero) and every rank reports the right scalars."""
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    one = bq.cumulative_integrate("profile", 0.0, 1800.0 / max(n, 2), n, backend="cuda")
    for world in _world_sizes():
        for replicate in (False, True):
            r = bq.cumulative_integrate("profile", 0.0, 1800.0 / max(n, 2), n, world=world, replicate=replicate,
                                        backend="cuda")
            for rank, o in enumerate(r.per_rank):
                assert o.status == 0 and (o.p_last, o.q_last, o.p_second_last) == (one.p_last, one.q_last, one.p_second_last)
                lo, hi = r.slices[rank]
                assert torch.equal(r.all_P[rank].cpu(), one.P[lo:hi].cpu()) and torch.equal(r.all_Q[rank].cpu(), one.Q[lo:hi].cpu())
