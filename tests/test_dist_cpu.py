"""Multi-process plumbing on CPU: gloo, world_size 2 and 4, one integrate() per rank.
Every rank must end with the same bits as a single-process run."""
import math
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCAN_N = 300_001


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, expected, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["B200QUAD_FORCE_CPU"] = "1"
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200quad as bq
    r = bq.integrate_result("sin", 0.0, math.pi, n, "left", backend="cpu", threads=2)
    p = bq.integrate("profile", 0.0, 1800.0, 180_000, "trap", backend="cpu", threads=2)
    # the distributed scan of 4main.c:90-221: whole tables everywhere, then slices only
    c = bq.cumulative_integrate("profile", 0.0, 1800.0 / SCAN_N, SCAN_N, replicate=True, backend="cpu", threads=2)
    s = bq.cumulative_integrate("vel", 0.0, 1800.0 / SCAN_N, SCAN_N, replicate=False, backend="cpu", threads=2)
    scan = (float(c.P.sum()), float(c.Q.sum()), c.p_last, c.q_last, c.p_second_last, tuple(c.P.shape),
            s.slice, float(s.P.sum()), s.p_last, s.p_second_last, c.world)
    q.put((rank, r.value, r.value_lo, r.world, p, scan))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_world_matches_single_process(world):
    import b200quad as bq
    n = 3_000_017
    one = bq.integrate_result("sin", 0.0, math.pi, n, "left", backend="cpu")
    prof = bq.integrate("profile", 0.0, 1800.0, 180_000, "trap", backend="cpu")
    c1 = bq.cumulative_integrate("profile", 0.0, 1800.0 / SCAN_N, SCAN_N, backend="cpu")
    s1 = bq.cumulative_integrate("vel", 0.0, 1800.0 / SCAN_N, SCAN_N, backend="cpu")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, one.value, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, v, lo, w, pv, scan in got:
        assert w == world
        assert v == one.value and lo == one.value_lo       # bit-identical on every rank
        assert pv == prof
        psum, qsum, pl, ql, p2, shape, sl, spsum, spl, sp2, cw = scan
        assert cw == world and shape == (SCAN_N,)
        assert (psum, qsum, pl, ql, p2) == (float(c1.P.sum()), float(c1.Q.sum()), c1.p_last, c1.q_last, c1.p_second_last)
        assert spsum == float(s1.P[sl[0]:sl[1]].sum()) and (spl, sp2) == (s1.p_last, s1.p_second_last)
