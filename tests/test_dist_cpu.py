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
    q.put((rank, r.value, r.value_lo, r.world, p))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_world_matches_single_process(world):
    import b200quad as bq
    n = 3_000_017
    one = bq.integrate_result("sin", 0.0, math.pi, n, "left", backend="cpu")
    prof = bq.integrate("profile", 0.0, 1800.0, 180_000, "trap", backend="cpu")
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
    for rank, v, lo, w, pv in got:
        assert w == world
        assert v == one.value and lo == one.value_lo       # bit-identical on every rank
        assert pv == prof
