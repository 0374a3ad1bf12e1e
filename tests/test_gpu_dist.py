"""One process per GPU (the torchrun layout): NCCL only bootstraps, the combine is in-kernel."""
import math
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

from conftest import gpu_count

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


SMALL_N = (200_000, 123_457)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import b200quad as bq
    vals = []
    for _ in range(50):                                   # back-to-back collective calls
        r = bq.integrate_result("sin", 0.0, math.pi, n, "left")
        vals.append((r.value, r.value_lo, r.status, r.world))
    c = bq.cumulative_integrate(n=3_000_000, dx=6e-4, replicate=True)
    c2 = bq.cumulative_integrate(n=3_000_000, dx=6e-4, replicate=False)
    # small replicated tables, twice with different n: P and Q of one call share an allocator segment
    # (and reuse cached blocks on the second call), which is where an IPC offset bug would show
    small = []
    for ns in SMALL_N:
        cs = bq.cumulative_integrate(n=ns, dx=1800.0 / ns, replicate=True)
        small.append((float(cs.P.sum()), float(cs.Q.sum()), float(cs.P[-1]), float(cs.Q[ns // 2])))
    # the torch op with the world size as an input (riemann.cpp:63-65): this rank's kernel of the
    # collective call, on torch's current stream, every rank's tensor holding the same bits
    from b200quad.ops import torch_ops
    t = torch_ops.integrate("sin", 0.0, math.pi, n, world=world)
    tP, tQ = torch_ops.cumulative("profile", 0.0, 6e-4, 3_000_000, world=world)
    op = (float(t), t.device.index, tuple(tP.shape), float(tP.sum()))
    q.put((rank, vals, c.p_last, c.p_second_last, float(c.P.sum()), float(c.Q[-1]), c2.slice,
           float(c2.P.sum()), small, op))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_torchrun_layout(world):
    if gpu_count() < world:
        pytest.skip("needs %d GPUs" % world)
    import b200quad as bq
    n = (1 << 27) + 3
    one = bq.integrate_result("sin", 0.0, math.pi, n, "left", world=1)
    c1 = bq.cumulative_integrate(n=3_000_000, dx=6e-4, world=1)
    small1 = []
    for ns in SMALL_N:
        cs = bq.cumulative_integrate(n=ns, dx=1800.0 / ns, world=1)
        small1.append((float(cs.P.sum()), float(cs.Q.sum()), float(cs.P[-1]), float(cs.Q[ns // 2])))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    total_slice = 0.0
    for rank, vals, p_last, p2, psum, qlast, sl, p2sum, small, op in got:
        assert small == small1
        assert op[0] == one.value and op[1] == rank                  # torch op, world = the process group
        assert op[2] == (sl[1] - sl[0],) and op[3] == p2sum          # its cumulative returns this rank's slice
        for v, lo, status, w in vals:
            assert status == 0 and w == world
            assert (v, lo) == (one.value, one.value_lo)
        assert p_last == c1.p_last and p2 == c1.p_second_last
        assert psum == float(c1.P.sum()) and qlast == float(c1.Q[-1])
        assert float(c1.P[sl[0]:sl[1]].sum()) == p2sum
