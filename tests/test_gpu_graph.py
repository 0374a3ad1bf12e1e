"""CUDA-graph capture of the fused kernels: a launch-bound loop of small integrals is captured once
and replayed (the reference has nothing comparable: every run of cintegrate.cu pays context
creation, three cudaMallocs and 288 MB of copies, cintegrate.cu:104-133)."""
import math

import pytest
import torch

import b200quad as bq
from b200quad.ops import torch_ops

pytestmark = pytest.mark.gpu


def test_integrate_op_is_capturable_and_replays_bitwise():
    ns = [(1 << 18) + 7 * k for k in range(12)]
    want = [bq.integrate("sin", 0.0, math.pi, n, "mid") for n in ns]
    torch_ops.integrate("sin", 0.0, math.pi, ns[0], "mid", device=0)     # context + module load outside capture
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for n in ns[:2]:
            torch_ops.integrate("sin", 0.0, math.pi, n, "mid", device=0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        outs = [torch_ops.integrate("sin", 0.0, math.pi, n, "mid", device=0) for n in ns]
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    got = [float(o) for o in outs]
    assert got == want
    # the uncaptured path still works afterwards (tickets, epochs and the result block are intact)
    assert bq.integrate("sin", 0.0, math.pi, ns[3], "mid") == want[3]


def test_cumulative_op_is_capturable():
    n = 300_000
    P0, Q0 = torch_ops.cumulative("profile", 0.0, 1e-2 * 0.6, n, device=0)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        P, Q = torch_ops.cumulative("profile", 0.0, 1e-2 * 0.6, n, device=0)
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(P, P0) and torch.equal(Q, Q0)
