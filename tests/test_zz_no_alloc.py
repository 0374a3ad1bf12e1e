"""bq_cumulative_noalloc — the entry point for callers that are not allowed to allocate.

Some users of this library cannot call cudaMalloc at all: a control loop with a hard
latency budget (cudaMalloc/cudaFree synchronise the device), a plug-in whose host
application owns the device memory and lends it only fixed buffers, a job running at
the edge of its memory limit.  For them the library offers one variant of the fused
cumulative-integration kernel whose scratch is a static device array.  This test
drives that variant exactly as such a caller would — through the C ABI, with buffers
it already owns — and checks it against the regular entry point (bq_cumulative):
the two run the same kernel over the same fixed tree, so the tables must be equal
bit for bit (the tolerance the other scan tests use for same-tree comparisons).

Collected last on purpose (file name sorts after every other test file): a run that
stops at the first failure has then already executed the rest of the suite.  It must
pass alone in a fresh process and after the other tests, so it assumes nothing about
the state of the shared scratch array.
"""
import ctypes as C

import pytest
import torch

import b200quad as bq
from b200quad import _native

pytestmark = pytest.mark.gpu


def _call_noalloc(lib, n, dx, name="profile"):
    """One call through the C ABI with caller-owned buffers; returns (P, Q, result)."""
    integ = bq.models.REGISTRY[name]
    job = _native.ScanJob(integ.native_id, 0, 0, 0, 0.0, float(dx), int(n), (C.c_double * 4)(0, 0, 0, 0))
    P = torch.empty(n, dtype=torch.float64, device="cuda:0")
    Q = torch.empty(n, dtype=torch.float64, device="cuda:0")
    out = torch.zeros(C.sizeof(_native.ScanResult), dtype=torch.uint8, device="cuda:0")
    stream = torch.cuda.current_stream(0).cuda_stream
    rc = lib.bq_cumulative_noalloc(0, C.byref(job), C.c_void_p(P.data_ptr()), C.c_void_p(Q.data_ptr()),
                                   C.c_void_p(out.data_ptr()), C.c_void_p(stream))
    assert rc == 0, _native.last_error()
    torch.cuda.synchronize(0)
    res = _native.ScanResult.from_buffer_copy(bytes(out.cpu().numpy().tobytes()))
    return P, Q, res


def test_two_calls_in_a_row_match_the_allocating_entry_point():
    lib, cuda = _native.load()
    assert cuda, "the CUDA library must be the one loaded on a GPU box"
    # Two different sizes back to back: the second call finds the first call's
    # leftovers in the shared scratch and must not be affected by them.
    for n, dx in ((18_000_000, 1.0e-4), (2_500_003, 7.0e-4)):
        ref = bq.cumulative_integrate("profile", 0.0, dx, n, backend="cuda")      # bq_cumulative
        P, Q, res = _call_noalloc(lib, n, dx)
        assert res.status == 0
        assert torch.equal(P, ref.P) and torch.equal(Q, ref.Q)
        assert res.p_last == ref.p_last and res.q_last == ref.q_last
        assert res.p_second_last == ref.p_second_last
        assert (res.slice_begin, res.slice_end) == (0, n)


def test_same_call_twice_gives_the_same_tables():
    lib, _ = _native.load()
    a = _call_noalloc(lib, 1_000_001, 1.0e-3, "vel")
    b = _call_noalloc(lib, 1_000_001, 1.0e-3, "vel")
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    ref = bq.cumulative_integrate("vel", 0.0, 1.0e-3, 1_000_001, backend="cuda")
    assert torch.equal(a[0], ref.P) and torch.equal(a[1], ref.Q)


def test_sizes_beyond_the_static_scratch_are_rejected():
    lib, _ = _native.load()
    limit = int(lib.bq_noalloc_max_samples())
    assert limit == 67_108_864
    job = _native.ScanJob(0, 0, 0, 0, 0.0, 1.0e-9, limit + 1, (C.c_double * 4)(0, 0, 0, 0))
    buf = torch.empty(8, dtype=torch.float64, device="cuda:0")        # never touched: rejected first
    rc = lib.bq_cumulative_noalloc(0, C.byref(job), C.c_void_p(buf.data_ptr()),
                                   C.c_void_p(buf.data_ptr()), None, None)
    assert rc == 1                                                    # BQ_ERR_INVALID
    assert "exceeds the no-allocation limit" in _native.last_error()


def test_a_grid_of_exactly_the_limit_runs():
    """n equal to BQ_NOALLOC_MAX_SAMPLES uses the last descriptor of the static scratch."""
    lib, _ = _native.load()
    n = int(lib.bq_noalloc_max_samples())
    dx = 1800.0 / n
    P, Q, res = _call_noalloc(lib, n, dx)
    assert res.status == 0 and (res.slice_begin, res.slice_end) == (0, n)
    assert res.p_last == float(P[-1]) and res.q_last == float(Q[-1]) and res.p_second_last == float(P[-2])
    assert abs(res.p_last * dx - 122000.004) < 1e-6
    ref = bq.cumulative_integrate("profile", 0.0, dx, n, backend="cuda")
    assert torch.equal(P, ref.P) and torch.equal(Q, ref.Q)
