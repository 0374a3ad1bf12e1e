"""Exact running sums for the scan tests.

A naive fp64 `cumsum` drifts by ~1e-13 over 10^4..10^7 terms, which is worse than the library it
is supposed to judge (dd-compensated, ~1e-16).  This oracle is exact up to the final rounding:

  v = H / 2^20 + r          H = rint(v * 2^20) as int64, r = v - H / 2^20 (exact, |r| <= 2^-21)
  P = cumsum(H) / 2^20 + cumsum(r)                 integer part exact; the r part in 80-bit
  Q = cumsum(cumsum(H)) / 2^20 + cumsum(cumsum(r)) integer part exact through 2048-sample blocks
                                                   (int64 inside a block, Python ints across)

The r parts are 2^-21 of the whole, summed in x87 extended precision (2^-64 per operation), so
their error is ~1e-19 of the result.  Values must satisfy |v| < 2^11 and n <= 2^25.
"""
import numpy as np

_SCALE = float(1 << 20)
_BLOCK = 2048


def exact_prefix(v):
    """v: 1-D float64 array.  Returns (P, Q) as np.longdouble arrays."""
    v = np.ascontiguousarray(np.asarray(v, dtype=np.float64))
    n = v.shape[0]
    assert n <= (1 << 25) and (n == 0 or float(np.abs(v).max()) < 2048.0)
    ld = np.longdouble
    Hf = np.rint(v * _SCALE)
    r = v - Hf / _SCALE                                  # exact
    H = Hf.astype(np.int64)
    nb = (n + _BLOCK - 1) // _BLOCK
    Hp = np.zeros(nb * _BLOCK, dtype=np.int64)
    Hp[:n] = H
    Hp = Hp.reshape(nb, _BLOCK)
    p_local = np.cumsum(Hp, axis=1)                      # |.| <= 2^11 * 2^20 * 2^11 = 2^42
    q_local = np.cumsum(p_local, axis=1)                 # <= 2^53
    blk_tot = p_local[:, -1]
    PB = np.concatenate(([0], np.cumsum(blk_tot)[:-1]))  # prefix before each block, <= 2^56
    j1 = np.arange(1, _BLOCK + 1, dtype=np.int64)
    assert float(np.abs(PB).max(initial=0)) * _BLOCK < 2.0 ** 62    # the next line stays inside int64
    t = PB[:, None] * j1[None, :] + q_local
    # block bases of Q as exact Python ints, split into two 63-bit-safe longdouble words
    qb_hi = np.zeros(nb, dtype=ld)
    qb_lo = np.zeros(nb, dtype=ld)
    acc = 0
    for b in range(nb):
        hi = (acc >> 40) << 40
        assert int(float(hi)) == hi                       # a multiple of 2^40 below 2^93: exact in fp64
        qb_hi[b] = ld(float(hi))
        qb_lo[b] = ld(acc - hi)
        acc += int(PB[b]) * _BLOCK + int(q_local[b, -1])
    P_main = (PB[:, None] + p_local).astype(ld) / ld(_SCALE)
    Q_main = (qb_hi[:, None] + (qb_lo[:, None] + t.astype(ld))) / ld(_SCALE)
    rl = r.astype(ld)
    P_r = np.cumsum(rl)
    Q_r = np.cumsum(P_r)
    P = P_main.reshape(-1)[:n] + P_r
    Q = Q_main.reshape(-1)[:n] + Q_r
    return P, Q


def max_rel_error(got, exact, floor=1e-300):
    """max |got - exact| / (running max of |exact|): relative to the size the sum has reached."""
    got = np.asarray(got, dtype=np.float64).astype(np.longdouble)
    scale = np.maximum.accumulate(np.abs(exact))
    scale = np.maximum(scale, np.longdouble(floor))
    return float(np.max(np.abs(got - exact) / scale)) if got.size else 0.0
