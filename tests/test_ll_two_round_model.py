"""Index arithmetic of the two-round in-kernel all-reduce
(``mgb_ll_allreduce_two_round`` in ``csrc/mgb_common.cuh``), re-stated in
Python: slot units, slice ownership, ragged and empty slices, and the claim
that the totals equal the one-round rank-order sum bit for bit.  The CUDA code
itself is exercised by the multi-GPU cases of ``tests/_dist_worker.py`` (4 and
8 GPUs, 660 / 1300 doubles); this model covers the shapes those do not."""
import numpy as np
import pytest


def simulate(m, W, rng):
    c = (m + W - 1) // W                       # slice length
    vals = rng.standard_normal((W, m))
    slots = np.full((W, W, 4 * c), np.nan)     # [dst][src][unit]
    # round 1: my contribution to slice d goes to rank d, units [0, 2c)
    for me in range(W):
        for i in range(2 * c * (W - 1)):
            pi, u = divmod(i, 2 * c)
            dst = (me + 1 + pi) % W
            j = dst * c + (u >> 1)
            if j < m:                           # lo word: +v, hi word: -v
                slots[dst, me, u] = vals[me, j] * (1 if u % 2 == 0 else -1)
    out = np.zeros((W, m))
    for me in range(W):
        lo_j = me * c
        mine = (c if lo_j + c <= m else m - lo_j) if lo_j < m else 0
        red = np.zeros(mine)
        for j in range(mine):
            acc = 0.0
            for s in range(W):                  # rank order
                if s == me:
                    acc += vals[me, lo_j + j]
                else:
                    lo, hi = slots[me, s, 2 * j], slots[me, s, 2 * j + 1]
                    assert not np.isnan(lo) and hi == -lo
                    acc += lo
            red[j] = acc
        # round 2: the sums of my slice go to everybody, units [2c, 4c)
        for i in range(2 * mine * (W - 1)):
            pi, u = divmod(i, 2 * mine)
            dst = (me + 1 + pi) % W
            slots[dst, me, 2 * c + u] = red[u >> 1] * (1 if u % 2 == 0 else -1)
        out[me, lo_j:lo_j + mine] = red
    for me in range(W):
        for s in range(W):
            if s == me:
                continue
            for j in range(c):
                g = s * c + j
                if g < m:
                    lo = slots[me, s, 2 * c + 2 * j]
                    assert not np.isnan(lo)
                    assert slots[me, s, 2 * c + 2 * j + 1] == -lo
                    out[me, g] = lo
    want = np.zeros(m)
    for s in range(W):
        want = want + vals[s]
    for me in range(W):
        assert np.array_equal(out[me], want)
    return 4 * c


@pytest.mark.parametrize("W", [3, 4, 5, 7, 8, 16])
def test_two_round_exchange_model(W):
    rng = np.random.default_rng(W)
    for m in (192, 193, 255, 660, 1301, 97 * W + 1):
        units = simulate(m, W, rng)
        assert units <= 2 * m          # fits the slots of the one-round form
