"""Range condition behind the packed CUDA fast path (fec_tdec_packed.cu): if
max|x| + max|y| <= FEC_PK_GMAX over a half-iteration (x = systematic + apriori, y = parity, tails included),
no saturating operation of the reference's windowed MAP decoder clips, so wrapping int16x2 arithmetic gives
the same numbers. Checked here on the instrumented CPU oracle (which is pinned bit-exact to the reference)
with random and adversarial inputs at the bound."""
import ctypes as C

import numpy as np
import pytest

from _libs import P, int_gauss

GMAX = 2000


def _patterns(K, bx, by, seed):
    g = int_gauss(2 * K + 12, seed)
    r = np.random.default_rng(seed)
    yield "uniform", (g[:K] % (2 * bx + 1) - bx), (g[K:2 * K] % (2 * by + 1) - by)
    yield "all+", np.full(K, bx), np.full(K, by)
    yield "all-", np.full(K, -bx), np.full(K, -by)
    yield "x+y-", np.full(K, bx), np.full(K, -by)
    yield "x-y+", np.full(K, -bx), np.full(K, by)
    s1, s2 = r.integers(0, 2, K) * 2 - 1, r.integers(0, 2, K) * 2 - 1
    yield "randsign", bx * s1, by * s2
    alt = np.where(np.arange(K) % 2 == 0, 1, -1)
    yield "alt", bx * alt, by * alt
    alt3 = np.where((np.arange(K) // 3) % 2 == 0, 1, -1)
    yield "alt3", bx * alt3, -by * alt3
    run = np.where((np.arange(K) // 41) % 2 == 0, 1, -1)
    yield "runs41", bx * run, by * s2


@pytest.mark.parametrize("K", [408, 800, 816, 1008, 2048, 5824, 6144])
def test_no_saturation_within_gmax(orc, K):
    orc.orc_dbg_sat_count.restype = C.c_long
    Pn = orc.orc_subblocks(K)
    assert Pn in (8, 16)
    for bx, by in [(GMAX // 2, GMAX // 2), (GMAX - 300, 300), (300, GMAX - 300), (GMAX, 0), (GMAX - 1, 1)]:
        for name, x, y in _patterns(K, bx, by, K + bx):
            x = np.ascontiguousarray(x, dtype=np.int16)
            y = np.ascontiguousarray(y, dtype=np.int16)
            m = min(bx, by)                                   # |tail| <= max|soft entry| <= both bounds' owner
            tx = np.array([m, -m, m], np.int16)
            tp = np.array([-m, m, m], np.int16)
            out = np.zeros(K, np.int16)
            orc.orc_dbg_reset()
            orc.orc_map(K, Pn, P(x), None, P(y), P(tx), P(tp), P(out))
            assert orc.orc_dbg_sat_count() == 0, (K, bx, by, name, orc.orc_dbg_peak_val())
            assert orc.orc_dbg_peak_val() <= 10000 + 10 * GMAX


def test_saturation_does_occur_far_above_gmax(orc):
    """sanity of the instrumentation: large inputs do clip"""
    orc.orc_dbg_sat_count.restype = C.c_long
    K = 6144
    x = np.full(K, 9000, np.int16)
    y = np.full(K, 9000, np.int16)
    t = np.zeros(3, np.int16)
    out = np.zeros(K, np.int16)
    orc.orc_dbg_reset()
    orc.orc_map(K, 16, P(x), None, P(y), P(t), P(t), P(out))
    assert orc.orc_dbg_sat_count() > 0
