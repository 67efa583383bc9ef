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


def pk_certificate(Ra, Rb, V, G):
    """the run-time range certificate of the packed kernel's TRK variant (fec_tdec_packed.cu pk_certificate)"""
    R = max(Ra, Rb)
    return 2 * R + 6 * G <= 32767 and Ra + Rb + 5 * G <= 32767 and 10000 + Rb + 5 * G <= 32767 and V <= 32767


@pytest.mark.parametrize("K", [408, 800, 816, 1008, 1056, 5824, 6144])
def test_certificate_implies_no_saturation(orc, K):
    """beyond GMAX the packed kernel relies on a certificate computed from measured state-metric magnitudes. Soundness,
    checked on the instrumented oracle: whenever the certificate holds, not one saturating operation of the reference
    clipped. Inputs run from just above GMAX to far beyond the certificate's reach, so that both outcomes occur."""
    orc.orc_dbg_sat_count.restype = C.c_long
    Pn = orc.orc_subblocks(K)
    held = clipped = 0
    for G in (2100, 2600, 3200, 3800, 4400, 5200, 6500, 9000, 14000):
        for bx, by in [(G // 2, G - G // 2), (G - 400, 400), (400, G - 400), (2 * G // 3, G - 2 * G // 3)]:
            pats = list(_patterns(K, bx, by, K + G + bx))
            r = np.random.default_rng(K + G + by)
            # codeword-like: mostly small values, rare extremes (what a demodulator produces)
            gx = np.clip(np.rint(r.standard_normal(K) * bx / 4), -bx, bx); gx[::97] = bx
            gy = np.clip(np.rint(r.standard_normal(K) * by / 4), -by, by); gy[::89] = -by
            pats.append(("gauss", gx, gy))
            for name, x, y in pats:
                x = np.ascontiguousarray(x, dtype=np.int16)
                y = np.ascontiguousarray(y, dtype=np.int16)
                m = min(bx, by)
                tx = np.array([m, -m, m], np.int16)
                tp = np.array([-m, m, m], np.int16)
                out = np.zeros(K, np.int16)
                orc.orc_dbg_reset()
                orc.orc_map(K, Pn, P(x), None, P(y), P(tx), P(tp), P(out))
                Ra = max(orc.orc_dbg_mag(2), orc.orc_dbg_mag(3))
                Rb = max(orc.orc_dbg_mag(0), orc.orc_dbg_mag(1))
                ok = pk_certificate(Ra, Rb, orc.orc_dbg_llr_range(), G)
                sat = orc.orc_dbg_sat_count()
                if ok:
                    held += 1
                    assert sat == 0, (K, G, bx, by, name, Ra, Rb, orc.orc_dbg_peak_val())
                    assert orc.orc_dbg_peak_val() <= 32767
                clipped += sat > 0
    assert held > 20 and clipped > 20, (held, clipped)
