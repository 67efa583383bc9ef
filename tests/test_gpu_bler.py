"""Statistical parity against the reference's GENERIC decoder (SURVEY 8c "secondary oracle", north_star: "against
turbodecoder_gen ... matching BER/BLER within stated tolerance on the same AWGN vectors").

tests/golden/bler_gen.npz holds, per (K, sigma) point, the block and bit error counts of the reference's full-length int16
max-log-MAP (srslte_tdec_init_manual(..., SRSLTE_TDEC_GENERIC), turbodecoder.c:173, turbodecoder_gen.c:59-203) and of its
AUTO SSE/AVX decoders on 2000 seeded code blocks after 8 half-iterations (oracle/gen_golden.py gen_bler). The CUDA path
decodes the SAME vectors. Stated tolerance:

  (1) against the AUTO decoders: error counts identical (the path is bit-exact with them);
  (2) against turbodecoder_gen, K >= 2048: |BLER_cuda - BLER_gen| <= 3 * sqrt(2 p (1 - p) / n) + 1 / n at every point
      (p the mean of the two; the binomial 3-sigma band of a difference), |BER_cuda - BER_gen| <= 25 % of BER_gen + 1e-5;
  (3) against turbodecoder_gen, K = 504 (8 windows of 63 steps with 40-step warm-ups: the reference's own SSE decoder gives
      up ~0.5 dB here, recorded in the fixture): the noise level at which BLER crosses 0.1 lies within 0.6 dB of the generic
      decoder's; for K >= 2048 within 0.05 dB.
"""
import importlib
import math

import numpy as np
import pytest

import _golden as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fec():
    pkg = importlib.import_module("empower-srslte_b200")
    f = pkg.FecCuda(0)
    yield f
    f.close()


_seen = {}


@pytest.mark.parametrize("case", list(G.bler_cases()), ids=lambda c: c["id"])
def test_bler_point_vs_generic_decoder(fec, case):
    K, n = case["K"], case["n"]
    bits, llr = G.bler_vectors(K, case["sigma_milli"], n, case["seed"])
    assert G.crc32(llr) == case["llrcrc"], "vector generator drifted from the one that made the fixture"
    res = fec.decode_cbs([K] * n, list(llr), case["nof"], early_stop=False)
    out = np.stack([r["out"] for r in res])
    err = np.unpackbits(out, axis=1) != bits
    blk, bit = int(err.any(axis=1).sum()), int(err.sum())
    _seen[(K, case["sigma_milli"])] = (blk / n, case["gen"][0] / n)
    assert (blk, bit) == case["auto"], ("differs from the reference's SSE/AVX decoder", blk, bit, case["auto"])
    if K >= 2048:
        p_c, p_g = blk / n, case["gen"][0] / n
        p = 0.5 * (p_c + p_g)
        assert abs(p_c - p_g) <= 3 * math.sqrt(2 * p * (1 - p) / n) + 1 / n, (p_c, p_g)
        b_c, b_g = bit / (n * K), case["gen"][1] / (n * K)
        assert abs(b_c - b_g) <= 0.25 * b_g + 1e-5, (b_c, b_g)


def _crossing_db(points, level=0.1):
    """noise level (dB, 20 log10 sigma) at which BLER crosses `level`, log-linear interpolation between the bracketing points"""
    pts = sorted(points)
    for (s0, p0), (s1, p1) in zip(pts, pts[1:]):
        if 0 < p0 <= level <= p1:
            f = math.log(level / p0) / math.log(p1 / p0)
            return 20 * math.log10((s0 + f * (s1 - s0)) / 1000.0)
    raise AssertionError(f"BLER {level} not bracketed by {pts}")


def test_waterfall_position_vs_generic_decoder():
    """runs after the points above (same module): the BLER = 0.1 crossing of the CUDA path against the generic decoder's"""
    if not _seen:
        pytest.skip("the per-point tests did not run")
    for K in sorted({k for k, _ in _seen}):
        cuda = [(s, v[0]) for (k, s), v in _seen.items() if k == K]
        gen = [(s, v[1]) for (k, s), v in _seen.items() if k == K]
        shift = abs(_crossing_db(cuda) - _crossing_db(gen))
        assert shift <= (0.6 if K < 2048 else 0.05), (K, shift)
        print(f"K={K}: BLER 0.1 crossing shifted by {shift:.3f} dB against turbodecoder_gen")
