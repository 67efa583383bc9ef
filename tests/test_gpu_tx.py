"""Transmit side on the GPU: fec_cuda_encode_tbs against the bits srslte_dlsch_encode emitted (golden
fixtures), against the oracle's encoder (pinned to the reference in test_oracle_vs_ref.py) and, at the bench's full
shape, through the encode -> decode round trip."""
import ctypes as C
import importlib
import os

import numpy as np
import pytest

import _golden as G
from _libs import P

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fec():
    pkg = importlib.import_module("empower-srslte_b200")
    f = pkg.FecCuda(0)
    yield f
    f.close()


def test_encode_golden(fec):
    """every (re)transmission of the golden transport blocks: the packed bits of the reference encoder (sch.c:188-296)"""
    z = np.load(os.path.join(G.GOLD, "tb.npz"))
    n = 0
    for ci, c in enumerate(G.tb_cases()):
        data = c["tbdata"][: c["tbs"] // 8]
        for tx, t in enumerate(c["tx"]):
            ret, e = fec.encode_tb(c["tbs"], c["Qm"], t["rv"], c["nbits"], data)
            want = z[f"e_{ci}_{tx}"]
            assert ret == 0
            assert (np.unpackbits(e)[: c["nbits"]] == np.unpackbits(want)[: c["nbits"]]).all(), (c["id"], tx)
            n += 1
    assert n >= 4


SHAPES = [(16, 2, 120), (40, 2, 288), (504, 2, 1200), (1000, 4, 1600), (6120, 4, 9600), (6200, 6, 14400), (12960, 2, 20000),
          (36696, 4, 57600), (61664, 6, 79998), (75376, 6, 90000), (75376, 6, 86400), (75376, 2, 400000)]


def test_encode_batch_vs_oracle(fec, orc):
    """one call with mixed shapes: one and many code blocks, every rv, E not a multiple of 8 (code blocks share output
    bytes), E > 3K+12 (the circular buffer wraps), all in one batch"""
    pkg = importlib.import_module("empower-srslte_b200")
    rng = np.random.default_rng(21)
    jobs = [(tbs, Qm, nb, rv) for (tbs, Qm, nb) in SHAPES for rv in range(4)]
    d = (pkg.TxTbDesc * len(jobs))()
    keep = []
    for i, (tbs, Qm, nb, rv) in enumerate(jobs):
        data = rng.integers(0, 256, tbs // 8).astype(np.uint8)
        out = np.full((nb + 7) // 8, 0xA5, np.uint8)          # stale content must not survive
        keep.append((data, out))
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits = tbs, Qm, rv, nb
        d[i].data, d[i].e_bits = data.ctypes.data, out.ctypes.data
    fec.encode_tbs(d)
    for i, (tbs, Qm, nb, rv) in enumerate(jobs):
        want = np.zeros((nb + 7) // 8, np.uint8)
        assert orc.orc_tb_encode(tbs, Qm, rv, nb, P(keep[i][0]), P(want)) == 0
        assert d[i].ret == 0
        assert (np.unpackbits(keep[i][1])[:nb] == np.unpackbits(want)[:nb]).all(), (tbs, Qm, nb, rv)


def test_encode_all_k_vs_oracle(fec, orc):
    """every one of the 188 code-block sizes as a single-block transport block (tbs = K - 24), rv 0 and 2"""
    pkg = importlib.import_module("empower-srslte_b200")
    from _libs import ALL_K
    rng = np.random.default_rng(22)
    jobs = [(K - 24, 2, 3 * K + 12 + 40, rv) for K in ALL_K for rv in (0, 2)]
    d = (pkg.TxTbDesc * len(jobs))()
    keep = []
    for i, (tbs, Qm, nb, rv) in enumerate(jobs):
        data = rng.integers(0, 256, tbs // 8).astype(np.uint8)
        out = np.zeros((nb + 7) // 8, np.uint8)
        keep.append((data, out))
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits = tbs, Qm, rv, nb
        d[i].data, d[i].e_bits = data.ctypes.data, out.ctypes.data
    fec.encode_tbs(d)
    bad = []
    for i, (tbs, Qm, nb, rv) in enumerate(jobs):
        want = np.zeros((nb + 7) // 8, np.uint8)
        assert orc.orc_tb_encode(tbs, Qm, rv, nb, P(keep[i][0]), P(want)) == 0
        if d[i].ret != 0 or not (np.unpackbits(keep[i][1])[:nb] == np.unpackbits(want)[:nb]).all():
            bad.append((tbs + 24, rv))
    assert not bad, bad


def test_encode_decode_round_trip_full_shape(fec):
    """BASELINE configs[1] shape, 296 transport blocks: encode on the GPU, noiseless LLRs at the demodulator's 64QAM scale
    (bit 1 -> +700), decode on the GPU: every payload comes back, one half-iteration per code block"""
    pkg = importlib.import_module("empower-srslte_b200")
    TBS, QM, NB, n = 75376, 6, 90000, 296
    rng = np.random.default_rng(23)
    data = rng.integers(0, 256, (n, TBS // 8)).astype(np.uint8)
    e = np.zeros((n, NB // 8), np.uint8)
    d = (pkg.TxTbDesc * n)()
    for i in range(n):
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits = TBS, QM, 0, NB
        d[i].data, d[i].e_bits = data[i].ctypes.data, e[i].ctypes.data
    fec.encode_tbs(d)
    assert all(d[i].ret == 0 for i in range(n))
    llr = (1400 * np.unpackbits(e, axis=1).astype(np.int16) - 700).astype(np.int16)
    out = np.zeros((n, TBS // 8 + 3 + 768), np.uint8)
    sb = fec.softbuffer_alloc(n)
    r = (pkg.TbDesc * n)()
    for i in range(n):
        r[i].tbs, r[i].Qm, r[i].rv, r[i].nof_e_bits, r[i].max_its = TBS, QM, 0, NB, 8
        r[i].softbuffer, r[i].new_data = sb + i, 1
        r[i].e_bits, r[i].data = llr[i].ctypes.data, out[i].ctypes.data
    fec.decode_tbs(r)
    assert all(r[i].ret == 0 for i in range(n))
    assert (out[:, : TBS // 8] == data).all()
    assert all(r[i].avg_its == 1 for i in range(n))


def test_encode_invalid(fec):
    data = np.zeros(10000, np.uint8)
    assert fec.encode_tb(75376, 6, 4, 90000, data)[0] == -2           # rv > 3
    assert fec.encode_tb(75376, 0, 0, 90000, data)[0] == -2           # Qm == 0
    assert fec.encode_tb(75380, 6, 0, 90000, data)[0] == -2           # not a multiple of 8
    assert fec.encode_tb(6200 - 8, 6, 0, 14400, data)[0] == -1        # needs filler bits: sch.c:202-205
    ret, e = fec.encode_tb(0, 2, 0, 64, data)                          # nothing to send
    assert ret == 0 and not e.any()
