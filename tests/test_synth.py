"""The numpy transport-block synthesiser used by bench.py (empower-srslte_b200/synth.py) against the reference's own
encoder: the golden fixtures hold the bits srslte_dlsch_encode emitted (sch.c:188-296) for every (re)transmission."""
import importlib

import numpy as np

import _golden as G

synth = importlib.import_module("empower-srslte_b200.synth")


def test_encoder_matches_reference_bits():
    n = 0
    z = np.load(G.os.path.join(G.GOLD, "tb.npz"))
    for ci, c in enumerate(G.tb_cases()):
        payload = np.ascontiguousarray(c["tbdata"][: c["tbs"] // 8])[None, :]
        for tx, t in enumerate(c["tx"]):
            want = np.unpackbits(z[f"e_{ci}_{tx}"])[: c["nbits"]]
            got = synth.encode_tbs(payload, c["tbs"], c["Qm"], t["rv"], c["nbits"])[0]
            assert (got == want).all(), (c["id"], tx)
            n += 1
    assert n >= 4


def test_segmentation_matches_library():
    pkg = importlib.import_module("empower-srslte_b200")
    for tbs in (40, 6120, 6144, 6200, 12960, 36696, 75376):
        a, b = synth.cbsegm(tbs), pkg.cbsegm(tbs)
        assert all(a[k] == b[k] for k in ("C", "K1", "K2", "C1", "C2", "F")), tbs


def test_crc_known_answer():
    """crc_test.h:40-48 style: CRC over data+CRC is zero, and the two LTE polynomials differ"""
    rng = np.random.default_rng(5)
    d = rng.integers(0, 256, (3, 77), dtype=np.uint8)
    for poly in (synth.CRC24A, synth.CRC24B):
        c = synth.crc24(poly, d)
        full = np.concatenate([d, np.stack([(c >> 16) & 255, (c >> 8) & 255, c & 255], axis=1).astype(np.uint8)], axis=1)
        assert (synth.crc24(poly, full) == 0).all()
    assert (synth.crc24(synth.CRC24A, d) != synth.crc24(synth.CRC24B, d)).any()


def test_llr_convention_and_contiguity():
    e = np.array([[0, 1, 1, 0]], np.uint8)
    llr = synth.awgn_llr(e, 0.0, np.random.default_rng(0))
    assert llr.flags["C_CONTIGUOUS"] and llr.dtype == np.int16 and llr.tolist() == [[-100, 100, 100, -100]]
