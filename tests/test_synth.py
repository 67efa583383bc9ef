"""The numpy transport-block synthesiser used by bench.py (empower-srslte_b200/synth.py) against the reference's own
encoder: the golden fixtures hold the bits srslte_dlsch_encode emitted (sch.c:188-296) for every (re)transmission."""
import importlib

import numpy as np

import _golden as G
from _libs import P

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


def test_symbol_synthesis_matches_reference(reflib):
    """scrambling sequence, modulation mapping and the UL channel interleaver of the symbol synthesiser against the
    reference's srslte_sequence_LTE_pr / srslte_mod_modulate / ulsch_deinterleave"""
    from _libs import aligned
    rng = np.random.default_rng(8)
    for c_init, n in ((1, 40), (0x2345678, 90000), (synth.cinit_pdsch(0x1234, 1, 7, 301), 777)):
        want = np.zeros(n, np.uint8)
        assert reflib.ref_sequence(c_init, n, P(want)) == 0
        assert (synth.gold_sequence(c_init, n) == want).all()
    assert synth.cinit_pdsch(1234, 1, 13, 77) == reflib.ref_cinit_pdsch(1234, 1, 13, 77)
    assert synth.cinit_pusch(1234, 13, 77) == reflib.ref_cinit_pusch(1234, 13, 77)
    for mod in (1, 2, 4, 6):
        bits = rng.integers(0, 2, 64 * 6 * mod).astype(np.uint8)
        want = aligned(2 * bits.size // mod, np.float32)
        assert reflib.ref_modulate(mod, P(bits), P(want), bits.size) == bits.size // mod
        got = synth.modulate(bits, mod)
        assert np.allclose(got.view(np.float32), want, atol=1e-6), mod
    # interleave here, de-interleave there
    for mod, prb, cols, nri in ((2, 3, 12, 0), (4, 5, 12, 6), (6, 4, 11, 3)):
        rows = 12 * prb
        ri_pos = [((rows - 1 - i // 4) + rows * (1, 10, 7, 4)[i % 4]) * mod + k for i in range(nri) for k in range(mod)]
        g = rng.integers(1, 1000, rows * cols * mod - len(ri_pos)).astype(np.int16)
        q = synth.ulsch_interleave(g, mod, cols, ri_pos)
        back = np.zeros(q.size, np.int16)
        pos = np.array(ri_pos + [0], np.uint32)
        assert reflib.ref_ulsch_deinterleave(P(q), mod, rows * cols, cols, P(back), P(pos), len(ri_pos)) == 0
        assert (back[1: g.size] == g[1:]).all() and (nri or back[0] == g[0])
