"""Pins the front-end part of the C restatement (soft demodulation, scrambling sequence, descrambling, CSI weighting,
UL-SCH channel de-interleaver: oracle/fec_oracle.c, section "front end") against the UNMODIFIED reference
(oracle/_ref, ref_shim_fe.c).  Skipped where the reference build is absent."""
import numpy as np
import pytest

from _libs import P, aligned


def _symbols(n, rng, big=False):
    s = aligned(2 * n, np.float32)
    s[:] = rng.standard_normal(2 * n).astype(np.float32) * (0.8 if not big else 30.0)
    if big:
        s[::97] = 1e12; s[5::101] = -1e12; s[7::89] = 46.8; s[11::83] = -46.9
    return s


@pytest.mark.parametrize("mod", [1, 2, 4, 6])
def test_demod_matches_reference(orc, reflib, mod):
    """every modulation, lengths that leave 0..3 symbols (0..15 values for QPSK) to the scalar tails, values that hit
    the rounding ties and (in the SIMD bodies) the saturation"""
    rng = np.random.default_rng(mod)
    for n in (1, 3, 4, 7, 8, 9, 16, 33, 250, 1001, 15000):
        for big in (False, True):
            sym = _symbols(n, rng, big)
            if not big:
                sym[: min(2 * n, 40)] = (np.arange(min(2 * n, 40)) - 20.5) / 1400.0     # x.5 / 700: exact rounding ties
            a, b = aligned(mod * n + 32, np.int16), aligned(mod * n + 32, np.int16)
            assert reflib.ref_demod_s(mod, P(sym), P(a), n) == 0
            assert orc.orc_demod_s(mod, P(sym), P(b), n) == 0
            if big and mod in (1, 2):
                # float -> short conversion of out-of-range values in the reference's scalar loops is undefined C: bodies only
                m = 16 * ((2 * n) // 16) if mod == 2 else 0
                assert (a[:m] == b[:m]).all(), (mod, n)
            elif big:
                m = mod * 4 * (n // 4)
                assert (a[:m] == b[:m]).all(), (mod, n)
            else:
                assert (a == b).all(), (mod, n, np.nonzero(a != b)[0][:8])


def test_sequence_and_seeds(orc, reflib):
    rng = np.random.default_rng(5)
    for c_init in [0, 1, 0x7FFFFFFF, 0x12345678 & 0x7FFFFFFF] + [int(v) for v in rng.integers(0, 2**31, 6)]:
        for n in (1, 31, 32, 1000, 90000):
            a, b = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
            assert reflib.ref_sequence(c_init, n, P(a)) == 0
            orc.orc_sequence(c_init, n, P(b))
            assert (a == b).all(), (c_init, n)
    for rnti, q, nslot, cell in ((1234, 0, 0, 1), (0xFFFF, 1, 18, 503), (70, 0, 5, 0), (0x1234, 1, 19, 277)):
        assert orc.orc_cinit_pdsch(rnti, q, nslot, cell) == reflib.ref_cinit_pdsch(rnti, q, nslot, cell)
        assert orc.orc_cinit_pusch(rnti, nslot, cell) == reflib.ref_cinit_pusch(rnti, nslot, cell)


def test_descramble(orc, reflib):
    rng = np.random.default_rng(6)
    for n in (5, 16, 17, 90000):
        x = aligned(n, np.int16)
        x[:] = rng.integers(-32768, 32768, n)
        x[0] = -32768
        y = x.copy()
        c = np.zeros(n, np.uint8)
        orc.orc_sequence(777, n, P(c))
        assert reflib.ref_descramble_s(777, P(x), n) == 0
        orc.orc_descramble_s(P(c), P(y), n)
        assert (x == y).all()


@pytest.mark.parametrize("mod", [2, 4, 6])
def test_csi_weighting(orc, reflib, mod):
    """the SSE body (with its swapped weights inside a symbol pair) and the scalar tail"""
    rng = np.random.default_rng(10 + mod)
    for nsym in (2, 3, 4, 5, 100, 101, 15000):
        e = aligned(mod * nsym, np.int16)
        e[:] = rng.integers(-3000, 3000, mod * nsym)
        csi = aligned(nsym + 8, np.float32)
        csi[:nsym] = rng.uniform(0.05, 2.0, nsym).astype(np.float32)
        a, b = e.copy(), e.copy()
        a2 = aligned(mod * nsym, np.int16); a2[:] = a
        assert reflib.ref_csi_s(mod, P(a2), P(csi), mod * nsym) == 0
        orc.orc_csi_s(mod, P(b), P(csi), mod * nsym)
        assert (a2 == b).all(), (mod, nsym, np.nonzero(a2 != b)[0][:8])


def test_ulsch_deinterleave(orc, reflib):
    rng = np.random.default_rng(12)
    for Qm, prb, cols, nri_sym in ((2, 1, 12, 0), (4, 6, 12, 0), (4, 6, 12, 5), (6, 10, 11, 9), (2, 25, 12, 16), (4, 96, 12, 40)):
        rows = 12 * prb
        Hp = rows * cols
        q = rng.integers(-2000, 2000, Hp * Qm).astype(np.int16)
        ri_cols = (1, 10, 7, 4)                                                      # uci.c:528-535 visiting order (3*i)%4
        pos = []
        for i in range(nri_sym):
            row, col = rows - 1 - i // 4, ri_cols[i % 4]
            pos += [row * Qm + rows * col * Qm + k for k in range(Qm)]
        pos = np.array(pos + [0], np.uint32)
        a, b = np.full(Hp * Qm, 7, np.int16), np.full(Hp * Qm, 7, np.int16)
        assert reflib.ref_ulsch_deinterleave(P(q), Qm, Hp, cols, P(a), P(pos), len(pos) - 1) == 0
        orc.orc_ulsch_deinterleave(P(q), Qm, Hp, cols, P(b), P(pos), len(pos) - 1)
        assert (a == b).all(), (Qm, prb, cols, nri_sym)
        if nri_sym:
            assert a[0] == q[max(pos[:-1])]       # the quirk: g[0] holds the last RI position's LLR


# ---- whole chains: reference from the symbols on vs the oracle's call sequence + orc_tb_decode ----
import ctypes as C  # noqa: E402

import _fe  # noqa: E402


def _orc_decode(orc, tbs, Qm, rv, llr, max_its, ot=None):
    own = ot is None
    if own:
        ot = C.c_void_p(orc.orc_tb_new())
        orc.orc_tb_reset(ot, tbs)
    data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
    crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
    ret = orc.orc_tb_decode(ot, tbs, Qm, rv, llr.size, P(np.ascontiguousarray(llr)), max_its, P(data), P(crc), P(its))
    if own:
        orc.orc_tb_free(ot)
    return ret, data, crc


@pytest.mark.parametrize("tbs,mod,nof_re,sigma,use_csi", [(75376, 6, 15000, 0.05, False), (75376, 6, 15000, 0.06, True),
                                                           (6200, 4, 2401, 0.12, False), (1000, 2, 1201, 0.3, True),
                                                           (12216, 4, 3600, 0.13, True)])
def test_dl_chain_vs_reference(orc, reflib, tbs, mod, nof_re, sigma, use_csi):
    """srslte_pdsch_codeword_decode (pdsch.c:778-834) on synthetic PDSCH symbols, two HARQ transmissions"""
    rng = np.random.default_rng(tbs + mod)
    payload = rng.integers(0, 256, (1, tbs // 8), dtype=np.uint8)
    rnti, sf_idx, cell = 0x4601, 5, 33
    c_init = orc.orc_cinit_pdsch(rnti, 0, 2 * sf_idx, cell)
    cw = C.c_void_p(reflib.ref_cw_new(100, cell))
    reflib.ref_cw_reset(cw, tbs)
    ot = C.c_void_p(orc.orc_tb_new())
    orc.orc_tb_reset(ot, tbs)
    for rv in (0, 2):
        sym = _fe.sym_buffer(_fe.make_dl(payload, tbs, mod, rv, nof_re, c_init, sigma, rng)[0])
        csi = None
        if use_csi:
            csi = aligned(nof_re, np.float32)
            csi[:] = rng.uniform(0.3, 1.7, nof_re).astype(np.float32)
        data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
        ack = C.c_int(0)
        e_ref = np.zeros(nof_re * mod, np.int16)
        crc_ref = np.zeros(13, np.uint8)
        r = reflib.ref_cw_decode(cw, tbs, mod, rv, nof_re, rnti, sf_idx, P(sym), P(csi) if use_csi else None, 8, P(data),
                                 C.byref(ack), P(e_ref), P(crc_ref))
        assert r == 0
        e, _ = _fe.oracle_frontend(orc, sym, mod, c_init, csi)
        assert (e == e_ref).all(), np.nonzero(e != e_ref)[0][:8]
        ret, d2, crc = _orc_decode(orc, tbs, mod, rv, e, 8, ot)
        assert (ret == 0) == bool(ack.value)
        assert (d2[: tbs // 8 + 3] == data[: tbs // 8 + 3]).all() and (crc == crc_ref).all()
        if ret == 0:
            assert (data[: tbs // 8] == payload[0]).all()
            break                       # ACKed: a HARQ process would start a new transport block
    reflib.ref_cw_free(cw)
    orc.orc_tb_free(ot)


@pytest.mark.parametrize("tbs,mod,prb,cols,uci", [(15264, 4, 40, 12, (0, 0)), (15264, 4, 40, 12, (1, 1)), (2216, 2, 10, 12, (2, 1)),
                                                  (36696, 4, 96, 12, (1, 0)), (6200, 6, 16, 11, (0, 1))])
def test_ul_chain_vs_reference(orc, reflib, tbs, mod, prb, cols, uci):
    """pusch.c:617-657 from the demodulator on, with HARQ-ACK / RI multiplexed: positions, zeroing, de-interleaver (g[0]
    quirk included), decode_tb"""
    rng = np.random.default_rng(tbs + prb)
    nof_re = 12 * prb * cols
    nb = nof_re * mod
    c_init = orc.orc_cinit_pusch(0x77, 8, 101)
    ul = C.c_void_p(reflib.ref_ul_new(100))
    reflib.ref_ul_reset(ul, tbs)
    uci_len = np.array([uci[0], uci[1], 0], np.uint32)
    ioff = np.array([9, 6, 7], np.uint32)
    payload = rng.integers(0, 256, (1, tbs // 8), dtype=np.uint8)

    def ref_run(sym, rv):
        data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
        out_uci, out_cqi = np.zeros(4, np.uint8), np.zeros(64, np.uint8)
        q, g = np.zeros(nb, np.int16), np.zeros(nb, np.int16)
        crc = np.zeros(13, np.uint8)
        ap, rp, cnt = np.zeros(12 * 288, np.uint32), np.zeros(12 * 288, np.uint32), np.zeros(2, np.uint32)
        r = reflib.ref_ul_decode(ul, tbs, mod, rv, prb, cols, c_init, P(sym), P(uci_len), P(ioff), 8, P(data), P(out_uci), P(out_cqi),
                                 P(q), P(g), P(crc), P(ap), P(rp), P(cnt))
        return r, data, q, g, crc, list(ap[: cnt[0]]), list(rp[: cnt[1]])

    # dry run on noise: which positions does the reference use for this configuration?
    dry = _fe.sym_buffer(rng.standard_normal((nof_re, 2)).astype(np.float32))
    _, _, _, _, _, ack_pos, ri_pos = ref_run(dry, 0)
    reflib.ref_ul_reset(ul, tbs)
    assert (len(ack_pos) > 0) == (uci[0] > 0) and (len(ri_pos) > 0) == (uci[1] > 0)
    rows = 12 * prb
    if cols == 12:      # normal CP without SRS: the column sets the GPU tests assume
        assert ack_pos == _fe.uci_positions(len(ack_pos) // mod, mod, rows, _fe.ACK_COLS)
        assert ri_pos == _fe.uci_positions(len(ri_pos) // mod, mod, rows, _fe.RI_COLS)
    sym, G = _fe.make_ul(payload, tbs, mod, 0, prb, cols, c_init, 0.08 if mod == 6 else 0.15, rng, ri_pos)
    sym = _fe.sym_buffer(sym[0])
    r, data, q_ref, g_ref, crc_ref, ap2, rp2 = ref_run(sym, 0)
    assert ap2 == ack_pos and rp2 == ri_pos
    g, raw = _fe.oracle_frontend(orc, sym, mod, c_init, None, cols, ack_pos, ri_pos)
    ng = nb - len(ri_pos)
    assert (g[:ng] == g_ref[:ng]).all(), np.nonzero(g[:ng] != g_ref[:ng])[0][:8]
    ret, d2, crc = _orc_decode(orc, tbs, mod, 0, g[:G], 8)
    assert ret == r and (d2[: tbs // 8 + 3] == data[: tbs // 8 + 3]).all() and (crc == crc_ref).all()
    assert r == 0 and (data[: tbs // 8] == payload[0]).all()
    reflib.ref_ul_free(ul)


# ---- 8-bit LLR front end (q->llr_is_8bit: srslte_demod_soft_demodulate_b, srslte_scrambling_sb_offset, int8 CSI weighting) ----
@pytest.mark.parametrize("mod", [1, 2, 4, 6])
def test_demod8_matches_reference(orc, reflib, mod):
    """int8 demodulators: SSE bodies over 8 symbols (16 values for QPSK) with rounding / truncation and two saturating packs,
    scalar tails with C conversions; lengths that leave 0..7 symbols to the tails, rounding ties, saturation"""
    rng = np.random.default_rng(40 + mod)
    for n in (1, 3, 7, 8, 9, 15, 16, 17, 33, 250, 1001, 15000):
        for big in (False, True):
            sym = _symbols(n, rng, big)
            if not big:
                sym[: min(2 * n, 40)] = (np.arange(min(2 * n, 40)) - 20.5) / 80.0          # x.5 / 40: exact rounding ties at 64QAM
            a, b = aligned(mod * n + 64, np.int8), aligned(mod * n + 64, np.int8)
            assert reflib.ref_demod_b(mod, P(sym), P(a), n) == 0
            assert orc.orc_demod_b(mod, P(sym), P(b), n) == 0
            if big:
                # out-of-range float -> int8 conversions in the reference's scalar loops are undefined C: SIMD bodies only
                m = 16 * ((2 * n) // 16) if mod == 2 else 0 if mod == 1 else mod * 8 * (n // 8)
                assert (a[:m] == b[:m]).all(), (mod, n, np.nonzero(a[:m] != b[:m])[0][:8])
            else:
                assert (a == b).all(), (mod, n, np.nonzero(a != b)[0][:8])


def test_descramble8(orc, reflib):
    rng = np.random.default_rng(46)
    for n in (5, 31, 32, 33, 90000):
        x = aligned(n, np.int8)
        x[:] = rng.integers(-128, 128, n)
        x[0] = -128
        y = x.copy()
        c = np.zeros(n, np.uint8)
        orc.orc_sequence(777, n, P(c))
        assert reflib.ref_descramble_b(777, P(x), n) == 0
        orc.orc_descramble_b(P(c), P(y), n)
        assert (x == y).all()


@pytest.mark.parametrize("mod", [2, 4, 6])
def test_csi_weighting8(orc, reflib, mod):
    rng = np.random.default_rng(50 + mod)
    for nsym in (2, 3, 5, 100, 101, 15000):
        e = aligned(mod * nsym, np.int8)
        e[:] = rng.integers(-128, 128, mod * nsym)
        csi = aligned(nsym + 8, np.float32)
        csi[:nsym] = rng.uniform(0.05, 2.0, nsym).astype(np.float32)
        a, b = e.copy(), e.copy()
        a2 = aligned(mod * nsym, np.int8); a2[:] = a
        assert reflib.ref_csi_b(mod, P(a2), P(csi), mod * nsym) == 0
        orc.orc_csi_b(mod, P(b), P(csi), mod * nsym)
        assert (a2 == b).all(), (mod, nsym, np.nonzero(a2 != b)[0][:8])
