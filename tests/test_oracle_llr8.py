"""8-bit LLR path on the CPU: the oracle's restatement of srslte_rm_turbo_rx_lut_8bit, the int8 windowed
decoders (sse8 / avx8) with their iteration driver and the 16-bit fallbacks, and decode_tb with llr_is_8bit, against the
golden outputs of the unmodified reference (tests/golden/llr8.npz) and, where oracle/_ref exists, the reference itself."""
import numpy as np
import pytest

import _golden as G
from _libs import ALL_K, P, aligned


def test_subblocks8_rule(orc):
    """turbodecoder.c:392-406 (AVX2 build)"""
    for K in ALL_K:
        want = 32 if (K % 32 == 0 and K > 2048) else 16 if (K % 16 == 0 and K > 800) else 8 if K > 400 else 0
        assert orc.orc_subblocks8(K) == want


@pytest.mark.parametrize("c", list(G.tdec8_cases()), ids=lambda c: c["id"])
def test_tdec8_golden(orc, c):
    K, nof = c["K"], c["nof"]
    ob = np.zeros((nof, K // 8), np.uint8)
    ol = np.zeros((nof, K), np.int16)
    assert orc.orc_tdec_halfits8(K, P(c["inp"]), nof, P(ob), P(ol)) == 0
    assert (ob == c["bytes"]).all()
    assert [G.crc32(ol[n]) for n in range(nof)] == [int(v) for v in c["llrcrc"]]


@pytest.mark.parametrize("c", list(G.tb8_cases()), ids=lambda c: c["id"])
def test_tb8_golden(orc, c):
    """return code, payload, per-CB CRC flags, average half-iterations and the int8 soft buffers after every transmission"""
    tbs = c["tbs"]
    seg = np.zeros(9, np.uint32)
    orc.orc_cbsegm(tbs, P(seg))
    C_, K1, K2, C1 = int(seg[1]), int(seg[2]), int(seg[3]), int(seg[6])
    t = orc.orc_tb_new()
    orc.orc_tb_reset(t, tbs)
    try:
        for tx in c["tx"]:
            d = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
            crc, its = np.zeros(16, np.uint8), np.zeros(16, np.uint32)
            ret = orc.orc_tb_decode8(t, tbs, c["Qm"], tx["rv"], c["nbits"], P(tx["llr"]), c["max_its"], P(d), P(crc), P(its))
            assert ret == tx["ret"]
            assert (crc[:C_] == tx["crc"][:C_]).all()
            assert int(its[:C_].sum()) // C_ == tx["avg_its"]                 # sch.c:424: nof_iterations /= C
            if ret == 0:
                assert (d[: tbs // 8] == c["payload"]).all() and (d[: tbs // 8 + 3] == tx["data"]).all()
            for cb in range(C_):
                K = K1 if cb < C1 else K2
                s = np.zeros(18600, np.int16)
                orc.orc_tb_get_soft(t, cb, P(s))
                assert G.crc32(G.soft8_view(s.view(np.int8), K, orc.orc_subblocks8(K))) == int(tx["softcrc"][cb]), (cb, K)
    finally:
        orc.orc_tb_free(t)


def test_tdec8_every_k_vs_ref(orc, reflib):
    """every code-block size of the int8 decoders (K > 800) and a sample of the fallbacks, saturating and benign inputs:
    decisions AND the LLR arrays of 4 half-iterations"""
    ref = reflib
    rng = np.random.default_rng(8)
    for K in ALL_K:
        if K <= 800 and K not in (40, 200, 400, 408, 608, 800):
            continue
        n = 3 * K + 12
        bits = rng.integers(0, 2, n)
        for inp in (rng.integers(-128, 128, n).astype(np.int8), rng.integers(-30, 31, n).astype(np.int8),
                    np.clip(np.rint(20 * (2 * bits - 1) + 12 * rng.standard_normal(n)), -128, 127).astype(np.int8),
                    np.clip(np.rint(60 * (2 * bits - 1) + 50 * rng.standard_normal(n)), -128, 127).astype(np.int8)):
            ob, ol = np.zeros((4, K // 8), np.uint8), np.zeros((4, K), np.int16)
            rb, rl = np.zeros((4, K // 8), np.uint8), np.zeros((4, K), np.int16)
            assert orc.orc_tdec_halfits8(K, P(inp), 4, P(ob), P(ol)) == 0
            assert ref.ref_tdec_halfits8(K, P(inp), 4, P(rb), P(rl)) == 0
            assert (ob == rb).all() and (ol == rl).all(), K


def test_rm_rx8_vs_ref(orc, reflib):
    """wrapping int8 accumulation into a non-zero soft buffer, E below / at / above 3K+12 (wrap code of rm_turbo.c:543-631)"""
    ref = reflib
    rng = np.random.default_rng(9)
    for K in (40, 408, 816, 1024, 2048, 2112, 5824, 6144):
        idx = orc.orc_cb_index(K)
        for rv in range(4):
            for E in (100, K + 5, 3 * K + 12, 3 * K + 500, 2 * (3 * K + 12) + 77):
                e = rng.integers(-128, 128, E).astype(np.int8)
                s0 = rng.integers(-128, 128, 2 * 18600).astype(np.int8)
                a, b = s0.copy(), s0.copy()
                assert orc.orc_rm_rx8(P(e), E, K, rv, 1, P(a)) == 0
                assert ref.ref_rm_rx8(P(e), E, idx, rv, P(b)) == 0
                assert (a == b).all(), (K, rv, E)


def test_tb8_vs_ref(orc, reflib):
    """srslte_dlsch_decode with llr_is_8bit over HARQ sequences: return code, payload, CRC flags, average iterations"""
    ref = reflib
    rng = np.random.default_rng(10)
    for tbs, Qm, nb, sig in ((75376, 6, 90000, 0.6), (61664, 6, 79998, 0.55), (6200, 6, 14400, 0.75), (2024, 2, 4000, 0.95),
                             (256, 2, 600, 0.8), (12960, 2, 20000, 0.85)):
        data = rng.integers(0, 256, tbs // 8).astype(np.uint8)
        to, tr = orc.orc_tb_new(), ref.ref_tb_new(100)
        orc.orc_tb_reset(to, tbs)
        ref.ref_tb_reset(tr, tbs)
        seg = np.zeros(9, np.uint32)
        orc.orc_cbsegm(tbs, P(seg))
        C_ = int(seg[1])
        for rv in (0, 2, 3, 1):
            eb = np.zeros((nb + 7) // 8, np.uint8)
            assert orc.orc_tb_encode(tbs, Qm, rv, nb, P(data), P(eb)) == 0
            bits = np.unpackbits(eb)[:nb].astype(np.float64)
            llr = aligned(nb + 64, np.int8)
            llr[:nb] = np.clip(np.rint(24 * ((2 * bits - 1) + sig * rng.standard_normal(nb))), -128, 127)
            do, dr = np.zeros(tbs // 8 + 3 + 1024, np.uint8), np.zeros(tbs // 8 + 3 + 1024, np.uint8)
            co, cr, io, ia = np.zeros(16, np.uint8), np.zeros(16, np.uint8), np.zeros(16, np.uint32), np.zeros(1, np.uint32)
            ro = orc.orc_tb_decode8(to, tbs, Qm, rv, nb, P(llr), 8, P(do), P(co), P(io))
            rr = ref.ref_tb_decode8(tr, tbs, Qm, rv, nb, P(llr), 8, P(dr), P(cr), P(ia))
            assert ro == rr and (co[:C_] == cr[:C_]).all() and int(io[:C_].sum()) // C_ == int(ia[0]), (tbs, rv)
            if ro == 0:
                assert (do[: tbs // 8 + 3] == dr[: tbs // 8 + 3]).all() and (do[: tbs // 8] == data).all()
                break
        orc.orc_tb_free(to)
        ref.ref_tb_free(tr)
