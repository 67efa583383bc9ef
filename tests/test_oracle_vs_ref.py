"""Pins the C restatement (oracle/fec_oracle.c) against the UNMODIFIED reference objects (oracle/_ref).

Runs wherever oracle/_ref/libsrslte_ref.so exists (built from /root/reference by oracle/Makefile);
skipped otherwise — the golden-vector tests (test_oracle_golden.py) cover that case.
"""
import ctypes as C

import numpy as np
import pytest

from _libs import ALL_K, P, awgn_llr


def test_cbsegm_and_sizes(orc, reflib):
    for tbs in list(range(8, 2000, 8)) + list(range(2000, 100000, 104)) + [75376, 6120, 6096, 6144, 149776]:
        a = np.zeros(9, np.uint32)
        b = np.zeros(9, np.uint32)
        ra, rb = orc.orc_cbsegm(tbs, P(a)), reflib.ref_cbsegm(tbs, P(b))
        assert ra == rb
        if rb == 0:
            assert (a == b).all(), tbs
    for K in range(30, 6200, 7):
        assert orc.orc_cb_index(K) == reflib.ref_cb_index(K)
    for K in ALL_K:
        assert orc.orc_subblocks(K) == reflib.ref_subblocks(K)
        assert orc.orc_cb_size(orc.orc_cb_index(K)) == K


def test_crc(orc, reflib):
    rng = np.random.default_rng(3)
    for n in (8, 24, 40, 5000, 5824, 6144, 75376, 75400):
        d = rng.integers(0, 256, n // 8 + 2).astype(np.uint8)
        for t in (0, 1):
            assert orc.orc_crc(t, P(d), n) == reflib.ref_crc(t, P(d), n)


def test_qpp_tables_all_k(orc, reflib):
    for K in ALL_K:
        for Pn in (1, 8, 16):
            if Pn > 1 and K % Pn:
                continue
            f1, r1, f2, r2 = (np.zeros(K, np.uint16) for _ in range(4))
            orc.orc_qpp_sb(K, Pn, P(f1), P(r1))
            reflib.ref_interleaver(K, Pn, P(f2), P(r2))
            assert (f1 == f2).all() and (r1 == r2).all(), (K, Pn)


def test_rm_rx_all_k_rv(orc, reflib):
    """all 188 K x 4 rv x {natural, sb} x {no wrap, wrap}: same scope as the reference's rm_turbo_test"""
    rng = np.random.default_rng(5)
    for K in ALL_K:
        idx = orc.orc_cb_index(K)
        for rv in range(4):
            for sb in (0, 1):
                for E in (int(rng.integers(100, 3 * K + 12)), 3 * K + 12 + int(rng.integers(1, 2 * K))):
                    e = rng.integers(-3000, 3000, E).astype(np.int16)
                    s1 = rng.integers(-20000, 20000, 18600).astype(np.int16)
                    s2 = s1.copy()
                    assert orc.orc_rm_rx(P(e), E, K, rv, sb, P(s1)) == 0
                    assert reflib.ref_rm_rx(P(e), E, idx, rv, sb, P(s2)) == 0
                    assert (s1 == s2).all(), (K, rv, sb, E)


def _halfits(orc, reflib, K, inp, nof):
    ob = np.zeros((nof, K // 8), np.uint8)
    ol = np.zeros((nof, K), np.int16)
    rb, rl = np.zeros_like(ob), np.zeros_like(ol)
    assert orc.orc_tdec_halfits(K, P(inp), nof, P(ob), P(ol)) == 0
    assert reflib.ref_tdec_halfits(K, P(inp), nof, 0, P(rb), P(rl)) == 0
    assert (ob == rb).all(), K
    assert (ol == rl).all(), K


@pytest.mark.parametrize("K", [40, 48, 104, 200, 400, 408, 512, 800, 816, 1024, 2048, 5824, 6144])
def test_tdec_random_llrs(orc, reflib, K):
    """uniform random LLRs incl. amplitudes that drive the saturating paths to +-32767"""
    rng = np.random.default_rng(K)
    for amp in (600, 6000, 30000):
        _halfits(orc, reflib, K, rng.integers(-amp, amp + 1, 3 * K + 12).astype(np.int16), 8)


def test_tdec_all_k_awgn(orc, reflib):
    rng = np.random.default_rng(17)
    for K in ALL_K:
        bits = rng.integers(0, 2, K).astype(np.uint8)
        coded = np.zeros(3 * K + 12, np.uint8)
        reflib.ref_encode_cb(K, P(bits), P(coded))
        _halfits(orc, reflib, K, awgn_llr(coded, float(rng.choice([0.6, 0.8, 1.0, 1.2])), rng), 4)


def test_decode_cb_from_soft_buffer(orc, reflib):
    """sch.c-style CB decode from an sb-layout soft buffer with CRC24B early stop"""
    rng = np.random.default_rng(23)
    for K in (40, 400, 408, 800, 816, 2048, 5824, 6144):
        idx = orc.orc_cb_index(K)
        bits = rng.integers(0, 2, K).astype(np.uint8)
        by = np.packbits(bits)
        crc = reflib.ref_crc(1, P(by), K - 24)
        by[-3:] = [(crc >> 16) & 255, (crc >> 8) & 255, crc & 255]
        bits = np.unpackbits(by)
        coded = np.zeros(3 * K + 12, np.uint8)
        reflib.ref_encode_cb(K, P(bits), P(coded))
        for sigma in (0.5, 0.9, 1.4):
            e = awgn_llr(coded, sigma, rng)   # rate 1/3, rv0 order does not matter: feed through RM
            s_ref = np.zeros(18600, np.int16)
            reflib.ref_rm_rx(P(e), e.size, idx, 0, 1, P(s_ref))
            s_orc = s_ref.copy()
            o1, o2 = np.zeros(K // 8, np.uint8), np.zeros(K // 8, np.uint8)
            n1, n2 = C.c_int(), C.c_int()
            ok1 = orc.orc_decode_cb(P(s_orc), K, 8, 1, P(o1), C.byref(n1))
            ok2 = reflib.ref_decode_cb(P(s_ref), K, 8, 1, P(o2), C.byref(n2))
            assert (ok1, n1.value) == (ok2, n2.value), (K, sigma)
            assert (o1 == o2).all()


def _tb_case(orc, reflib, tbs, Qm, nbits, sigmas, rvs, max_its, seed):
    rng = np.random.default_rng(seed)
    rt = C.c_void_p(reflib.ref_tb_new(100))
    ot = C.c_void_p(orc.orc_tb_new())
    data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
    data[: tbs // 8] = rng.integers(0, 256, tbs // 8)
    reflib.ref_tb_reset(rt, tbs)
    orc.orc_tb_reset(ot, tbs)
    seg = np.zeros(9, np.uint32)
    orc.orc_cbsegm(tbs, P(seg))
    Cn = int(seg[1])
    for rv, sigma in zip(rvs, sigmas):
        e = np.zeros(nbits, np.uint8)
        assert reflib.ref_tb_encode(rt, tbs, Qm, rv, nbits, P(data), P(e)) == 0
        llr = awgn_llr(e, sigma, rng)
        d1, d2, d3 = (np.zeros(tbs // 8 + 3 + 1024, np.uint8) for _ in range(3))
        c1, c2 = np.zeros(13, np.uint8), np.zeros(13, np.uint8)
        i1, i2 = np.zeros(13, np.uint32), np.zeros(13, np.uint32)
        r1 = orc.orc_tb_decode(ot, tbs, Qm, rv, nbits, P(llr), max_its, P(d1), P(c1), P(i1))
        r2 = reflib.ref_tb_decode_detail(rt, tbs, Qm, rv, nbits, P(llr), max_its, P(d2), P(c2), P(i2))
        assert r1 == r2
        assert (c1[:Cn] == c2[:Cn]).all() and (i1[:Cn] == i2[:Cn]).all(), (i1, i2)
        assert (d1[: tbs // 8 + 3] == d2[: tbs // 8 + 3]).all()
        for cb in range(Cn):   # soft buffers identical at every LUT-addressed position (and, here, everywhere)
            s1, s2 = np.zeros(18600, np.int16), np.zeros(18600, np.int16)
            orc.orc_tb_get_soft(ot, cb, P(s1))
            reflib.ref_tb_get_soft(rt, cb, P(s2))
            K = int(seg[2])
            if orc.orc_subblocks(K):   # trap 7: the reference copies tails into the pad gaps; ignore pads
                for base in (0, K + 32, 2 * (K + 32)):
                    assert (s1[base:base + K] == s2[base:base + K]).all()
                assert (s1[3 * (K + 32):3 * (K + 32) + 12] == s2[3 * (K + 32):3 * (K + 32) + 12]).all()
            else:
                assert (s1[:3 * K + 12] == s2[:3 * K + 12]).all()
        if r2 == 0:
            assert (d2[: tbs // 8] == data[: tbs // 8]).all()
            break   # HARQ ends on ACK (decoding again without a reset would copy the reference's uninitialised CB cache)
    reflib.ref_tb_free(rt)
    orc.orc_tb_free(ot)


def test_tb_decode_mcs28_harq(orc, reflib):
    """config 2 shape (TBS 75376, 13 x K=5824, E quirk) with HARQ retransmissions rv 0,2,3,1"""
    _tb_case(orc, reflib, 75376, 6, 90000, [0.62, 0.62, 0.62, 0.62], [0, 2, 3, 1], 8, 1)
    _tb_case(orc, reflib, 75376, 6, 90000, [0.45], [0], 8, 2)
    _tb_case(orc, reflib, 75376, 6, 86400, [0.42, 0.5], [0, 2], 8, 3)   # TM4 2-codeword shape


def test_tb_decode_small_and_single_cb(orc, reflib):
    _tb_case(orc, reflib, 1000, 2, 2400, [0.7, 0.7], [0, 1], 8, 4)      # C=1, CRC24A early stop, K=1024
    _tb_case(orc, reflib, 16, 2, 288, [0.9], [0], 10, 5)                # K=40, P=0
    _tb_case(orc, reflib, 6200, 4, 9600, [0.75, 0.75], [0, 3], 8, 6)    # C=2
    _tb_case(orc, reflib, 12216, 2, 14400, [0.95, 0.9, 0.9], [0, 2, 3], 8, 7)   # low rate + wrap (E>3K+12) after combining


def test_tb_encode_vs_ref(orc, reflib):
    """transmit side: orc_tb_encode == srslte_dlsch_encode for one- and many-code-block transport blocks,
    all redundancy versions, E not a multiple of 8 (code blocks share bytes of the packed output)"""
    rng = np.random.default_rng(11)
    reflib.ref_tb_new.restype = C.c_void_p
    h = C.c_void_p(reflib.ref_tb_new(100))
    for tbs, Qm, nbits in ((40, 2, 288), (504, 2, 1200), (6120, 4, 9600), (6200, 6, 14400), (12960, 2, 20000), (75376, 6, 90000), (75376, 6, 86400),
                           (36696, 4, 57600), (61664, 6, 79998)):
        for rv in range(4):
            data = np.zeros(tbs // 8 + 8, np.uint8)
            data[: tbs // 8] = rng.integers(0, 256, tbs // 8)
            d2 = data.copy()
            want = np.zeros(nbits + 16, np.uint8)
            # the reference keeps the circular buffer of rv 0 in its tx soft buffer: give it the rv-0 call first
            assert reflib.ref_tb_encode(h, tbs, Qm, 0, nbits, P(d2), P(want)) == 0
            if rv:
                assert reflib.ref_tb_encode(h, tbs, Qm, rv, nbits, P(d2), P(want)) == 0
            got = np.zeros((nbits + 7) // 8, np.uint8)
            assert orc.orc_tb_encode(tbs, Qm, rv, nbits, P(data), P(got)) == 0
            assert (np.unpackbits(got)[:nbits] == want[:nbits]).all(), (tbs, Qm, nbits, rv)
    reflib.ref_tb_free(h)
