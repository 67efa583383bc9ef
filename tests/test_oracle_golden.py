"""The C restatement against the committed golden vectors (generated from the unmodified reference by
oracle/gen_golden.py). Runs anywhere gcc exists; no GPU, no /root/reference."""
import ctypes as C

import numpy as np
import pytest

import _golden as G
from _libs import ALL_K, P, int_gauss


def test_golden_tables(orc):
    z = G.tables()
    for K, Pn, cf, cr in z["qpp"]:
        f, r = np.zeros(int(K), np.uint16), np.zeros(int(K), np.uint16)
        assert orc.orc_qpp_sb(int(K), int(Pn), P(f), P(r)) == 0
        assert (G.crc32(f), G.crc32(r)) == (int(cf), int(cr)), (K, Pn)
    for n, a, b in z["crc"]:
        d = (int_gauss(int(n) // 8, int(n)) & 255).astype(np.uint8)
        assert orc.orc_crc(0, P(d), int(n)) == int(a) and orc.orc_crc(1, P(d), int(n)) == int(b)
    for row in z["cbsegm"]:
        o = np.zeros(9, np.uint32)
        assert orc.orc_cbsegm(int(row[0]), P(o)) == int(row[1])
        if row[1] == 0:
            assert [int(x) for x in o] == [int(x) for x in row[2:]]


def test_golden_rm_all_k_rv(orc):
    z = G.tables()
    for K, rv, sb, E, c in z["rm"]:
        K, rv, sb, E = int(K), int(rv), int(sb), int(E)
        e = (int_gauss(E, K * 8 + rv) % 2001 - 1000).astype(np.int16)
        s = (int_gauss(18600, K) % 20001 - 10000).astype(np.int16)
        assert orc.orc_rm_rx(P(e), E, K, rv, sb, P(s)) == 0
        assert G.crc32(s) == int(c), (K, rv, sb, E)


def test_crc_known_answers_reference_crc_test(orc):
    """crc_test.h:40-48: 5001 rand()%2 bits after srand(1) -> CRC24A 0x1C5C97, CRC24B 0x36D1F0 (glibc rand).
    The path only ever checks whole bytes, so the KAT is applied to the bit-serial definition the
    byte routine restates: 5001 bits followed by nothing, MSB first."""
    libc = C.CDLL("libc.so.6")
    libc.srand(1)
    bits = np.array([libc.rand() % 2 for _ in range(5001)], np.uint8)
    for t, poly, want in ((0, 0x1864CFB, 0x1C5C97), (1, 0x1800063, 0x36D1F0)):
        r = 0
        for b in list(bits) + [0] * 24:
            r = (r << 1) | int(b)
            if r & 0x1000000:
                r ^= poly
        assert r == want
        # whole-byte prefix must agree with the oracle's byte routine
        by = np.packbits(bits[:5000])
        r = 0
        for b in list(bits[:5000]) + [0] * 24:
            r = (r << 1) | int(b)
            if r & 0x1000000:
                r ^= poly
        assert orc.orc_crc(t, P(by), 5000) == r


@pytest.mark.parametrize("case", list(G.tdec_cases()), ids=lambda c: c["id"])
def test_golden_tdec(orc, case):
    K, nof = case["K"], case["nof"]
    ob, ol = np.zeros((nof, K // 8), np.uint8), np.zeros((nof, K), np.int16)
    assert orc.orc_tdec_halfits(K, P(case["inp"]), nof, P(ob), P(ol)) == 0
    assert (ob == case["bytes"]).all()
    assert [G.crc32(ol[n]) for n in range(nof)] == [int(x) for x in case["llrcrc"]]
    if case["llr"] is not None:
        assert (ol == case["llr"]).all()


@pytest.mark.parametrize("case", list(G.tb_cases()), ids=lambda c: c["id"])
def test_golden_tb(orc, case):
    tbs = case["tbs"]
    t = C.c_void_p(orc.orc_tb_new())
    orc.orc_tb_reset(t, tbs)
    seg = np.zeros(9, np.uint32)
    orc.orc_cbsegm(tbs, P(seg))
    Cn = int(seg[1])
    for tx in case["tx"]:
        d = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
        crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
        ret = orc.orc_tb_decode(t, tbs, case["Qm"], tx["rv"], case["nbits"], P(tx["llr"]), case["max_its"], P(d), P(crc), P(its))
        assert ret == tx["ret"]
        assert (crc[:Cn] == tx["crc"][:Cn]).all() and (its[:Cn] == tx["its"][:Cn]).all()
        assert (d[: tbs // 8 + 3] == tx["data"]).all()
        for cb in range(Cn):
            s = np.zeros(18600, np.int16)
            orc.orc_tb_get_soft(t, cb, P(s))
            K = int(seg[2]) if cb < int(seg[6]) else int(seg[3])
            assert G.soft_crc_ref_layout(s, K, orc.orc_subblocks(K)) == int(tx["softcrc"][cb])
        if ret == 0:
            assert (d[: tbs // 8] == case["tbdata"]).all()
    orc.orc_tb_free(t)


@pytest.mark.parametrize("case", list(G.fe_dl_cases()), ids=lambda c: c["id"])
def test_golden_frontend_dl(orc, case):
    """demodulate + descramble (+ CSI weights): LLRs of the reference's pdsch.c:795-813"""
    import _fe
    e, _ = _fe.oracle_frontend(orc, case["sym"], case["mod"], case["c_init"], case["csi"])
    assert G.crc32(e) == case["crc"]
    if case["e"] is not None:
        assert (e == case["e"]).all()


@pytest.mark.parametrize("case", list(G.fe_ul_cases()), ids=lambda c: c["id"])
def test_golden_frontend_ul(orc, case):
    """pusch.c:617-657: HARQ-ACK positions zeroed, descrambling, channel de-interleaver around the RI positions"""
    import _fe
    g, _ = _fe.oracle_frontend(orc, case["sym"], case["mod"], case["c_init"], None, case["cols"], case["ack_pos"], case["ri_pos"])
    ng = g.size - len(case["ri_pos"])
    assert (g[:64] == case["g0"]).all()
    assert G.crc32(g[:ng]) == case["crc_g"]


def test_tb_encode_golden(orc):
    """transmit side: the oracle's encoder reproduces the bits srslte_dlsch_encode emitted for every
    (re)transmission of the golden transport blocks (sch.c:188-296)"""
    import os
    import _golden as G
    z = np.load(os.path.join(G.GOLD, "tb.npz"))
    n = 0
    for ci, c in enumerate(G.tb_cases()):
        data = np.ascontiguousarray(c["tbdata"][: c["tbs"] // 8])
        for tx, t in enumerate(c["tx"]):
            want = z[f"e_{ci}_{tx}"]
            got = np.zeros((c["nbits"] + 7) // 8, np.uint8)
            assert orc.orc_tb_encode(c["tbs"], c["Qm"], t["rv"], c["nbits"], P(data), P(got)) == 0
            assert (np.unpackbits(got)[: c["nbits"]] == np.unpackbits(want)[: c["nbits"]]).all(), (c["id"], tx)
            n += 1
    assert n >= 4


def test_oracle_bler_point_matches_reference_auto_decoder(orc):
    """one point of tests/golden/bler_gen.npz (K = 504, the size where the windowed decoder departs most from the generic
    one): the oracle's block / bit error counts on the seeded vectors equal those of the reference's AUTO decoder"""
    case = [c for c in G.bler_cases() if c["K"] == 504 and c["sigma_milli"] == 1050][0]
    K = case["K"]
    bits, llr = G.bler_vectors(K, case["sigma_milli"], case["n"], case["seed"])
    assert G.crc32(llr) == case["llrcrc"]
    ob = np.zeros((case["nof"], K // 8), np.uint8)
    blk = bit = 0
    for i in range(case["n"]):
        assert orc.orc_tdec_halfits(K, P(np.ascontiguousarray(llr[i])), case["nof"], P(ob), None) == 0
        e = np.unpackbits(ob[-1]) != bits[i]
        blk += int(e.any()); bit += int(e.sum())
    assert (blk, bit) == case["auto"]
