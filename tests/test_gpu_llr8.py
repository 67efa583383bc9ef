"""8-bit LLR path on the GPU: fec_cuda_set_llr_8bit + the decode entry points against the golden outputs
of the unmodified reference (tests/golden/llr8.npz), the oracle (pinned to the reference in test_oracle_llr8.py) and, at the
bench's full shape, through the encode -> int8 LLR -> decode round trip."""
import importlib

import numpy as np
import pytest

import _golden as G
from _libs import ALL_K, P

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fec8():
    pkg = importlib.import_module("empower-srslte_b200")
    f = pkg.FecCuda(0)
    f.set_llr_8bit(True)
    yield f
    f.close()


def test_tdec8_golden(fec8):
    """decisions after every half-iteration count of the golden snapshots (all four decoder classes)"""
    cases = list(G.tdec8_cases())
    for nof in range(1, 9):
        sel = [c for c in cases if c["nof"] >= nof]
        res = fec8.decode_cbs([c["K"] for c in sel], [c["inp"].astype(np.int16) for c in sel], nof, early_stop=False)
        bad = [(c["id"], nof) for c, r in zip(sel, res) if not (r["out"] == c["bytes"][nof - 1]).all()]
        assert not bad, bad


def test_tdec8_every_k_vs_oracle(fec8, orc):
    """every code-block size: saturating (full-range) and demodulator-like int8 inputs, 1..4 half-iterations, one batch per count"""
    rng = np.random.default_rng(41)
    Ks, llrs, want = [], [], []
    for K in ALL_K:
        n = 3 * K + 12
        bits = rng.integers(0, 2, n)
        for inp in (rng.integers(-128, 128, n).astype(np.int8),
                    np.clip(np.rint(24 * (2 * bits - 1) + 16 * rng.standard_normal(n)), -128, 127).astype(np.int8)):
            ob = np.zeros((4, K // 8), np.uint8)
            assert orc.orc_tdec_halfits8(K, P(inp), 4, P(ob), None) == 0
            Ks.append(K); llrs.append(inp.astype(np.int16)); want.append(ob)
    for nof in (1, 2, 3, 4):
        res = fec8.decode_cbs(Ks, llrs, nof, early_stop=False)
        bad = [(K, nof) for K, r, w in zip(Ks, res, want) if not (r["out"] == w[nof - 1]).all()]
        assert not bad, bad[:10]


def test_tdec8_early_stop_vs_oracle(fec8, orc):
    """noisy codewords with CRC-24B early stop: iteration counts and payloads, int8 classes 16 and 32"""
    synth = importlib.import_module("empower-srslte_b200.synth")
    rng = np.random.default_rng(42)
    Ks, llrs, exp = [], [], []
    for K in (816, 1024, 1056, 2048, 2112, 3136, 5824, 6144):
        for sigma in (0.3, 0.45, 0.6):
            body = rng.integers(0, 256, (1, K // 8 - 3), dtype=np.uint8)
            blk = np.concatenate([body, synth._crc_bytes(synth.crc24(synth.CRC24B, body))], axis=1)
            coded = synth.turbo_encode(np.unpackbits(blk, axis=1))[0]
            x = np.clip(np.rint(24 * ((2.0 * coded - 1.0) + sigma * rng.standard_normal(coded.shape))), -128, 127).astype(np.int8)
            ob = np.zeros((10, K // 8), np.uint8)
            assert orc.orc_tdec_halfits8(K, P(x), 10, P(ob), None) == 0
            ok = [n + 1 for n in range(10) if orc.orc_crc(1, P(np.ascontiguousarray(ob[n])), K) == 0]
            n = ok[0] if ok else 10
            Ks.append(K); llrs.append(x.astype(np.int16)); exp.append((n, ob[n - 1].copy(), bool(ok)))
    res = fec8.decode_cbs(Ks, llrs, 10, crc_type=1, early_stop=True)
    for K, r, (n, ob, ok) in zip(Ks, res, exp):
        assert r["its"] == n and bool(r["crc_ok"]) == ok and (r["out"] == ob).all(), (K, r["its"], n)
    assert sum(1 for e in exp if e[2]) >= len(exp) // 2


@pytest.mark.parametrize("c", list(G.tb8_cases()), ids=lambda c: c["id"])
def test_tb8_golden(fec8, orc, c):
    """HARQ sequences of srslte_dlsch_decode with llr_is_8bit: return code, payload, CRC flags, average iterations; the
    int8 soft buffers after every transmission against the oracle's (whose checksums are pinned to the reference's)"""
    pkg = importlib.import_module("empower-srslte_b200")
    tbs = c["tbs"]
    seg = np.zeros(9, np.uint32)
    orc.orc_cbsegm(tbs, P(seg))
    C_, K1, K2, C1 = int(seg[1]), int(seg[2]), int(seg[3]), int(seg[6])
    sb = fec8.softbuffer_alloc(1)
    to = orc.orc_tb_new()
    orc.orc_tb_reset(to, tbs)
    try:
        for ti, tx in enumerate(c["tx"]):
            out = np.zeros(tbs // 8 + 3 + 768, np.uint8)
            d = (pkg.TbDesc * 1)()
            d[0].tbs, d[0].Qm, d[0].rv, d[0].nof_e_bits, d[0].max_its = tbs, c["Qm"], tx["rv"], c["nbits"], c["max_its"]
            d[0].softbuffer, d[0].new_data = sb, 1 if ti == 0 else 0
            llr = np.ascontiguousarray(tx["llr"])
            d[0].e_bits, d[0].data = llr.ctypes.data, out.ctypes.data
            fec8.decode_tbs(d)
            assert d[0].ret == tx["ret"]
            assert [int(d[0].cb_crc[i]) for i in range(C_)] == [int(v) for v in tx["crc"][:C_]]
            assert int(d[0].avg_its) == tx["avg_its"]
            if tx["ret"] == 0:
                assert (out[: tbs // 8 + 3] == tx["data"]).all() and (out[: tbs // 8] == c["payload"]).all()
            do = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
            co, io = np.zeros(16, np.uint8), np.zeros(16, np.uint32)
            orc.orc_tb_decode8(to, tbs, c["Qm"], tx["rv"], c["nbits"], P(llr), c["max_its"], P(do), P(co), P(io))
            for cb in range(C_):
                K = K1 if cb < C1 else K2
                P8 = orc.orc_subblocks8(K)
                s = np.zeros(18600, np.int16)
                orc.orc_tb_get_soft(to, cb, P(s))
                want = G.soft8_view(s.view(np.int8), K, P8).astype(np.int16)
                got = fec8.softbuffer_read(sb, cb, K)
                assert G.crc32(want.astype(np.int8)) == int(tx["softcrc"][cb])
                assert (G.soft8_view(got, K, P8) == want).all(), (cb, K)
    finally:
        orc.orc_tb_free(to)


def test_rm_rx8_all_k_vs_oracle(fec8, orc):
    """the int8 rate de-matcher (srslte_rm_turbo_rx_lut_8bit, rm_turbo.c:432-469) for every one of the 188 sizes x 4 rv,
    lengths below and above 3K+12, accumulating with int8 wrap-around into a non-empty buffer; checker: the oracle, which
    tests/test_oracle_llr8.py pins to the compiled reference on the same grid"""
    rng = np.random.default_rng(88)
    bad = []
    for K in ALL_K:
        P8 = orc.orc_subblocks8(K)
        for rv in range(4):
            E = int(rng.choice([K + 7, 3 * K + 12, 3 * K + 12 + 333, 2 * (3 * K + 12) + 5]))
            e = rng.integers(-128, 128, E).astype(np.int8)
            s0 = np.zeros(18600, np.int8)
            s0[: 3 * K + 108] = rng.integers(-128, 128, 3 * K + 108)
            got, want = s0.copy(), np.zeros(2 * 18600, np.int8)
            want[:18600] = s0
            assert fec8.lib.fec_cuda_rm_rx(fec8.h, P(e), E, ALL_K.index(K), rv, 1, P(got)) == 0
            assert orc.orc_rm_rx8(P(e), E, K, rv, 1, P(want)) == 0
            if not (G.soft8_view(got, K, P8) == G.soft8_view(want, K, P8)).all():
                bad.append((K, rv, E))
    assert not bad, bad[:20]


def test_tb8_round_trip_full_shape(fec8, orc):
    """BASELINE configs[1] shape, 296 transport blocks: encode on the GPU, int8 LLRs with noise, decode in 8-bit mode: every
    payload comes back at low noise; at high noise a retransmission combines into the int8 soft buffers and rescues blocks;
    return codes and per-code-block half-iterations of a sample equal the oracle's over the same two transmissions"""
    pkg = importlib.import_module("empower-srslte_b200")
    TBS, QM, NB, n = 75376, 6, 90000, 296
    rng = np.random.default_rng(43)
    data = rng.integers(0, 256, (n, TBS // 8)).astype(np.uint8)
    sb = fec8.softbuffer_alloc(n)
    out = np.zeros((n, TBS // 8 + 3 + 768), np.uint8)

    def transmit(rv, sigma, new):
        e = np.zeros((n, NB // 8), np.uint8)
        d = (pkg.TxTbDesc * n)()
        for i in range(n):
            d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits = TBS, QM, rv, NB
            d[i].data, d[i].e_bits = data[i].ctypes.data, e[i].ctypes.data
        fec8.encode_tbs(d)
        bits = np.unpackbits(e, axis=1)[:, :NB].astype(np.float32)
        llr = np.clip(np.rint(24 * ((2 * bits - 1) + sigma * rng.standard_normal(bits.shape, dtype=np.float32))), -128, 127).astype(np.int8)
        r = (pkg.TbDesc * n)()
        for i in range(n):
            r[i].tbs, r[i].Qm, r[i].rv, r[i].nof_e_bits, r[i].max_its = TBS, QM, rv, NB, 8
            r[i].softbuffer, r[i].new_data = sb + i, new
            r[i].e_bits, r[i].data = llr[i].ctypes.data, out[i].ctypes.data
        fec8.decode_tbs(r)
        return r, llr

    r, _ = transmit(0, 0.2, 1)        # the int8 decoders have an error floor: at sigma 0.3 the reference itself loses ~1 block in 2000
    assert all(r[i].ret == 0 for i in range(n))
    assert (out[:, : TBS // 8] == data).all()
    r1, llr1 = transmit(0, 0.6, 1)                    # too noisy for one transmission
    ok1 = [r1[i].ret == 0 for i in range(n)]
    its1 = [[int(r1[i].cb_its[k]) for k in range(13)] for i in range(6)]
    r2, llr2 = transmit(2, 0.6, 0)                    # the retransmission combines
    ok2 = [r2[i].ret == 0 for i in range(n)]
    assert sum(ok1) < n // 2 and sum(ok2) > sum(ok1) and sum(ok2) > n // 2, (sum(ok1), sum(ok2))
    assert all((out[i, : TBS // 8] == data[i]).all() for i in range(n) if ok2[i])
    for i in range(6):
        t = orc.orc_tb_new()
        orc.orc_tb_reset(t, TBS)
        do, co, io = np.zeros(TBS // 8 + 3 + 1024, np.uint8), np.zeros(16, np.uint8), np.zeros(16, np.uint32)
        ra = orc.orc_tb_decode8(t, TBS, QM, 0, NB, P(np.ascontiguousarray(llr1[i])), 8, P(do), P(co), P(io))
        assert (ra == 0) == ok1[i] and [int(v) for v in io[:13]] == its1[i], i
        rb = orc.orc_tb_decode8(t, TBS, QM, 2, NB, P(np.ascontiguousarray(llr2[i])), 8, P(do), P(co), P(io))
        assert (rb == 0) == ok2[i] and [int(v) for v in io[:13]] == [int(r2[i].cb_its[k]) for k in range(13)], i
        orc.orc_tb_free(t)


def test_mode_switch_leaves_16bit_path_untouched(orc):
    """the same context decodes a 16-bit code block identically before and after a spell in 8-bit mode"""
    pkg = importlib.import_module("empower-srslte_b200")
    f = pkg.FecCuda(0)
    try:
        rng = np.random.default_rng(44)
        K = 6144
        inp = rng.integers(-300, 301, 3 * K + 12).astype(np.int16)
        ob = np.zeros((3, K // 8), np.uint8)
        assert orc.orc_tdec_halfits(K, P(inp), 3, P(ob), None) == 0
        a = f.decode_cbs([K], [inp], 3, early_stop=False)[0]["out"].copy()
        f.set_llr_8bit(True)
        i8 = rng.integers(-128, 128, 3 * K + 12).astype(np.int8)
        o8 = np.zeros((3, K // 8), np.uint8)
        assert orc.orc_tdec_halfits8(K, P(i8), 3, P(o8), None) == 0
        assert (f.decode_cbs([K], [i8.astype(np.int16)], 3, early_stop=False)[0]["out"] == o8[2]).all()
        f.set_llr_8bit(False)
        b = f.decode_cbs([K], [inp], 3, early_stop=False)[0]["out"]
        assert (a == ob[2]).all() and (b == ob[2]).all()
    finally:
        f.close()


def test_invalid_in_8bit_mode(fec8):
    """LLRs wider than int8 through the int16 code-block entry point, and the symbol front end, are refused in 8-bit mode"""
    K = 1024
    bad = np.zeros(3 * K + 12, np.int16)
    bad[5] = 200
    with pytest.raises(RuntimeError):
        fec8.decode_cbs([K], [bad], 2, early_stop=False)
    ok = fec8.decode_cbs([K], [np.zeros(3 * K + 12, np.int16)], 1, early_stop=False)
    assert not ok[0]["out"].any()
