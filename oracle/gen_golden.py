#!/usr/bin/env python
"""Generate tests/golden/*.npz from the UNMODIFIED reference objects (oracle/_ref/libsrslte_ref.so).

Run in the container where /root/reference is mounted:   python oracle/gen_golden.py
Inputs are derived from integer-only generators in tests/_libs.py (bit-identical on every box), so the
fixtures only need to hold the reference's OUTPUTS (plus encoder bits, which need the reference encoder).
"""
import ctypes as C
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _libs import ALL_K, P, aligned, int_awgn_llr, int_gauss, ref  # noqa: E402

R = ref()
assert R is not None, "needs oracle/_ref (make -C oracle ref)"
OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def crc32(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def gen_tdec():
    """srslte_tdec_iteration snapshots on natural-layout inputs (turbodecoder_test path, AUTO impl)."""
    out = {}
    cases = []
    for K, kind, par, nof in [(40, "awgn", 900, 6), (104, "rand", 600, 8), (400, "awgn", 800, 8), (408, "awgn", 700, 8),
                              (512, "rand", 6000, 8), (800, "awgn", 900, 8), (816, "awgn", 800, 8), (1024, "rand", 30000, 8),
                              (2048, "awgn", 700, 8), (5824, "awgn", 620, 8), (6144, "awgn", 1540, 4), (6144, "awgn", 800, 8),
                              (6144, "rand", 30000, 8)]:
        seed = 1000 + len(cases)
        if kind == "awgn":
            bits = (int_gauss(K, seed) & 1).astype(np.uint8)
            coded = np.zeros(3 * K + 12, np.uint8)
            R.ref_encode_cb(K, P(bits), P(coded))
            inp = int_awgn_llr(coded, par, seed + 1)
            out[f"coded_{len(cases)}"] = np.packbits(coded)
        else:
            inp = (int_gauss(3 * K + 12, seed) % (2 * par + 1) - par).astype(np.int16)
        ob = np.zeros((nof, K // 8), np.uint8)
        ol = np.zeros((nof, K), np.int16)
        assert R.ref_tdec_halfits(K, P(inp), nof, 0, P(ob), P(ol)) == 0
        out[f"bytes_{len(cases)}"] = ob
        out[f"llrcrc_{len(cases)}"] = np.array([crc32(ol[n]) for n in range(nof)], np.uint32)
        if K <= 1024:
            out[f"llr_{len(cases)}"] = ol
        out[f"incrc_{len(cases)}"] = np.array([crc32(inp)], np.uint32)
        cases.append((K, 0 if kind == "awgn" else 1, par, nof, seed))
    out["cases"] = np.array(cases, np.int64)
    np.savez_compressed(os.path.join(OUT, "tdec.npz"), **out)


def gen_tables():
    """CRC KATs, QPP and RM tables as checksums over all 188 K (x 4 rv)."""
    out = {}
    qpp = []
    for K in ALL_K:
        for Pn in (1, 8, 16):
            if Pn > 1 and K % Pn:
                continue
            f, r = np.zeros(K, np.uint16), np.zeros(K, np.uint16)
            R.ref_interleaver(K, Pn, P(f), P(r))
            qpp.append((K, Pn, crc32(f), crc32(r)))
    out["qpp"] = np.array(qpp, np.int64)
    rm = []
    for K in ALL_K:
        idx = R.ref_cb_index(K)
        for rv in range(4):
            for sb in (0, 1):
                for E in (3 * K + 12, 2 * K + 6, 5 * K):
                    e = (int_gauss(E, K * 8 + rv) % 2001 - 1000).astype(np.int16)
                    s = (int_gauss(18600, K) % 20001 - 10000).astype(np.int16)
                    assert R.ref_rm_rx(P(e), E, idx, rv, sb, P(s)) == 0
                    rm.append((K, rv, sb, E, crc32(s)))
    out["rm"] = np.array(rm, np.int64)
    crc = []
    for n in (8, 24, 40, 5000, 5824, 6144, 75376, 75400):
        d = (int_gauss(n // 8, n) & 255).astype(np.uint8)
        crc.append((n, R.ref_crc(0, P(d), n), R.ref_crc(1, P(d), n)))
    out["crc"] = np.array(crc, np.int64)
    seg = []
    for tbs in list(range(16, 3000, 8)) + list(range(3000, 150000, 104)) + [75376, 149776]:
        o = np.zeros(9, np.uint32)
        ret = R.ref_cbsegm(tbs, P(o))
        seg.append([tbs, ret] + [int(x) for x in o])
    out["cbsegm"] = np.array(seg, np.int64)
    np.savez_compressed(os.path.join(OUT, "tables.npz"), **out)


def gen_tb():
    """srslte_dlsch_encode -> integer AWGN -> srslte_dlsch_decode sequences (HARQ rv order 0,2,3,1)."""
    out = {}
    cases = []
    for tbs, Qm, nbits, sigma, max_its, ntx in [(75376, 6, 90000, 450, 8, 1), (75376, 6, 90000, 620, 8, 4),
                                                (75376, 6, 86400, 500, 8, 2), (75376, 6, 90000, 0, 10, 1),
                                                (1000, 2, 2400, 700, 8, 2), (16, 2, 288, 900, 10, 1),
                                                (6200, 4, 9600, 750, 8, 2), (12216, 2, 14400, 950, 8, 4),
                                                (3240, 2, 14400, 1300, 8, 3)]:
        ci = len(cases)
        seed = 5000 + 10 * ci
        t = C.c_void_p(R.ref_tb_new(100))
        data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
        data[: tbs // 8] = (int_gauss(tbs // 8, seed) & 255).astype(np.uint8)
        R.ref_tb_reset(t, tbs)
        ntx_done = 0
        for tx, rv in enumerate([0, 2, 3, 1][:ntx]):
            e = np.zeros(nbits, np.uint8)
            assert R.ref_tb_encode(t, tbs, Qm, rv, nbits, P(data), P(e)) == 0
            llr = int_awgn_llr(e, sigma, seed + 1 + tx) if sigma else (700 * (2 * e.astype(np.int16) - 1)).astype(np.int16)
            d = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
            crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
            d2 = np.zeros_like(d)
            ret = R.ref_tb_decode_detail(t, tbs, Qm, rv, nbits, P(llr), max_its, P(d), P(crc), P(its))
            out[f"e_{ci}_{tx}"] = np.packbits(e)
            out[f"data_{ci}_{tx}"] = d[: tbs // 8 + 3].copy()
            out[f"crc_{ci}_{tx}"] = crc
            out[f"its_{ci}_{tx}"] = its
            out[f"ret_{ci}_{tx}"] = np.array([ret], np.int64)
            seg = np.zeros(9, np.uint32)
            R.ref_cbsegm(tbs, P(seg))
            sc = []
            for cb in range(int(seg[1])):
                s = np.zeros(18600, np.int16)
                R.ref_tb_get_soft(t, cb, P(s))
                K = int(seg[2]) if cb < int(seg[6]) else int(seg[3])
                sc.append(crc32(s[:3 * K + 12]) if R.ref_subblocks(K) == 0 else
                          crc32(np.concatenate([s[0:K], s[K + 32:2 * K + 32], s[2 * K + 64:3 * K + 64], s[3 * K + 96:3 * K + 108]])))
            out[f"softcrc_{ci}_{tx}"] = np.array(sc, np.uint32)
            ntx_done += 1
            if ret == 0:
                break
        out[f"tbdata_{ci}"] = data[: tbs // 8].copy()
        cases.append((tbs, Qm, nbits, sigma, max_its, ntx_done, seed))
        R.ref_tb_free(t)
    out["cases"] = np.array(cases, np.int64)
    np.savez_compressed(os.path.join(OUT, "tb.npz"), **out)


def fe_symbols(n, seed):
    """deterministic equalised symbols: integer pseudo-Gaussians scaled in float64, rounded once to float32"""
    from _libs import aligned
    a = aligned(2 * n, np.float32)
    a[:] = (int_gauss(2 * n, seed).astype(np.float64) * (0.75 / 65536.0)).astype(np.float32)
    return a


def fe_csi(n, seed):
    from _libs import aligned
    a = aligned(n, np.float32)
    a[:] = (0.2 + (np.abs(int_gauss(n, seed)) % 60000).astype(np.float64) / 40000.0).astype(np.float32)
    return a


def gen_frontend():
    """LLRs the reference hands to decode_tb from equalised symbols: DL = demodulate + descramble (+ csi_correction),
    UL = pusch.c:617-657 with HARQ-ACK / RI positions (q after descrambling and g after ulsch_deinterleave)."""
    out = {}
    dl, ul = [], []
    for mod, n, use_csi, c_init in [(1, 37, 0, 11), (2, 1201, 0, 0x1234567), (2, 1201, 1, 0x1234567), (4, 2401, 1, 99), (4, 403, 0, 7),
                                    (6, 15000, 0, 0x2C0A21), (6, 15000, 1, 0x2C0A21), (6, 1001, 1, 5), (6, 7, 0, 1)]:
        i = len(dl)
        sym = fe_symbols(n, 700 + i)
        e = np.zeros(n * mod + 32, np.int16)
        assert R.ref_demod_s(mod, P(sym), P(e), n) == 0
        assert R.ref_descramble_s(c_init, P(e), n * mod) == 0
        if use_csi:
            assert R.ref_csi_s(mod, P(e), P(fe_csi(n, 900 + i)), n * mod) == 0
        e = e[: n * mod]
        out[f"dl_crc_{i}"] = np.array([crc32(e)], np.uint32)
        if n <= 1300:
            out[f"dl_e_{i}"] = e.copy()
        dl.append((mod, n, use_csi, c_init, 700 + i, 900 + i))
    for mod, prb, cols, nack, nri, c_init in [(2, 6, 12, 0, 0, 3), (4, 10, 12, 1, 1, 0x55AA55), (4, 25, 12, 2, 0, 77), (6, 12, 11, 0, 1, 1234567)]:
        i = len(ul)
        n = 12 * prb * cols
        sym = fe_symbols(n, 800 + i)
        t = C.c_void_p(R.ref_ul_new(100))
        uci_len = np.array([nack, nri, 0], np.uint32)
        ioff = np.array([9, 6, 7], np.uint32)
        data, o1, o2 = np.zeros(4096, np.uint8), np.zeros(4, np.uint8), np.zeros(64, np.uint8)
        q, g = np.zeros(n * mod, np.int16), np.zeros(n * mod, np.int16)
        ap, rp, cnt = np.zeros(12 * 288, np.uint32), np.zeros(12 * 288, np.uint32), np.zeros(2, np.uint32)
        R.ref_ul_decode(t, 0, mod, 0, prb, cols, c_init, P(sym), P(uci_len), P(ioff), 8, P(data), P(o1), P(o2), P(q), P(g), None,
                        P(ap), P(rp), P(cnt))
        R.ref_ul_free(t)
        ng = n * mod - int(cnt[1])
        out[f"ul_ack_{i}"], out[f"ul_ri_{i}"] = ap[: cnt[0]].copy(), rp[: cnt[1]].copy()
        out[f"ul_crc_{i}"] = np.array([crc32(q), crc32(g[:ng])], np.uint32)
        out[f"ul_g0_{i}"] = g[:64].copy()
        ul.append((mod, prb, cols, nack, nri, c_init, 800 + i))
    out["dl"], out["ul"] = np.array(dl, np.int64), np.array(ul, np.int64)
    np.savez_compressed(os.path.join(OUT, "frontend.npz"), **out)


def llr8_input(coded, sigma_milli, seed, amp=24):
    """int8 LLRs as the 8-bit demodulator scale delivers them: amp*(2b-1) + noise, clipped to int8"""
    return np.clip(int_awgn_llr(coded, sigma_milli, seed, amp=amp), -128, 127).astype(np.int8)


def soft8_crc(s8, K, P8):
    if P8 == 0:
        return crc32(s8[:3 * K + 12])
    return crc32(np.concatenate([s8[0:K], s8[K + 32:2 * K + 32], s8[2 * K + 64:3 * K + 64], s8[3 * K + 96:3 * K + 108]]))


def gen_llr8():
    """8-bit LLR path: srslte_tdec_iteration_8bit snapshots on natural-layout int8 inputs (all four decoder
    classes: avx8 P=32, sse8 P=16, and the 16-bit fallbacks P=8 / 0) and srslte_dlsch_decode sequences with llr_is_8bit.
    Transport blocks whose code blocks fall to the P=8 fallback are left out: in soft-buffer layout the reference widens only
    the first 3K+12 of the 3K+108 entries (turbodecoder.c:427-431,455-459) and decodes stale memory behind them."""
    out = {}
    cases = []
    for K, kind, par, nof in [(40, "awgn", 700, 6), (400, "awgn", 800, 6), (408, "awgn", 750, 8), (800, "awgn", 800, 8),
                              (816, "awgn", 700, 8), (1008, "awgn", 750, 8), (1024, "rand", 127, 8), (2048, "awgn", 650, 8),
                              (2112, "awgn", 700, 8), (3136, "rand", 60, 8), (5824, "awgn", 600, 8), (6144, "awgn", 700, 8),
                              (6144, "rand", 127, 6)]:
        i = len(cases)
        seed = 9000 + i
        if kind == "awgn":
            bits = (int_gauss(K, seed) & 1).astype(np.uint8)
            coded = np.zeros(3 * K + 12, np.uint8)
            R.ref_encode_cb(K, P(bits), P(coded))
            inp = llr8_input(coded, par, seed + 1)
            out[f"coded_{i}"] = np.packbits(coded)
        else:
            inp = (int_gauss(3 * K + 12, seed) % (2 * par + 1) - par).astype(np.int8)
        ob = np.zeros((nof, K // 8), np.uint8)
        ol = np.zeros((nof, K), np.int16)
        assert R.ref_tdec_halfits8(K, P(inp), nof, P(ob), P(ol)) == 0
        out[f"bytes_{i}"] = ob
        out[f"llrcrc_{i}"] = np.array([crc32(ol[n]) for n in range(nof)], np.uint32)
        out[f"incrc_{i}"] = np.array([crc32(inp)], np.uint32)
        cases.append((K, 0 if kind == "awgn" else 1, par, nof, seed))
    out["cases"] = np.array(cases, np.int64)
    tbc = []
    for tbs, Qm, nbits, sigma, max_its, ntx in [(75376, 6, 90000, 450, 8, 1), (75376, 6, 90000, 620, 8, 4), (36696, 4, 57600, 600, 8, 2),
                                                (6200, 6, 14400, 700, 8, 2), (1000, 4, 1600, 550, 8, 2), (1800, 2, 3000, 900, 8, 3),
                                                (256, 2, 600, 700, 10, 2), (12960, 2, 20000, 800, 8, 4)]:
        ci = len(tbc)
        seed = 9500 + 10 * ci
        t = C.c_void_p(R.ref_tb_new(100))
        data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
        data[: tbs // 8] = (int_gauss(tbs // 8, seed) & 255).astype(np.uint8)
        R.ref_tb_reset(t, tbs)
        seg = np.zeros(9, np.uint32)
        R.ref_cbsegm(tbs, P(seg))
        assert R.ref_subblocks8(int(seg[2])) != 8 and (int(seg[7]) == 0 or R.ref_subblocks8(int(seg[3])) != 8)
        done = 0
        for tx, rv in enumerate([0, 2, 3, 1][:ntx]):
            e = np.zeros(nbits, np.uint8)
            assert R.ref_tb_encode(t, tbs, Qm, rv, nbits, P(data), P(e)) == 0
            llr = aligned(nbits + 64, np.int8)
            llr[:nbits] = llr8_input(e, sigma, seed + 1 + tx)
            d = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
            crc, avg = np.zeros(13, np.uint8), np.zeros(1, np.uint32)
            ret = R.ref_tb_decode8(t, tbs, Qm, rv, nbits, P(llr), max_its, P(d), P(crc), P(avg))
            out[f"tb_e_{ci}_{tx}"] = np.packbits(e)
            out[f"tb_data_{ci}_{tx}"] = d[: tbs // 8 + 3].copy()
            out[f"tb_crc_{ci}_{tx}"] = crc
            out[f"tb_ret_{ci}_{tx}"] = np.array([ret, int(avg[0])], np.int64)
            sc = []
            for cb in range(int(seg[1])):
                s = np.zeros(18600, np.int16)
                R.ref_tb_get_soft(t, cb, P(s))
                K = int(seg[2]) if cb < int(seg[6]) else int(seg[3])
                sc.append(soft8_crc(s.view(np.int8), K, R.ref_subblocks8(K)))
            out[f"tb_softcrc_{ci}_{tx}"] = np.array(sc, np.uint32)
            done += 1
            if ret == 0:
                break
        out[f"tb_payload_{ci}"] = data[: tbs // 8].copy()
        tbc.append((tbs, Qm, nbits, sigma, max_its, done, seed))
        R.ref_tb_free(t)
    out["tb_cases"] = np.array(tbc, np.int64)
    np.savez_compressed(os.path.join(OUT, "llr8.npz"), **out)


def gen_bler():
    """row g1 (north_star: "against turbodecoder_gen ... matching BER/BLER within stated tolerance on the same AWGN vectors"):
    block / bit error counts of the reference's generic full-length int16 decoder (srslte_tdec_init_manual(...,
    SRSLTE_TDEC_GENERIC), turbodecoder.c:173, turbodecoder_gen.c:59-203) and of its AUTO SSE/AVX decoders after 8
    half-iterations, 2000 seeded code blocks per point, sigma across the waterfall of each K"""
    import _golden as G
    GENERIC = 1                                   # srslte_tdec_impl_type_t, turbodecoder_impl.h:33-42
    rows = []
    grid = {504: (1000, 1050, 1100, 1150, 1200), 2048: (1050, 1075, 1100, 1125, 1150), 5824: (1075, 1100, 1125), 6144: (1075, 1100, 1125)}
    for K, sms in grid.items():
        for sm in sms:
            n, seed, nof = 2000, 900000 + K * 10 + sm, 8
            bits, llr = G.bler_vectors(K, sm, n, seed)
            cnt = []
            for mode in (GENERIC, 0):
                ob = np.zeros((n, K // 8), np.uint8)
                assert R.ref_tdec_batch(K, P(llr), n, nof, mode, P(ob)) == 0
                err = np.unpackbits(ob, axis=1) != bits
                cnt += [int(err.any(axis=1).sum()), int(err.sum())]
            rows.append((K, sm, n, seed, nof, crc32(llr)) + tuple(cnt))
            print("bler", rows[-1], flush=True)
    np.savez_compressed(os.path.join(OUT, "bler_gen.npz"), cases=np.array(rows, np.int64))


if __name__ == "__main__":
    if "llr8" in sys.argv[1:]:        # only the 8-bit fixtures (the others stay byte-identical)
        gen_llr8()
    elif "bler" in sys.argv[1:]:      # only the BLER statistics against the generic decoder
        gen_bler()
    else:
        gen_bler()
        gen_frontend()
        gen_tdec()
        gen_tables()
        gen_tb()
        gen_llr8()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
