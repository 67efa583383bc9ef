"""Synthetic DL-SCH transport blocks for bench.py and the property tests (host side, numpy only).

The transmit half of the chain is NOT part of the accelerated path; this module
exists so that the GPU box, where /root/reference is absent, can make valid AWGN-corrupted transport
blocks of the BASELINE.json shapes: CRC24A attach -> code-block segmentation + CRC24B -> rate-1/3 turbo
encoding -> rate matching (36.212 5.1.1-5.1.4.1) -> BPSK-style LLRs `rint(amp*(+-1 + sigma*n))`, the
convention of the reference's own tests (lib/src/phy/fec/test/turbodecoder_test.c:245-252).

It is written from the 3GPP text, vectorised ACROSS transport blocks (the per-bit recursions loop over
K, every numpy op works on all code blocks at once), and pinned against the reference encoder
(srslte_dlsch_encode, lib/src/phy/phch/sch.c:188-296) by tests/test_synth.py through the golden
transport blocks in tests/golden/tb.npz.
"""
import math

import numpy as np

CRC24A = 0x1864CFB
CRC24B = 0x1800063

_K_TABLE = (list(range(40, 513, 8)) + list(range(528, 1025, 16)) + list(range(1056, 2049, 32))
            + list(range(2112, 6145, 64)))

# 36.212 table 5.1.3-3 (f1, f2) in K order
_F1F2 = None


def _qpp_params():
    """(f1, f2) per K, read from the C table the library itself is built from (csrc/lte_qpp_table.h)."""
    global _F1F2
    if _F1F2 is None:
        import os
        import re
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "lte_qpp_table.h")
        txt = open(path).read()
        rows = re.findall(r"\{\s*(\d+)\s*,\s*(\d+)\s*,\s*(\d+)\s*\}", txt)
        _F1F2 = {int(k): (int(a), int(b)) for k, a, b in rows}
        assert len(_F1F2) == 188
    return _F1F2


def cbsegm(tbs):
    """36.212 5.1.2 -> dict(C, K1, K2, C1, C2, F). Same arithmetic as srslte_cbsegm (cbsegm.c:58-108)."""
    B = tbs + 24
    if B <= 6144:
        C, Bp = 1, B
    else:
        C = int(math.ceil(np.float32(B) / np.float32(6120)))
        Bp = B + 24 * C
    need = (Bp - 1) // C + 1
    i1 = next(i for i, k in enumerate(_K_TABLE) if k >= need)
    K1 = _K_TABLE[i1]
    K2, C1, C2 = 0, 1, 0
    if C > 1:
        K2 = _K_TABLE[i1 - 1]
        C2 = (C * K1 - Bp) // (K1 - K2)
        C1 = C - C2
    return dict(C=C, K1=K1, K2=K2, C1=C1, C2=C2, F=C1 * K1 + C2 * K2 - Bp)


def _crc_table(poly):
    t = np.zeros(256, np.uint32)
    for b in range(256):
        r = b << 16
        for _ in range(8):
            r <<= 1
            if r & 0x1000000:
                r ^= poly
        t[b] = r & 0xFFFFFF
    return t


_TABS = {}


def crc24(poly, data):
    """CRC-24 (MSB first, init 0, no final xor) of each ROW of a uint8 byte matrix -> uint32 vector."""
    if poly not in _TABS:
        _TABS[poly] = _crc_table(poly)
    tab = _TABS[poly]
    data = np.atleast_2d(data)
    r = np.zeros(data.shape[0], np.uint32)
    for j in range(data.shape[1]):
        idx = ((r >> np.uint32(16)) ^ data[:, j]) & np.uint32(0xFF)
        r = ((r << np.uint32(8)) & np.uint32(0xFFFFFF)) ^ tab[idx]
    return r


def _crc_bytes(c):
    return np.stack([(c >> np.uint32(16)) & 0xFF, (c >> np.uint32(8)) & 0xFF, c & 0xFF], axis=1).astype(np.uint8)


def qpp(K):
    f1, f2 = _qpp_params()[K]
    i = np.arange(K, dtype=np.int64)
    return ((f1 * i + f2 * i * i) % K).astype(np.int64)


def _rsc(bits):
    """constituent encoder g0 = 1+D^2+D^3 (feedback), g1 = 1+D+D^3 on each row of bits[n, K]
    -> (parity[n, K], tail_x[n, 3], tail_z[n, 3])"""
    n, K = bits.shape
    z = np.empty((n, K), np.uint8)
    r0 = np.zeros(n, np.uint8)
    r1 = np.zeros(n, np.uint8)
    r2 = np.zeros(n, np.uint8)
    bt = np.ascontiguousarray(bits.T)
    zt = np.empty((K, n), np.uint8)
    for k in range(K):
        a = bt[k] ^ r1 ^ r2
        zt[k] = a ^ r0 ^ r2
        r2, r1, r0 = r1, r0, a
    z[:] = zt.T
    tx = np.empty((n, 3), np.uint8)
    tz = np.empty((n, 3), np.uint8)
    for k in range(3):
        x = r1 ^ r2                      # input that forces the feedback bit to 0
        tx[:, k] = x
        tz[:, k] = r0 ^ r2
        r2, r1, r0 = r1, r0, np.zeros(n, np.uint8)
    return z, tx, tz


def turbo_encode(bits):
    """bits[n, K] (0/1) -> coded[n, 3K+12] in the order [s0 p0 p'0 s1 p1 p'1 ...] + 12 tail bits
    (xK zK xK+1 zK+1 xK+2 zK+2 x'K z'K x'K+1 z'K+1 x'K+2 z'K+2), 36.212 5.1.3.2."""
    n, K = bits.shape
    p0, tx0, tz0 = _rsc(bits)
    p1, tx1, tz1 = _rsc(bits[:, qpp(K)])
    out = np.empty((n, 3 * K + 12), np.uint8)
    out[:, 0:3 * K:3] = bits
    out[:, 1:3 * K:3] = p0
    out[:, 2:3 * K:3] = p1
    for k in range(3):
        out[:, 3 * K + 2 * k] = tx0[:, k]
        out[:, 3 * K + 2 * k + 1] = tz0[:, k]
        out[:, 3 * K + 6 + 2 * k] = tx1[:, k]
        out[:, 3 * K + 6 + 2 * k + 1] = tz1[:, k]
    return out


_PERM = np.array([0, 16, 8, 24, 4, 20, 12, 28, 2, 18, 10, 26, 6, 22, 14, 30,
                  1, 17, 9, 25, 5, 21, 13, 29, 3, 19, 11, 27, 7, 23, 15, 31])


def rm_order(K):
    """36.212 5.1.4.1: the circular buffer of one code block with <NULL>s removed.
    Returns (seq, k0cnt): seq[j] = flat coded index (3*i + stream) of the j-th non-NULL buffer entry,
    k0cnt[rv] = number of non-NULL entries before k0(rv)."""
    D = K + 4
    R = (D - 1) // 32 + 1
    KP = 32 * R
    ND = KP - D
    w = np.arange(3 * KP)
    y = np.empty(3 * KP, np.int64)
    stream = np.empty(3 * KP, np.int64)
    a = w[:KP]
    y[:KP] = (a % R) * 32 + _PERM[a // R]
    stream[:KP] = 0
    m = (w[KP:] - KP) // 2
    odd = ((w[KP:] - KP) & 1) == 1
    y1 = (m % R) * 32 + _PERM[m // R]
    y2 = (_PERM[m // R] + 32 * (m % R) + 1) % KP
    y[KP:] = np.where(odd, y2, y1)
    stream[KP:] = np.where(odd, 2, 1)
    valid = y >= ND
    seq = (3 * (y - ND) + stream)[valid]
    before = np.concatenate([[0], np.cumsum(valid)])
    k0 = [R * (2 * int(math.ceil(np.float32(3 * KP) / np.float32(8 * R))) * rv + 2) for rv in range(4)]
    return seq, [int(before[k]) for k in k0]


def rate_match(coded, K, rv, E):
    """coded[n, 3K+12] -> e[n, E] (srslte_rm_turbo_tx semantics, rm_turbo.c:95-162)"""
    seq, k0 = rm_order(K)
    N = seq.size
    idx = seq[(k0[rv] + np.arange(E)) % N]
    return coded[:, idx]


def encode_tbs(payload, tbs, Qm, rv, nof_e_bits):
    """payload[n, tbs/8] uint8 -> e_bits[n, nof_e_bits] uint8 (0/1), the bits srslte_dlsch_encode emits
    (sch.c:188-296; per-CB E split of the ENCODER: sch.c:241 `i <= C-gamma-1`)."""
    payload = np.atleast_2d(payload)
    n = payload.shape[0]
    seg = cbsegm(tbs)
    assert seg["F"] == 0 and seg["C2"] == 0, "shapes with filler bits / two block sizes are rejected by the reference decoder"
    C, K = seg["C"], seg["K1"]
    tb = np.concatenate([payload, _crc_bytes(crc24(CRC24A, payload))], axis=1)       # (tbs+24)/8 bytes
    rlen = (K - 24) // 8 if C > 1 else K // 8
    Gp = nof_e_bits // Qm
    gamma = Gp % C
    out = np.empty((n, nof_e_bits), np.uint8)
    wp = 0
    for cb in range(C):
        blk = tb[:, cb * rlen:(cb + 1) * rlen]
        if C > 1:
            blk = np.concatenate([blk, _crc_bytes(crc24(CRC24B, blk))], axis=1)
        coded = turbo_encode(np.unpackbits(blk, axis=1))
        E = Qm * (Gp // C) if cb <= C - gamma - 1 else Qm * ((Gp - 1) // C + 1)
        out[:, wp:wp + E] = rate_match(coded, K, rv, E)
        wp += E
    assert wp == nof_e_bits
    return out


def awgn_llr(e_bits, sigma, rng, amp=100.0, out=None):
    """LLR = rint(amp*(+-1 + sigma*n)) as int16, bit 1 -> +amp"""
    x = np.ascontiguousarray(e_bits, dtype=np.float32) if e_bits.dtype == np.float32 else np.array(e_bits, dtype=np.float32, order="C")
    x *= 2.0
    x -= 1.0
    x += np.float32(sigma) * rng.standard_normal(e_bits.shape, dtype=np.float32)
    x *= np.float32(amp)
    np.rint(x, out=x)
    np.clip(x, -32768, 32767, out=x)
    if out is None:
        return np.ascontiguousarray(x.astype(np.int16))
    out[...] = x
    return out


# ------------------------------------------------------------------------------------------------------------------
# Symbols for the front end: scrambling + modulation (+ UL channel interleaver), 36.211 7.1/7.2,
# 36.212 5.2.2.8.  Host-side synthesis only; pinned against the reference's modulator and sequence generator in
# tests/test_synth.py.

def gold_sequence(c_init, n):
    """c(0..n-1) of 36.211 7.2 as a uint8 array; generated 28 bits per step on Python integers"""
    def run(state, taps, total):
        x, have = state, 31
        while have < total:
            new = 0
            for t in taps:
                new ^= x >> (have - 31 + t)
            x |= (new & 0x0FFFFFFF) << have
            have += 28
        return x
    total = 1600 + n + 31
    x1 = run(1, (0, 3), total)
    x2 = run(c_init & 0x7FFFFFFF, (0, 1, 2, 3), total)
    c = ((x1 ^ x2) >> 1600) & ((1 << n) - 1)
    return np.unpackbits(np.frombuffer(c.to_bytes((n + 7) // 8, "little"), np.uint8), bitorder="little")[:n].copy()


def cinit_pdsch(rnti, q, nslot, cell_id):
    return ((rnti << 14) + (q << 13) + ((nslot // 2) << 9) + cell_id) & 0xFFFFFFFF


def cinit_pusch(rnti, nslot, cell_id):
    return ((rnti << 14) + ((nslot // 2) << 9) + cell_id) & 0xFFFFFFFF


def modulate(bits, mod_bits):
    """36.211 7.1 mapping; bits: (..., n*mod_bits) of 0/1 -> complex64 (..., n)"""
    b = 1.0 - 2.0 * bits.reshape(bits.shape[:-1] + (-1, mod_bits)).astype(np.float64)
    if mod_bits == 1:
        i = q = b[..., 0] / math.sqrt(2)
    elif mod_bits == 2:
        i, q = b[..., 0] / math.sqrt(2), b[..., 1] / math.sqrt(2)
    elif mod_bits == 4:
        i, q = b[..., 0] * (2 - b[..., 2]) / math.sqrt(10), b[..., 1] * (2 - b[..., 3]) / math.sqrt(10)
    else:
        i = b[..., 0] * (4 - b[..., 2] * (2 - b[..., 4])) / math.sqrt(42)
        q = b[..., 1] * (4 - b[..., 3] * (2 - b[..., 5])) / math.sqrt(42)
    return (i + 1j * q).astype(np.complex64)


def ulsch_interleave(g_bits, mod_bits, cols, ri_pos=()):
    """UL-SCH channel interleaver without control multiplexing payloads (36.212 5.2.2.8): g -> q; RI positions are left 0"""
    n = g_bits.shape[-1] + len(ri_pos)
    rows = n // mod_bits // cols
    taken = np.zeros(n, bool)
    taken[list(ri_pos)] = True
    order = (np.arange(cols)[None, :, None] * rows + np.arange(rows)[:, None, None]) * mod_bits + np.arange(mod_bits)[None, None, :]
    order = order.reshape(-1)                        # q positions in the order g is consumed (row by row)
    order = order[~taken[order]]
    q = np.zeros(g_bits.shape[:-1] + (n,), g_bits.dtype)
    q[..., order] = g_bits
    return q


def awgn_symbols(sym, sigma, rng):
    """complex AWGN with standard deviation sigma per real dimension, float32 (re, im) pairs out"""
    out = np.empty(sym.shape + (2,), np.float32)
    out[..., 0] = sym.real + sigma * rng.standard_normal(sym.shape)
    out[..., 1] = sym.imag + sigma * rng.standard_normal(sym.shape)
    return out


ACK_COLS = (2, 3, 8, 9)          # HARQ-ACK / RI columns of the channel interleaver, normal CP (uci.c:499-546)
RI_COLS = (1, 4, 7, 10)


def uci_positions(n_sym_uci, mod, rows, cols_set):
    """q-bit positions of n_sym_uci control symbols: rows from the bottom, columns in the (3*i)%4 visiting order
    (uci.c:499-546: HARQ-ACK {2,3,8,9}, RI {1,4,7,10}, normal CP)"""
    pos = []
    for i in range(n_sym_uci):
        row, col = rows - 1 - i // 4, cols_set[(3 * i) % 4]
        pos += [row * mod + rows * col * mod + k for k in range(mod)]
    return pos
