"""Helpers shared by the front-end tests: oracle chain from equalised symbols to decode_tb's LLRs, UL control
positions, and synthetic DL / UL transmissions (symbols) built with the numpy synthesiser."""
import importlib

import numpy as np

from _libs import P, aligned

synth = importlib.import_module("empower-srslte_b200.synth")


uci_positions, ACK_COLS, RI_COLS = synth.uci_positions, synth.ACK_COLS, synth.RI_COLS


def oracle_frontend(orc, sym, mod, c_init, csi=None, nof_symb=0, ack_pos=(), ri_pos=()):
    """the reference's call sequence on the oracle's restatements: DL pdsch.c:795-813, UL pusch.c:617-657 + sch.c:903-958.
    Returns (llr as decode_tb gets it [DL e / UL g], raw q LLRs at ack_pos + ri_pos)"""
    sym = np.ascontiguousarray(sym, np.float32).reshape(-1)
    n = sym.size // 2
    q = np.zeros(n * mod, np.int16)
    assert orc.orc_demod_s(mod, P(sym), P(q), n) == 0
    raw = q[list(ack_pos) + list(ri_pos)].copy()
    q[list(ack_pos)] = 0
    c = np.zeros(n * mod, np.uint8)
    orc.orc_sequence(c_init, n * mod, P(c))
    orc.orc_descramble_s(P(c), P(q), n * mod)
    if nof_symb == 0:
        if csi is not None:
            orc.orc_csi_s(mod, P(q), P(np.ascontiguousarray(csi, np.float32)), n * mod)
        return q, raw
    g = np.zeros(n * mod, np.int16)
    ri = np.array(list(ri_pos) + [0], np.uint32)
    orc.orc_ulsch_deinterleave(P(q), mod, n, nof_symb, P(g), P(ri), len(ri_pos))
    return g, raw


def make_dl(payload, tbs, mod, rv, nof_re, c_init, sigma, rng):
    """payload (ntb, tbs/8) -> noisy equalised symbols float32 (ntb, nof_re, 2)"""
    e = synth.encode_tbs(payload, tbs, mod, rv, nof_re * mod)
    c = synth.gold_sequence(c_init, nof_re * mod)
    return synth.awgn_symbols(synth.modulate(e ^ c[None, :], mod), sigma, rng)


def make_ul(payload, tbs, mod, rv, prb, cols, c_init, sigma, rng, ri_pos=(), n_cqi=0):
    """UL-SCH data multiplexed behind n_cqi (random) CQI bits, channel-interleaved around the RI positions"""
    nof_re = 12 * prb * cols
    G = nof_re * mod - len(ri_pos) - n_cqi
    e = synth.encode_tbs(payload, tbs, mod, rv, G)
    cqi = rng.integers(0, 2, (payload.shape[0], n_cqi)).astype(e.dtype)
    q = synth.ulsch_interleave(np.concatenate([cqi, e], axis=1), mod, cols, ri_pos)
    c = synth.gold_sequence(c_init, nof_re * mod)
    return synth.awgn_symbols(synth.modulate(q ^ c[None, :], mod), sigma, rng), G


def sym_buffer(sym):
    """16-byte aligned float32 copy of (nof_re, 2) symbols"""
    flat = np.ascontiguousarray(sym, np.float32).reshape(-1)
    a = aligned(flat.size, np.float32)
    a[:] = flat
    return a
