"""Readers for tests/golden/*.npz (produced by oracle/gen_golden.py from the unmodified reference)."""
import os
import zlib

import numpy as np

from _libs import int_awgn_llr, int_gauss

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def crc32(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def tdec_cases():
    """yield dict(K, inp int16[3K+12], nof, bytes[nof][K/8], llrcrc[nof], llr or None)"""
    z = np.load(os.path.join(GOLD, "tdec.npz"))
    for i, (K, kind, par, nof, seed) in enumerate(z["cases"]):
        K, par, nof, seed = int(K), int(par), int(nof), int(seed)
        if kind == 0:
            coded = np.unpackbits(z[f"coded_{i}"])[: 3 * K + 12]
            inp = int_awgn_llr(coded, par, seed + 1)
        else:
            inp = (int_gauss(3 * K + 12, seed) % (2 * par + 1) - par).astype(np.int16)
        assert crc32(inp) == int(z[f"incrc_{i}"][0]), "input generator drifted"
        yield dict(K=K, inp=inp, nof=nof, bytes=z[f"bytes_{i}"], llrcrc=z[f"llrcrc_{i}"],
                   llr=z[f"llr_{i}"] if f"llr_{i}" in z.files else None, id=f"K{K}-{'awgn' if kind == 0 else 'rand'}{par}")


def tb_cases():
    """yield dict(tbs,Qm,nbits,max_its, tx=[dict(rv, llr, data, crc, its, ret, softcrc)], tbdata)"""
    z = np.load(os.path.join(GOLD, "tb.npz"))
    for ci, (tbs, Qm, nbits, sigma, max_its, ntx, seed) in enumerate(z["cases"]):
        tbs, Qm, nbits, sigma, max_its, ntx, seed = (int(v) for v in (tbs, Qm, nbits, sigma, max_its, ntx, seed))
        txs = []
        for tx in range(ntx):
            e = np.unpackbits(z[f"e_{ci}_{tx}"])[:nbits]
            llr = int_awgn_llr(e, sigma, seed + 1 + tx) if sigma else (700 * (2 * e.astype(np.int16) - 1)).astype(np.int16)
            txs.append(dict(rv=[0, 2, 3, 1][tx], llr=llr, data=z[f"data_{ci}_{tx}"], crc=z[f"crc_{ci}_{tx}"],
                            its=z[f"its_{ci}_{tx}"], ret=int(z[f"ret_{ci}_{tx}"][0]), softcrc=z[f"softcrc_{ci}_{tx}"]))
        yield dict(tbs=tbs, Qm=Qm, nbits=nbits, max_its=max_its, tx=txs, tbdata=z[f"tbdata_{ci}"],
                   id=f"tbs{tbs}-s{sigma}-n{nbits}")


def tables():
    return np.load(os.path.join(GOLD, "tables.npz"))


def soft_crc_ref_layout(s, K, P):
    """checksum of a reference-layout soft buffer over LUT-addressed positions only (SURVEY trap 7)"""
    if P == 0:
        return crc32(s[: 3 * K + 12])
    return crc32(np.concatenate([s[0:K], s[K + 32:2 * K + 32], s[2 * K + 64:3 * K + 64], s[3 * K + 96:3 * K + 108]]))


def fe_symbols(n, seed):
    """same generator as oracle/gen_golden.py: 16-byte aligned float32 (re, im) pairs"""
    from _libs import aligned
    a = aligned(2 * n, np.float32)
    a[:] = (int_gauss(2 * n, seed).astype(np.float64) * (0.75 / 65536.0)).astype(np.float32)
    return a


def fe_csi(n, seed):
    from _libs import aligned
    a = aligned(n, np.float32)
    a[:] = (0.2 + (np.abs(int_gauss(n, seed)) % 60000).astype(np.float64) / 40000.0).astype(np.float32)
    return a


def fe_dl_cases():
    """yield dict(mod, n, sym, csi or None, c_init, crc, e or None): LLRs the reference hands to srslte_dlsch_decode2"""
    z = np.load(os.path.join(GOLD, "frontend.npz"))
    for i, (mod, n, use_csi, c_init, s1, s2) in enumerate(z["dl"]):
        mod, n, c_init = int(mod), int(n), int(c_init)
        yield dict(mod=mod, n=n, sym=fe_symbols(n, int(s1)), csi=fe_csi(n, int(s2)) if use_csi else None, c_init=c_init,
                   crc=int(z[f"dl_crc_{i}"][0]), e=z[f"dl_e_{i}"] if f"dl_e_{i}" in z.files else None,
                   id=f"dl-qm{mod}-n{n}-{'csi' if use_csi else 'nocsi'}")


def fe_ul_cases():
    """yield dict(mod, prb, cols, sym, c_init, ack_pos, ri_pos, crc_q, crc_g, g0): pusch.c:617-657 outputs"""
    z = np.load(os.path.join(GOLD, "frontend.npz"))
    for i, (mod, prb, cols, nack, nri, c_init, s1) in enumerate(z["ul"]):
        mod, prb, cols = int(mod), int(prb), int(cols)
        yield dict(mod=mod, prb=prb, cols=cols, sym=fe_symbols(12 * prb * cols, int(s1)), c_init=int(c_init),
                   ack_pos=[int(v) for v in z[f"ul_ack_{i}"]], ri_pos=[int(v) for v in z[f"ul_ri_{i}"]],
                   crc_q=int(z[f"ul_crc_{i}"][0]), crc_g=int(z[f"ul_crc_{i}"][1]), g0=z[f"ul_g0_{i}"],
                   id=f"ul-qm{mod}-prb{prb}-ack{int(nack)}-ri{int(nri)}")


def llr8_input(coded, sigma_milli, seed, amp=24):
    """int8 LLRs of the 8-bit fixtures (same generator as oracle/gen_golden.py)"""
    return np.clip(int_awgn_llr(coded, sigma_milli, seed, amp=amp), -128, 127).astype(np.int8)


def soft8_view(s8, K, P8):
    """the entries of an int8 reference soft buffer that carry data (sub-block layout leaves 32-entry gaps)"""
    if P8 == 0:
        return s8[:3 * K + 12]
    return np.concatenate([s8[0:K], s8[K + 32:2 * K + 32], s8[2 * K + 64:3 * K + 64], s8[3 * K + 96:3 * K + 108]])


def tdec8_cases():
    """8-bit decoder snapshots: dict(K, inp int8[3K+12], nof, bytes[nof][K/8], llrcrc[nof])"""
    z = np.load(os.path.join(GOLD, "llr8.npz"))
    for i, (K, kind, par, nof, seed) in enumerate(z["cases"]):
        K, par, nof, seed = int(K), int(par), int(nof), int(seed)
        if kind == 0:
            inp = llr8_input(np.unpackbits(z[f"coded_{i}"])[: 3 * K + 12], par, seed + 1)
        else:
            inp = (int_gauss(3 * K + 12, seed) % (2 * par + 1) - par).astype(np.int8)
        assert crc32(inp) == int(z[f"incrc_{i}"][0]), "input generator drifted"
        yield dict(K=K, inp=inp, nof=nof, bytes=z[f"bytes_{i}"], llrcrc=z[f"llrcrc_{i}"], id=f"K{K}-{'awgn' if kind == 0 else 'rand'}{par}")


def tb8_cases():
    """8-bit transport-block sequences: dict(tbs,Qm,nbits,max_its, tx=[dict(rv, llr int8, data, crc, ret, avg_its, softcrc)], payload)"""
    z = np.load(os.path.join(GOLD, "llr8.npz"))
    for ci, (tbs, Qm, nbits, sigma, max_its, ntx, seed) in enumerate(z["tb_cases"]):
        tbs, Qm, nbits, sigma, max_its, ntx, seed = (int(v) for v in (tbs, Qm, nbits, sigma, max_its, ntx, seed))
        txs = []
        for tx in range(ntx):
            e = np.unpackbits(z[f"tb_e_{ci}_{tx}"])[:nbits]
            txs.append(dict(rv=[0, 2, 3, 1][tx], llr=llr8_input(e, sigma, seed + 1 + tx), data=z[f"tb_data_{ci}_{tx}"],
                            crc=z[f"tb_crc_{ci}_{tx}"], ret=int(z[f"tb_ret_{ci}_{tx}"][0]), avg_its=int(z[f"tb_ret_{ci}_{tx}"][1]),
                            softcrc=z[f"tb_softcrc_{ci}_{tx}"]))
        yield dict(tbs=tbs, Qm=Qm, nbits=nbits, max_its=max_its, tx=txs, payload=z[f"tb_payload_{ci}"], id=f"tbs{tbs}-s{sigma}")


def bler_vectors(K, sigma_milli, n, seed):
    """n random code blocks of size K, turbo-encoded at rate 1/3 (natural layout [s p p'] x K + 12 tail) and sent over an
    integer AWGN channel: LLR = 100*(2b-1) + 100*sigma*g, g = sum of twelve 16-bit uniforms (Irwin-Hall, std 65536) from
    numpy's PCG64 integer stream. Returns (bits[n,K], llr[n,3K+12] int16). Same generator in oracle/gen_golden.py (fixture
    bler_gen.npz, which records crc32(llr) so that a drifting generator is noticed) and in the GPU test."""
    import importlib
    synth = importlib.import_module("empower-srslte_b200.synth")
    rng = np.random.Generator(np.random.PCG64(seed))
    bits = rng.integers(0, 2, size=(n, K), dtype=np.uint8)
    coded = synth.turbo_encode(bits)
    g = np.zeros(coded.shape, np.int64)
    for _ in range(12):
        g += rng.integers(0, 65536, size=coded.shape, dtype=np.uint16)
    g -= 6 * 65535
    noise = (100 * sigma_milli * g) // (1000 * 65536)
    llr = np.clip(100 * (2 * coded.astype(np.int64) - 1) + noise, -32768, 32767).astype(np.int16)
    return bits, llr


def bler_cases():
    """rows of tests/golden/bler_gen.npz: dict(K, sigma_milli, n, seed, nof, llrcrc, gen=(block errors, bit errors) of the
    reference's full-length int16 max-log-MAP (SRSLTE_TDEC_GENERIC, turbodecoder_gen.c:59-203), auto=(...) of its AUTO
    SSE/AVX decoders, both after nof half-iterations on these vectors)"""
    z = np.load(os.path.join(GOLD, "bler_gen.npz"))
    for K, sm, n, seed, nof, crc, gb, gbit, ab, abit in z["cases"]:
        yield dict(K=int(K), sigma_milli=int(sm), n=int(n), seed=int(seed), nof=int(nof), llrcrc=int(crc),
                   gen=(int(gb), int(gbit)), auto=(int(ab), int(abit)), id=f"K{int(K)}-s{int(sm)}")
