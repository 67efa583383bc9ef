"""GPU parity of the front end in 8-bit LLR mode (q->llr_is_8bit, pdsch.c:795-813): srslte_demod_soft_demodulate_b,
srslte_scrambling_sb_offset and the int8 branch of csi_correction in front of the 8-bit rate de-matcher and decoders, through
fec_cuda_frontend / fec_cuda_decode_syms after fec_cuda_set_llr_8bit. Checker: the CPU oracle, which tests/test_oracle_frontend.py
pins to the unmodified reference for exactly these functions. Bit-exact."""
import ctypes as C
import importlib

import numpy as np
import pytest

import _fe
from _libs import P

pytestmark = pytest.mark.gpu

pkg = importlib.import_module("empower-srslte_b200")
synth = importlib.import_module("empower-srslte_b200.synth")


@pytest.fixture(scope="module")
def fec8():
    f = pkg.FecCuda(0)
    f.set_llr_8bit(True)
    yield f
    f.close()


def oracle_fe8(orc, sym, mod, c_init, csi):
    n = sym.size // 2
    buf = _fe.sym_buffer(np.asarray(sym, np.float32).reshape(-1, 2))
    e = np.zeros(n * mod + 64, np.int8)
    assert orc.orc_demod_b(mod, P(buf), P(e), n) == 0
    c = np.zeros(n * mod, np.uint8)
    orc.orc_sequence(c_init & 0x7FFFFFFF, n * mod, P(c))
    orc.orc_descramble_b(P(c), P(e), n * mod)
    if csi is not None:
        orc.orc_csi_b(mod, P(e), P(np.ascontiguousarray(csi, np.float32)), n * mod)
    return e[: n * mod]


@pytest.mark.parametrize("mod", [1, 2, 4, 6])
def test_frontend8_sizes_tails_and_saturation(fec8, orc, mod):
    """lengths that leave symbols to the reference's scalar tails (8 symbols / 16 QPSK values per SSE step), rounding ties,
    values that saturate both packs, tile boundaries of the kernel (2048 symbols), CSI on and off"""
    rng = np.random.default_rng(300 + mod)
    for n in (1, 2, 3, 7, 8, 9, 17, 2047, 2048, 2049, 4099, 20001):
        sym = (rng.standard_normal(2 * n) * 0.8).astype(np.float32)
        k = min(2 * n, 40)
        sym[:k] = (np.arange(k) - 20.5) / 80.0
        if mod in (4, 6) and n >= 16:
            sym[0] = 1e12; sym[1] = -60.0; sym[2] = 3.3          # inside the SIMD body: int32 indefinite, both packs saturate
        for use_csi in (False, True):
            if use_csi and mod == 1:
                continue
            csi = rng.uniform(0.1, 1.9, n).astype(np.float32) if use_csi else None
            got, _ = fec8.frontend(sym, mod, 0x2F2E3D4 + n, csi=csi, llr8=True)
            want = oracle_fe8(orc, sym, mod, 0x2F2E3D4 + n, csi)
            assert (got == want).all(), (mod, n, use_csi, np.nonzero(got != want)[0][:8])


def test_frontend8_refuses_uplink(fec8):
    with pytest.raises(RuntimeError):
        fec8.frontend(np.zeros(2 * 144, np.float32), 4, 1, nof_symb=12, llr8=True)


@pytest.mark.parametrize("tbs,mod,nof_re,sigma,use_csi", [(75376, 6, 15000, 0.05, False), (75376, 6, 15000, 0.05, True), (6200, 4, 2401, 0.12, False)])
def test_decode_syms_8bit_dl_harq(fec8, orc, tbs, mod, nof_re, sigma, use_csi):
    """PDSCH codeword decode from symbols in 8-bit mode: a batch of HARQ processes over two transmissions (int8 combining)
    against the oracle chain demod_b -> descramble_b -> csi_b -> decode_tb with llr_is_8bit"""
    ntb = 4
    rng = np.random.default_rng(tbs + 5 * mod)
    payload = rng.integers(0, 256, (ntb, tbs // 8), dtype=np.uint8)
    sb0 = fec8.softbuffer_alloc(ntb)
    ots = [C.c_void_p(orc.orc_tb_new()) for _ in range(ntb)]
    nbits = nof_re * mod
    for txi, rv in enumerate((0, 2)):
        for ot in ots:
            if txi == 0:
                orc.orc_tb_reset(ot, tbs)
        e = synth.encode_tbs(payload, tbs, mod, rv, nbits)
        d = (pkg.TbDesc * ntb)()
        s = (pkg.SymDesc * ntb)()
        keep, wants = [], []
        for i in range(ntb):
            c_init = synth.cinit_pdsch(0x1234 + i, 0, 2 * ((i + txi) % 10), 7)
            sym = synth.awgn_symbols(synth.modulate(e[i] ^ synth.gold_sequence(c_init, nbits), mod)[None], sigma * (1.6 if txi == 0 else 1.0), rng)[0]
            csi = rng.uniform(0.6, 1.4, nof_re).astype(np.float32) if use_csi else None
            buf = _fe.sym_buffer(sym)
            out = np.zeros(tbs // 8 + 3 + 768, np.uint8)
            keep.append((buf, out, csi))
            d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its = tbs, mod, rv, nbits, 8
            d[i].softbuffer, d[i].new_data, d[i].data = sb0 + i, int(txi == 0), out.ctypes.data
            s[i].symbols, s[i].nof_re, s[i].mod_bits, s[i].c_init = buf.ctypes.data, nof_re, mod, c_init
            if csi is not None:
                s[i].csi = csi.ctypes.data
            llr = oracle_fe8(orc, sym.reshape(-1), mod, c_init, csi)
            data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
            crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
            ret = orc.orc_tb_decode8(ots[i], tbs, mod, rv, nbits, P(np.ascontiguousarray(llr)), 8, P(data), P(crc), P(its))
            wants.append((ret, data, list(crc), list(its)))
        fec8.decode_syms(d, s)
        for i in range(ntb):
            ret, data, crc, its = wants[i]
            assert d[i].ret == ret, (txi, i, d[i].ret, ret)
            assert list(d[i].cb_its) == its and list(d[i].cb_crc) == crc, (txi, i)
            assert (keep[i][1][: tbs // 8 + 3] == data[: tbs // 8 + 3]).all(), (txi, i)
    for ot in ots:
        orc.orc_tb_free(ot)
