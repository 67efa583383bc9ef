"""GPU parity of the front end (SURVEY 8f rows f1 + f2): equalised symbols -> soft demodulation -> descrambling -> CSI
weights (DL) / HARQ-ACK zeroing + channel de-interleaver around the RI positions (UL) -> decode_tb, through the C ABI
(fec_cuda_frontend, fec_cuda_decode_syms).  Checkers: the golden vectors generated from the unmodified reference and
the CPU oracle (itself pinned to the reference in test_oracle_frontend.py).  Bit-exact."""
import ctypes as C
import importlib

import numpy as np
import pytest

import _fe
import _golden as G
from _libs import P

pytestmark = pytest.mark.gpu

pkg = importlib.import_module("empower-srslte_b200")
synth = importlib.import_module("empower-srslte_b200.synth")


@pytest.fixture(scope="module")
def fec():
    f = pkg.FecCuda(0)
    yield f
    f.close()


@pytest.mark.parametrize("case", list(G.fe_dl_cases()), ids=lambda c: c["id"])
def test_frontend_dl_golden(fec, case):
    e, _ = fec.frontend(case["sym"], case["mod"], case["c_init"], csi=case["csi"])
    assert G.crc32(e) == case["crc"]
    if case["e"] is not None:
        assert (e == case["e"]).all()


@pytest.mark.parametrize("case", list(G.fe_ul_cases()), ids=lambda c: c["id"])
def test_frontend_ul_golden(fec, orc, case):
    g, uci = fec.frontend(case["sym"], case["mod"], case["c_init"], nof_symb=case["cols"], ack_pos=case["ack_pos"],
                          ri_pos=case["ri_pos"], nof_cqi_bits=24)
    ng = g.size - len(case["ri_pos"])
    assert (g[:64] == case["g0"]).all()
    assert G.crc32(g[:ng]) == case["crc_g"]
    # raw control LLRs (what srslte_uci_decode_ack_ri reads) and the CQI head of g against the oracle
    want_g, want_raw = _fe.oracle_frontend(orc, case["sym"], case["mod"], case["c_init"], None, case["cols"], case["ack_pos"], case["ri_pos"])
    assert (uci[: want_raw.size] == want_raw).all()
    assert (uci[want_raw.size:] == want_g[:24]).all()


@pytest.mark.parametrize("mod", [1, 2, 4, 6])
def test_frontend_sizes_tails_and_saturation(fec, orc, mod):
    """lengths that leave symbols to the reference's scalar tails, rounding ties, values that saturate in the SIMD bodies,
    tile boundaries of the kernel (2048 symbols), CSI on and off"""
    rng = np.random.default_rng(100 + mod)
    for n in (1, 2, 3, 5, 8, 17, 2047, 2048, 2049, 4099, 20001):
        sym = (rng.standard_normal(2 * n) * 0.8).astype(np.float32)
        k = min(2 * n, 40)
        sym[:k] = (np.arange(k) - 20.5) / 1400.0
        if mod in (4, 6) and n >= 8:
            sym[8 * (n // 8 // 2)] = 1e12
            sym[8 * (n // 8 // 2) + 1] = -60.0
        for use_csi in (False, True):
            if use_csi and mod == 1:
                continue
            csi = rng.uniform(0.1, 1.9, n).astype(np.float32) if use_csi else None
            got, _ = fec.frontend(sym, mod, 0x1F2E3D4 + n, csi=csi)
            want, _ = _fe.oracle_frontend(orc, sym, mod, 0x1F2E3D4 + n, csi)
            assert (got == want).all(), (mod, n, use_csi, np.nonzero(got != want)[0][:8])


def test_frontend_argument_checks(fec):
    sym = np.zeros(2 * 24, np.float32)
    for kw in (dict(mod_bits=3), dict(mod_bits=6, nof_symb=5), dict(mod_bits=1, nof_symb=12), dict(mod_bits=2, nof_symb=12, ri_pos=[48])):
        with pytest.raises(RuntimeError):
            fec.frontend(sym, kw.pop("mod_bits"), 1, **kw)
    with pytest.raises(RuntimeError):
        fec.frontend(np.zeros(2 * 30000, np.float32), 6, 1)          # longer than MAX_SEQ_LEN (sequence.c:38)


def _run_syms(fec, tbs, tbQm, items, new_data, sb0, max_its=8, dev=False):
    """items: list of dict(sym (nof_re,2) float32, mod, c_init, csi, nof_symb, ack_pos, ri_pos, n_cqi, nof_e_bits, rv)"""
    n = len(items)
    d = (pkg.TbDesc * n)()
    s = (pkg.SymDesc * n)()
    keep = []
    for i, it in enumerate(items):
        buf = _fe.sym_buffer(it["sym"])
        out = np.zeros(tbs // 8 + 3 + 768, np.uint8)
        ack = np.ascontiguousarray(it.get("ack_pos", ()), dtype=np.uint32)
        ri = np.ascontiguousarray(it.get("ri_pos", ()), dtype=np.uint32)
        uci = np.zeros(ack.size + ri.size + it.get("n_cqi", 0) + 8, np.int16)
        keep.append((buf, out, ack, ri, uci))
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its = tbs, tbQm, it["rv"], it["nof_e_bits"], max_its
        d[i].softbuffer, d[i].new_data, d[i].data = sb0 + i, int(new_data), out.ctypes.data
        s[i].symbols, s[i].nof_re, s[i].mod_bits, s[i].c_init = buf.ctypes.data, buf.size // 2, it["mod"], it["c_init"]
        s[i].nof_symb = it.get("nof_symb", 0)
        if it.get("csi") is not None:
            c = np.ascontiguousarray(it["csi"], np.float32)
            keep.append(c)
            s[i].csi = c.ctypes.data
        s[i].ack_pos, s[i].ri_pos, s[i].nof_ack_pos, s[i].nof_ri_pos = ack.ctypes.data, ri.ctypes.data, ack.size, ri.size
        s[i].nof_cqi_bits, s[i].uci_out = it.get("n_cqi", 0), uci.ctypes.data
    fec.decode_syms(d, s)
    return d, [k[1] for k in keep if isinstance(k, tuple)], [k[4] for k in keep if isinstance(k, tuple)]


def _oracle_decode(orc, ot, tbs, Qm, rv, llr, max_its=8):
    data = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
    crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
    ret = orc.orc_tb_decode(ot, tbs, Qm, rv, llr.size, P(np.ascontiguousarray(llr)), max_its, P(data), P(crc), P(its))
    return ret, data, crc, its


@pytest.mark.parametrize("tbs,mod,nof_re,sigma,use_csi", [(75376, 6, 15000, 0.062, False), (75376, 6, 15000, 0.075, True),
                                                           (6200, 4, 2401, 0.17, False), (1000, 2, 1201, 0.45, True)])
def test_decode_syms_dl_harq(fec, orc, tbs, mod, nof_re, sigma, use_csi):
    """PDSCH codeword decode from symbols, a batch of HARQ processes over two transmissions, against the oracle chain"""
    ntb = 5
    rng = np.random.default_rng(tbs + 3 * mod)
    payload = rng.integers(0, 256, (ntb, tbs // 8), dtype=np.uint8)
    sb0 = fec.softbuffer_alloc(ntb)
    ots = [C.c_void_p(orc.orc_tb_new()) for _ in range(ntb)]
    for ot in ots:
        orc.orc_tb_reset(ot, tbs)
    live = list(range(ntb))
    acked = 0
    for tx, rv in enumerate((0, 2)):
        c_inits = [synth.cinit_pdsch(0x100 + i, 0, 2 * ((tx + i) % 10), 42) for i in live]
        items = []
        for s_, i in enumerate(live):
            sym = _fe.make_dl(payload[i:i + 1], tbs, mod, rv, nof_re, c_inits[s_], sigma, rng)[0]
            csi = rng.uniform(0.4, 1.6, nof_re).astype(np.float32) if use_csi else None
            items.append(dict(sym=sym, mod=mod, c_init=c_inits[s_], csi=csi, nof_e_bits=nof_re * mod, rv=rv))
        # soft buffers of the live processes are not contiguous any more after the first round: one call per process id run
        res = []
        for s_, i in enumerate(live):
            dd, outs, _ = _run_syms(fec, tbs, mod, [items[s_]], tx == 0, sb0 + i)
            res.append((dd[0], outs[0]))
        nxt = []
        for s_, i in enumerate(live):
            e, _ = _fe.oracle_frontend(orc, items[s_]["sym"], mod, c_inits[s_], items[s_]["csi"])
            ret, data, crc, its = _oracle_decode(orc, ots[i], tbs, mod, rv, e)
            dd, out = res[s_]
            assert dd.ret == ret, (tx, i)
            assert (out[: tbs // 8 + 3] == data[: tbs // 8 + 3]).all(), (tx, i)
            assert list(dd.cb_its) == list(its) and list(dd.cb_crc) == list(crc)
            if ret == 0:
                assert (out[: tbs // 8] == payload[i]).all()
                acked += 1
            else:
                nxt.append(i)
        live = nxt
        if not live:
            break
    assert acked >= 1
    for ot in ots:
        orc.orc_tb_free(ot)


def test_decode_syms_dl_batch_and_device_io(fec, orc):
    """one batched call over 40 transport blocks of mixed shapes is not possible (tbs is per TB, fine) - here: 40 x MCS28
    in ONE call, host symbols vs device symbols give identical results, and both equal the oracle"""
    import torch
    tbs, mod, nof_re, ntb = 75376, 6, 15000, 40
    rng = np.random.default_rng(77)
    payload = rng.integers(0, 256, (ntb, tbs // 8), dtype=np.uint8)
    c_inits = [synth.cinit_pdsch(61 + i, 0, 2 * (i % 10), 7) for i in range(ntb)]
    syms = np.stack([_fe.make_dl(payload[i:i + 1], tbs, mod, 0, nof_re, c_inits[i], 0.068, rng)[0] for i in range(ntb)])
    sb0 = fec.softbuffer_alloc(ntb)
    items = [dict(sym=syms[i], mod=mod, c_init=c_inits[i], nof_e_bits=nof_re * mod, rv=0) for i in range(ntb)]
    d1, o1, _ = _run_syms(fec, tbs, mod, items, True, sb0)
    # device-resident symbols and outputs
    dsym = torch.from_numpy(syms).cuda()
    dout = torch.zeros((ntb, tbs // 8 + 3 + 768), dtype=torch.uint8, device="cuda")
    d = (pkg.TbDesc * ntb)()
    s = (pkg.SymDesc * ntb)()
    for i in range(ntb):
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its = tbs, mod, 0, nof_re * mod, 8
        d[i].softbuffer, d[i].new_data, d[i].data = sb0 + i, 1, dout[i].data_ptr()
        s[i].symbols, s[i].nof_re, s[i].mod_bits, s[i].c_init = dsym[i].data_ptr(), nof_re, mod, c_inits[i]
    fec.decode_syms(d, s, pkg.F_DEVICE_IO)
    o2 = dout.cpu().numpy()
    ok = 0
    for i in range(ntb):
        assert d1[i].ret == d[i].ret and list(d1[i].cb_its) == list(d[i].cb_its)
        assert (o1[i][: tbs // 8 + 3] == o2[i, : tbs // 8 + 3]).all()
        if i < 6:
            ot = C.c_void_p(orc.orc_tb_new())
            orc.orc_tb_reset(ot, tbs)
            e, _ = _fe.oracle_frontend(orc, syms[i], mod, c_inits[i])
            ret, data, crc, its = _oracle_decode(orc, ot, tbs, mod, 0, e)
            orc.orc_tb_free(ot)
            assert d1[i].ret == ret and list(d1[i].cb_its) == list(its)
            assert (o1[i][: tbs // 8 + 3] == data[: tbs // 8 + 3]).all()
        if d1[i].ret == 0:
            assert (o1[i][: tbs // 8] == payload[i]).all()
            ok += 1
    assert ok > ntb // 2
    assert fec.last_frontend_ms() > 0


@pytest.mark.parametrize("tbs,mod,prb,n_ack_sym,n_ri_sym,n_cqi", [(15264, 4, 40, 0, 0, 0), (15264, 4, 40, 8, 4, 0), (2216, 2, 10, 20, 8, 32),
                                                                 (36696, 4, 96, 12, 0, 64), (0, 4, 6, 4, 4, 40)])
def test_decode_syms_ul_with_control(fec, orc, tbs, mod, prb, n_ack_sym, n_ri_sym, n_cqi):
    """PUSCH from the demodulator on (pusch.c:617-657): control positions as srslte_uci_decode_ack_ri places them
    (uci.c:499-546), G = H' - Q'_ri - Q'_cqi, data behind the CQI bits; tbs == 0 = control-only PUSCH"""
    cols, rows = 12, 12 * prb
    rng = np.random.default_rng(5 * prb + n_cqi)
    ack_pos = _fe.uci_positions(n_ack_sym, mod, rows, _fe.ACK_COLS)
    ri_pos = _fe.uci_positions(n_ri_sym, mod, rows, _fe.RI_COLS)
    c_init = synth.cinit_pusch(0x46, 6, 9)
    ntb = 3
    tb_eff = tbs if tbs else 1000
    payload = rng.integers(0, 256, (ntb, tb_eff // 8), dtype=np.uint8)
    sym, G = _fe.make_ul(payload, tb_eff, mod, 0, prb, cols, c_init, 0.16, rng, ri_pos, n_cqi)
    sb0 = fec.softbuffer_alloc(ntb)
    items = [dict(sym=sym[i], mod=mod, c_init=c_init, nof_symb=cols, ack_pos=ack_pos, ri_pos=ri_pos, n_cqi=n_cqi, nof_e_bits=G if tbs else 0, rv=0)
             for i in range(ntb)]
    d, outs, ucis = _run_syms(fec, tbs, mod, items, True, sb0)
    for i in range(ntb):
        g, raw = _fe.oracle_frontend(orc, sym[i], mod, c_init, None, cols, ack_pos, ri_pos)
        assert (ucis[i][: raw.size] == raw).all(), i
        assert (ucis[i][raw.size: raw.size + n_cqi] == g[:n_cqi]).all(), i
        if tbs == 0:
            assert d[i].ret == 0
            continue
        ot = C.c_void_p(orc.orc_tb_new())
        orc.orc_tb_reset(ot, tbs)
        ret, data, crc, its = _oracle_decode(orc, ot, tbs, mod, 0, g[n_cqi: n_cqi + G])
        orc.orc_tb_free(ot)
        assert d[i].ret == ret and list(d[i].cb_its) == list(its) and list(d[i].cb_crc) == list(crc)
        assert (outs[i][: tbs // 8 + 3] == data[: tbs // 8 + 3]).all()
        if n_ack_sym <= 8:
            assert ret == 0 and (outs[i][: tbs // 8] == payload[i]).all()


def test_decode_syms_ul_mixed_control_layouts_in_one_batch(fec, orc):
    """the host pass reuses the sorted position lists of an allocation for the following allocations with the same layout
    (fe_same_uci, csrc/fec_cuda_api.cu): one batch that alternates between three layouts (also one without any control
    information, and equal layouts given through different arrays) must decode every block as the oracle does"""
    tbs, mod, prb, cols = 15264, 4, 40, 12
    rows = 12 * prb
    rng = np.random.default_rng(4242)
    layouts = {"A": (8, 4, 0), "B": (0, 0, 0), "C": (12, 0, 64), "D": (8, 4, 32)}       # D: A's positions, other CQI length
    order = ["A", "A", "B", "A", "C", "C", "B", "D", "A", "D"]
    c_init = synth.cinit_pusch(0x51, 4, 3)
    payload = rng.integers(0, 256, (len(order), tbs // 8), dtype=np.uint8)
    items, want = [], []
    for i, key in enumerate(order):
        n_ack_sym, n_ri_sym, n_cqi = layouts[key]
        ack_pos = _fe.uci_positions(n_ack_sym, mod, rows, _fe.ACK_COLS)
        ri_pos = _fe.uci_positions(n_ri_sym, mod, rows, _fe.RI_COLS)
        sym, G = _fe.make_ul(payload[i:i + 1], tbs, mod, 0, prb, cols, c_init, 0.16, rng, ri_pos, n_cqi)
        items.append(dict(sym=sym[0], mod=mod, c_init=c_init, nof_symb=cols, ack_pos=list(ack_pos), ri_pos=list(ri_pos), n_cqi=n_cqi, nof_e_bits=G, rv=0))
        want.append((ack_pos, ri_pos, n_cqi, G))
    sb0 = fec.softbuffer_alloc(len(order))
    d, outs, ucis = _run_syms(fec, tbs, mod, items, True, sb0)
    for i, (ack_pos, ri_pos, n_cqi, G) in enumerate(want):
        g, raw = _fe.oracle_frontend(orc, items[i]["sym"], mod, c_init, None, cols, ack_pos, ri_pos)
        assert (ucis[i][: raw.size] == raw).all(), (i, order[i])
        assert (ucis[i][raw.size: raw.size + n_cqi] == g[:n_cqi]).all(), (i, order[i])
        ot = C.c_void_p(orc.orc_tb_new())
        orc.orc_tb_reset(ot, tbs)
        ret, data, crc, its = _oracle_decode(orc, ot, tbs, mod, 0, g[n_cqi: n_cqi + G])
        orc.orc_tb_free(ot)
        assert d[i].ret == ret and list(d[i].cb_its) == list(its) and list(d[i].cb_crc) == list(crc), (i, order[i])
        assert (outs[i][: tbs // 8 + 3] == data[: tbs // 8 + 3]).all(), (i, order[i])
        assert ret == 0 and (outs[i][: tbs // 8] == payload[i]).all(), (i, order[i])


@pytest.mark.parametrize("cols,ack_cols,ri_cols", [(11, (2, 3, 8, 9), (1, 4, 7, 10)),      # normal CP, last symbol taken by SRS
                                                    (10, (1, 2, 6, 7), (0, 3, 5, 8)),      # extended CP (uci.c:499-546)
                                                    (9, (1, 2, 6, 7), (0, 3, 5, 8))])      # extended CP with SRS
def test_decode_syms_ul_other_column_counts(fec, orc, cols, ack_cols, ri_cols):
    """the UL tile is FE_SEG / cols whole rows of the channel interleaver: every column count LTE has, not only the 12 of the
    bench (rows per tile, run lengths of the scrambling words and the transpose all depend on it)"""
    tbs, mod, prb, n_ack_sym, n_ri_sym, n_cqi = 15264, 4, 50, 8, 4, 24
    rows = 12 * prb
    rng = np.random.default_rng(100 + cols)
    ack_pos = _fe.uci_positions(n_ack_sym, mod, rows, ack_cols)
    ri_pos = _fe.uci_positions(n_ri_sym, mod, rows, ri_cols)
    c_init = synth.cinit_pusch(0x77, 8, 5)
    ntb = 3
    payload = rng.integers(0, 256, (ntb, tbs // 8), dtype=np.uint8)
    sym, G = _fe.make_ul(payload, tbs, mod, 0, prb, cols, c_init, 0.14, rng, ri_pos, n_cqi)
    sb0 = fec.softbuffer_alloc(ntb)
    items = [dict(sym=sym[i], mod=mod, c_init=c_init, nof_symb=cols, ack_pos=ack_pos, ri_pos=ri_pos, n_cqi=n_cqi, nof_e_bits=G, rv=0) for i in range(ntb)]
    d, outs, ucis = _run_syms(fec, tbs, mod, items, True, sb0)
    for i in range(ntb):
        g, raw = _fe.oracle_frontend(orc, sym[i], mod, c_init, None, cols, ack_pos, ri_pos)
        assert (ucis[i][: raw.size] == raw).all(), i
        assert (ucis[i][raw.size: raw.size + n_cqi] == g[:n_cqi]).all(), i
        ot = C.c_void_p(orc.orc_tb_new())
        orc.orc_tb_reset(ot, tbs)
        ret, data, crc, its = _oracle_decode(orc, ot, tbs, mod, 0, g[n_cqi: n_cqi + G])
        orc.orc_tb_free(ot)
        assert d[i].ret == ret and list(d[i].cb_its) == list(its) and list(d[i].cb_crc) == list(crc)
        assert (outs[i][: tbs // 8 + 3] == data[: tbs // 8 + 3]).all()
        assert ret == 0 and (outs[i][: tbs // 8] == payload[i]).all()
