"""GPU parity tests proper: the CUDA path through the C ABI vs the golden vectors (generated from the
unmodified reference) and vs the CPU oracle on seeded inputs. Bit-exact: decoded bytes, per-CB CRC flags,
per-CB half-iteration counts, soft-buffer contents."""
import ctypes as C
import importlib

import numpy as np
import pytest

import _golden as G
from _libs import ALL_K, P, int_awgn_llr, int_gauss

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fec():
    pkg = importlib.import_module("empower-srslte_b200")
    f = pkg.FecCuda(0)
    yield f
    f.close()


@pytest.mark.parametrize("case", list(G.tdec_cases()), ids=lambda c: c["id"])
def test_tdec_golden(fec, case):
    """srslte_tdec_run_all shape: exact half-iteration counts, no early stop; bytes after every count"""
    K = case["K"]
    for nof in range(1, case["nof"] + 1):
        r = fec.decode_cbs([K], [case["inp"]], nof, early_stop=False)[0]
        assert r["its"] == nof
        assert (r["out"] == case["bytes"][nof - 1]).all(), (K, nof)


def test_tdec_all_k_vs_oracle(fec, orc):
    """every one of the 188 LTE sizes, one batch (exercises per-K grouping and all three decoder classes)"""
    Ks, llrs, want = [], [], []
    for i, K in enumerate(ALL_K):
        inp = (int_gauss(3 * K + 12, 7000 + i) % 1201 - 600).astype(np.int16)
        ob = np.zeros((5, K // 8), np.uint8)
        assert orc.orc_tdec_halfits(K, P(inp), 5, P(ob), None) == 0
        Ks.append(K); llrs.append(inp); want.append(ob[4])
    res = fec.decode_cbs(Ks, llrs, 5, early_stop=False)
    bad = [K for K, r, w in zip(Ks, res, want) if not (r["out"] == w).all()]
    assert not bad, bad


def test_tdec_saturation_vs_oracle(fec, orc):
    """amplitudes that drive the saturating adds to the rails"""
    for K in (408, 800, 816, 2048, 6144):
        for amp in (6000, 30000):
            inp = (int_gauss(3 * K + 12, K + amp) % (2 * amp + 1) - amp).astype(np.int16)
            ob = np.zeros((8, K // 8), np.uint8)
            orc.orc_tdec_halfits(K, P(inp), 8, P(ob), None)
            for nof in (1, 2, 5, 8):
                r = fec.decode_cbs([K], [inp], nof, early_stop=False)[0]
                assert (r["out"] == ob[nof - 1]).all(), (K, amp, nof)


def test_tdec_at_range_bound_vs_oracle(fec, orc):
    """inputs sitting exactly on the packed kernel's range condition (max|soft| = GMAX/2 = 1000, adversarial
    sign patterns): the fast path must still be bit-exact; later half-iterations exercise the hand-over"""
    Ks, llrs, want = [], [], []
    for K in (408, 800, 816, 1008, 5824, 6144):
        n = 3 * K + 12
        r = np.random.default_rng(K)
        pats = [np.full(n, 1000), np.full(n, -1000), np.where(np.arange(n) % 2 == 0, 1000, -1000),
                np.where((np.arange(n) // 3) % 2 == 0, 1000, -1000), 1000 * (r.integers(0, 2, n) * 2 - 1),
                np.where(np.arange(n) % 3 == 0, 1000, -1000), r.integers(-1000, 1001, n), 500 * (r.integers(0, 2, n) * 2 - 1)]
        for pt in pats:
            inp = np.ascontiguousarray(pt, dtype=np.int16)
            ob = np.zeros((3, K // 8), np.uint8)
            assert orc.orc_tdec_halfits(K, P(inp), 3, P(ob), None) == 0
            for nof in (1, 2, 3):
                Ks.append(K); llrs.append(inp); want.append((nof, ob[nof - 1].copy()))
    for nof in (1, 2, 3):
        sel = [i for i, w in enumerate(want) if w[0] == nof]
        res = fec.decode_cbs([Ks[i] for i in sel], [llrs[i] for i in sel], nof, early_stop=False)
        bad = [(Ks[i], nof) for i, r in zip(sel, res) if not (r["out"] == want[i][1]).all()]
        assert not bad, bad


def _noisy_codeword(K, amp, sigma, seed):
    """natural-layout int16 LLRs [s p p'] x K + 12 tail of a random turbo codeword (CRC24B attached), bit 1 -> +amp"""
    synth = importlib.import_module("empower-srslte_b200.synth")
    rng = np.random.default_rng(seed)
    body = rng.integers(0, 256, (1, K // 8 - 3), dtype=np.uint8)
    blk = np.concatenate([body, synth._crc_bytes(synth.crc24(synth.CRC24B, body))], axis=1)
    coded = synth.turbo_encode(np.unpackbits(blk, axis=1))[0]
    x = amp * ((2.0 * coded - 1.0) + sigma * rng.standard_normal(coded.shape))
    return np.ascontiguousarray(np.clip(np.rint(x), -32768, 32767).astype(np.int16))


def test_tdec_beyond_apriori_range_vs_oracle(fec, orc):
    """LLR scales of the reference's own demodulator and above (max|soft| 600 .. 2250, i.e. G up to 4500 and beyond): the
    code blocks leave the a-priori range of the packed kernel and run in its certificate variant (reference normalisation
    schedule + run-time range certificate) or, when the certificate cannot cover them, in the exact kernel. Decisions
    of every half-iteration must be bit-exact either way; the certificate variant must actually carry the realistic
    cases (few hand-overs to the exact kernel)."""
    synth = importlib.import_module("empower-srslte_b200.synth")
    cases = []
    for K in (408, 800, 816, 1008, 1056, 2048, 5824, 6144):       # P=8 (L=51, 100), P=16 with L odd / = 2 mod 4 / = 0 mod 4
        n = 3 * K + 12
        r = np.random.default_rng(K + 7)
        for amp in (1100, 1500, 2250):                              # adversarial constant-magnitude patterns
            cases.append((K, np.where(np.arange(n) % 2 == 0, amp, -amp).astype(np.int16), "alt%d" % amp))
            cases.append((K, (amp * (r.integers(0, 2, n) * 2 - 1)).astype(np.int16), "rnd%d" % amp))
            cases.append((K, r.integers(-amp, amp + 1, n).astype(np.int16), "uni%d" % amp))
        # noisy codewords at the scale of the reference's 64QAM demodulator (max|LLR| ~ 700..1300), several half-iterations needed
        for amp, sigma in ((150, 0.9), (200, 0.8), (250, 0.75), (250, 1.0)):
            cases.append((K, _noisy_codeword(K, amp, sigma, K + amp), "cw%d_%g" % (amp, sigma)))
    NH = 8
    want = []
    for K, inp, _ in cases:
        ob = np.zeros((NH, K // 8), np.uint8)
        assert orc.orc_tdec_halfits(K, P(inp), NH, P(ob), None) == 0
        want.append(ob)
    tot_trk = tot_exact = 0
    for nof in (1, 2, 3, 4, 6, 8):
        res = fec.decode_cbs([c[0] for c in cases], [c[1] for c in cases], nof, early_stop=False)
        bad = [(c[0], c[2], nof) for c, r, w in zip(cases, res, want) if not (r["out"] == w[nof - 1]).all()]
        assert not bad, bad
        t, e = fec.last_path_counts()
        tot_trk += t; tot_exact += e
    assert tot_trk > 0, "the certificate kernel never ran"
    # realistic scales with CRC early stop (as decode_tb_cb runs them): the certificate variant must carry (nearly) all of them
    idx = [i for i, c in enumerate(cases) if c[2].startswith("cw")]
    exp_its = []
    for i in idx:
        K = cases[i][0]
        ok = [n + 1 for n in range(NH) if orc.orc_crc(1, P(np.ascontiguousarray(want[i][n])), K) == 0]
        exp_its.append(ok[0] if ok else NH)
    res = fec.decode_cbs([cases[i][0] for i in idx], [cases[i][1] for i in idx], NH, crc_type=1, early_stop=True)
    for i, r, n in zip(idx, res, exp_its):
        assert r["its"] == n and (r["out"] == want[i][n - 1]).all(), (cases[i][0], cases[i][2], r["its"], n)
    t, e = fec.last_path_counts()
    print("certificate kernel: %d of %d realistic code blocks (half-iterations %s), exact kernel: %d" % (t, len(idx), exp_its, e))
    assert t >= len(idx) // 2 and e <= len(idx) // 4, (t, e, len(idx))


def test_tdec_continues_in_certificate_kernel_vs_oracle(fec, orc):
    """code blocks that START inside the a-priori range of the packed kernel (max|soft| <= 800) and leave it in mid-decode
    because their extrinsics grow: the certificate kernel goes on from the dumped extrinsic array and half-iteration count
    instead of restarting. The decisions of every half-iteration count must equal the oracle's, with and without CRC
    early stop; every hand-over counted here is a mid-decode one (none of these blocks is routed there at refill)."""
    cases = []
    for K in (408, 800, 816, 1008, 1056, 2048, 5824, 6144):
        for amp, sigma, clip in ((160, 1.0, 800), (220, 0.95, 800), (120, 1.05, 640), (250, 0.8, 750)):
            x = _noisy_codeword(K, amp, sigma, 31 * K + amp)
            cases.append((K, np.ascontiguousarray(np.clip(x, -clip, clip)), "cw%d_%g" % (amp, sigma)))
    assert all(int(np.abs(c[1].astype(np.int32)).max()) <= 800 for c in cases)
    NH = 10
    want = []
    for K, inp, _ in cases:
        ob = np.zeros((NH, K // 8), np.uint8)
        assert orc.orc_tdec_halfits(K, P(inp), NH, P(ob), None) == 0
        want.append(ob)
    tot_trk = 0
    for nof in (2, 3, 4, 5, 7, 10):
        res = fec.decode_cbs([c[0] for c in cases], [c[1] for c in cases], nof, early_stop=False)
        bad = [(c[0], c[2], nof) for c, r, w in zip(cases, res, want) if not (r["out"] == w[nof - 1]).all()]
        assert not bad, bad
        t, e = fec.last_path_counts()
        tot_trk += t
    assert tot_trk >= len(cases), "no code block was handed over in mid-decode (%d)" % tot_trk
    exp_its = []
    for i, (K, _, _) in enumerate(cases):
        ok = [n + 1 for n in range(NH) if orc.orc_crc(1, P(np.ascontiguousarray(want[i][n])), K) == 0]
        exp_its.append(ok[0] if ok else NH)
    res = fec.decode_cbs([c[0] for c in cases], [c[1] for c in cases], NH, crc_type=1, early_stop=True)
    for c, r, w, n in zip(cases, res, want, exp_its):
        assert r["its"] == n and (r["out"] == w[n - 1]).all(), (c[0], c[2], r["its"], n)
    print("mid-decode hand-overs: %d over 6 runs of %d blocks; early-stop half-iterations %s" % (tot_trk, len(cases), exp_its))


def test_rm_rx_golden_all_rows(fec):
    """every row of the fixture the unmodified reference produced (oracle/gen_golden.py:71-81): 188 K x 4 rv x {natural,
    sub-block layout} x 3 lengths E (below, at and above 3K+12). The per-K rank table is where a gather breaks (ND, R, the
    wrap at k0), so all of them — the scope of rm_turbo_test.c:140-193. The buffer is exactly the extent the layout uses,
    between canaries (the reference's own test allocates 3K+12 entries, rm_turbo_test.c:52,186)."""
    z = G.tables()
    assert len(z["rm"]) == 188 * 4 * 2 * 3
    bad = []
    for K, rv, sb, E, c in z["rm"]:
        K, rv, sb, E = int(K), int(rv), int(sb), int(E)
        e = (int_gauss(E, K * 8 + rv) % 2001 - 1000).astype(np.int16)
        s = (int_gauss(18600, K) % 20001 - 10000).astype(np.int16)
        ext = 3 * (K + 32) + 12 if sb and K > 400 else 3 * K + 12
        buf = np.full(ext + 128, 0x5A5A, np.int16)
        buf[64:64 + ext] = s[:ext]
        rc = fec.lib.fec_cuda_rm_rx(fec.h, P(e), E, ALL_K.index(K), rv, sb, C.c_void_p(buf.ctypes.data + 128))
        assert rc == 0
        assert (buf[:64] == 0x5A5A).all() and (buf[64 + ext:] == 0x5A5A).all(), ("wrote outside the extent", K, rv, sb, E)
        s[:ext] = buf[64:64 + ext]
        if G.crc32(s) != int(c):
            bad.append((K, rv, sb, E))
    assert not bad, bad[:20]


def test_rm_rx_invalid(fec):
    s = np.zeros(18600, np.int16)
    assert fec.rm_rx(np.zeros(10, np.int16), 188, 0, s) == -2
    assert fec.rm_rx(np.zeros(10, np.int16), 0, 4, s) == -2


@pytest.mark.parametrize("case", list(G.tb_cases()), ids=lambda c: c["id"])
def test_tb_golden(fec, case):
    tbs = case["tbs"]
    sb = fec.softbuffer_alloc(1)
    pkg = importlib.import_module("empower-srslte_b200")
    seg = pkg.cbsegm(tbs)
    for i, tx in enumerate(case["tx"]):
        r = fec.decode_tb(tbs, case["Qm"], tx["rv"], tx["llr"], sb, new_data=(i == 0), max_its=case["max_its"])
        assert r["ret"] == tx["ret"]
        assert (r["cb_crc"] == tx["crc"][: seg["C"]]).all()
        assert (r["cb_its"] == tx["its"][: seg["C"]]).all(), (r["cb_its"], tx["its"])
        assert (r["data"] == tx["data"]).all()
        for cb in range(seg["C"]):
            K = seg["K1"] if cb < seg["C1"] else seg["K2"]
            if tx["its"][cb] == 0:
                continue      # CB skipped (already OK): the reference leaves its soft buffer untouched too
            s = fec.softbuffer_read(sb, cb, K)
            assert G.soft_crc_ref_layout(s, K, pkg.load().fec_cuda_subblocks(K)) == int(tx["softcrc"][cb]), (i, cb)
        if r["ret"] == 0:
            assert (r["data"][: tbs // 8] == case["tbdata"]).all()


def test_tb_batch_vs_oracle(fec, orc):
    """a mixed batch (several TBS / HARQ processes in one call) against the CPU oracle"""
    pkg = importlib.import_module("empower-srslte_b200")
    cases = [c for c in G.tb_cases()]
    n = len(cases)
    sb0 = fec.softbuffer_alloc(n)
    ots = [C.c_void_p(orc.orc_tb_new()) for _ in cases]
    for c, ot in zip(cases, ots):
        orc.orc_tb_reset(ot, c["tbs"])
    maxtx = max(len(c["tx"]) for c in cases)
    for txi in range(maxtx):
        live = [j for j, c in enumerate(cases) if txi < len(c["tx"])]
        d = (pkg.TbDesc * len(live))()
        outs = []
        for slot, j in enumerate(live):
            c, tx = cases[j], cases[j]["tx"][txi]
            out = np.zeros(c["tbs"] // 8 + 3 + 768, np.uint8)
            outs.append(out)
            d[slot].tbs, d[slot].Qm, d[slot].rv, d[slot].nof_e_bits, d[slot].max_its = c["tbs"], c["Qm"], tx["rv"], c["nbits"], c["max_its"]
            d[slot].softbuffer, d[slot].new_data = sb0 + j, int(txi == 0)
            d[slot].e_bits, d[slot].data = tx["llr"].ctypes.data, out.ctypes.data
        fec.decode_tbs(d)
        for slot, j in enumerate(live):
            c, tx = cases[j], cases[j]["tx"][txi]
            dd = np.zeros(c["tbs"] // 8 + 3 + 1024, np.uint8)
            crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
            ret = orc.orc_tb_decode(ots[j], c["tbs"], c["Qm"], tx["rv"], c["nbits"], P(tx["llr"]), c["max_its"], P(dd), P(crc), P(its))
            assert d[slot].ret == ret
            assert (outs[slot][: c["tbs"] // 8 + 3] == dd[: c["tbs"] // 8 + 3]).all()
            assert list(d[slot].cb_its) == list(its) and list(d[slot].cb_crc) == list(crc)


def test_invalid_inputs(fec):
    """decode_tb argument checks (sch.c:444-457): filler bits / unknown soft buffer / null pointers -> -2"""
    pkg = importlib.import_module("empower-srslte_b200")
    sb = fec.softbuffer_alloc(1)
    e = np.zeros(1000, np.int16)
    r = fec.decode_tb(0, 2, 0, e, sb, True)
    assert r["ret"] == 0                                   # tbs == 0: nothing to do, success
    r = fec.decode_tb(1000, 2, 0, e, sb + 12345, True)
    assert r["ret"] == -2
    r = fec.decode_tb(6120 + 8, 2, 0, np.zeros(20000, np.int16), sb, True)   # B'=6176 -> F != 0 not supported
    seg = pkg.cbsegm(6128)
    assert (seg["F"] != 0) == (r["ret"] == -2)


def test_submit_wait_two_batches_in_flight(fec):
    """fec_cuda_submit_tbs / fec_cuda_wait: two batches in flight (own soft buffers) give the golden results"""
    pkg = importlib.import_module("empower-srslte_b200")
    cases = list(G.tb_cases())
    reps = 40                                   # enough transport blocks per batch to span several sub-batches
    batches = []
    for half in range(2):
        mine = [c for i, c in enumerate(cases) if i % 2 == half] * reps
        sb0 = fec.softbuffer_alloc(len(mine))
        d = (pkg.TbDesc * len(mine))()
        outs = []
        for slot, c in enumerate(mine):
            tx = c["tx"][0]
            out = np.zeros(c["tbs"] // 8 + 3 + 768, np.uint8)
            outs.append(out)
            d[slot].tbs, d[slot].Qm, d[slot].rv, d[slot].nof_e_bits, d[slot].max_its = c["tbs"], c["Qm"], tx["rv"], c["nbits"], c["max_its"]
            d[slot].softbuffer, d[slot].new_data = sb0 + slot, 1
            d[slot].e_bits, d[slot].data = tx["llr"].ctypes.data, out.ctypes.data
        batches.append((mine, d, outs))
    t0 = fec.submit_tbs(batches[0][1])
    t1 = fec.submit_tbs(batches[1][1])
    assert t1 == t0 + 1
    fec.wait(t0)
    fec.wait(t1)
    for mine, d, outs in batches:
        for slot, c in enumerate(mine):
            tx = c["tx"][0]
            assert d[slot].ret == tx["ret"]
            assert (outs[slot][: c["tbs"] // 8 + 3] == tx["data"][: c["tbs"] // 8 + 3]).all()
            assert list(d[slot].cb_its)[: len(tx["its"])] == [int(v) for v in tx["its"]][:13]
