"""GPU parity on the workload shapes of BASELINE.json beyond the golden fixtures: HARQ retransmission batches
(config 3), every code-block size with CRC early stop (config 4), UL-SCH shapes (config 5), and size-independent
properties at the full MCS28 shape (config 2).  Checker: the CPU oracle, bit for bit."""
import ctypes as C
import importlib

import numpy as np
import pytest

from _libs import ALL_K, P

pytestmark = pytest.mark.gpu

pkg = importlib.import_module("empower-srslte_b200")
synth = importlib.import_module("empower-srslte_b200.synth")


@pytest.fixture(scope="module")
def fec():
    f = pkg.FecCuda(0)
    yield f
    f.close()


def _oracle_tb(orc, ot, tbs, Qm, rv, nbits, llr, max_its):
    dd = np.zeros(tbs // 8 + 3 + 1024, np.uint8)
    crc, its = np.zeros(13, np.uint8), np.zeros(13, np.uint32)
    ret = orc.orc_tb_decode(ot, tbs, Qm, rv, nbits, P(np.ascontiguousarray(llr)), max_its, P(dd), P(crc), P(its))
    return ret, dd, crc, its


def _run_harq(fec, orc, tbs, Qm, nbits, sigma, ntb, seed, max_its=8, amp=100.0):
    """ntb HARQ processes, rv sequence 0,2,3,1 until ACK, all processes of a round in ONE batched call"""
    rng = np.random.default_rng(seed)
    payload = rng.integers(0, 256, (ntb, tbs // 8), dtype=np.uint8)
    sb0 = fec.softbuffer_alloc(ntb)
    ots = [C.c_void_p(orc.orc_tb_new()) for _ in range(ntb)]
    for ot in ots:
        orc.orc_tb_reset(ot, tbs)
    live = list(range(ntb))
    stats = dict(acked=0, rounds=0, skipped_cbs=0)
    for txi, rv in enumerate((0, 2, 3, 1)):
        if not live:
            break
        e = synth.encode_tbs(payload[live], tbs, Qm, rv, nbits)
        llr = synth.awgn_llr(e, sigma, rng, amp)
        d = (pkg.TbDesc * len(live))()
        outs = []
        for s, i in enumerate(live):
            out = np.zeros(tbs // 8 + 3 + 768, np.uint8)
            outs.append(out)
            d[s].tbs, d[s].Qm, d[s].rv, d[s].nof_e_bits, d[s].max_its = tbs, Qm, rv, nbits, max_its
            d[s].softbuffer, d[s].new_data = sb0 + i, int(txi == 0)
            d[s].e_bits, d[s].data = llr[s].ctypes.data, out.ctypes.data
        fec.decode_tbs(d)
        pc = fec.last_path_counts()
        stats["trk"] = stats.get("trk", 0) + pc[0]; stats["exact"] = stats.get("exact", 0) + pc[1]
        nxt = []
        for s, i in enumerate(live):
            ret, dd, crc, its = _oracle_tb(orc, ots[i], tbs, Qm, rv, nbits, llr[s], max_its)
            assert d[s].ret == ret, (txi, i)
            assert (outs[s][: tbs // 8 + 3] == dd[: tbs // 8 + 3]).all(), (txi, i)
            assert list(d[s].cb_its) == list(its) and list(d[s].cb_crc) == list(crc), (txi, i)
            stats["skipped_cbs"] += int(sum(1 for c in range(13) if crc[c] and its[c] == 0))
            if ret == 0:
                assert (outs[s][: tbs // 8] == payload[i]).all()
                stats["acked"] += 1
            else:
                nxt.append(i)
        live = nxt
        stats["rounds"] += 1
    for ot in ots:
        orc.orc_tb_free(ot)
    return stats


def test_harq_at_demodulator_llr_scale(fec, orc):
    """the same HARQ flow with LLRs at the scale of the reference's 64QAM demodulator (max|LLR| ~ 800, doubled by soft
    combining): code blocks leave the a-priori range of the packed kernel and run in its certificate variant — launched by
    the host when the hand-over counts come back (first call) and right behind the first kernel afterwards — or in the exact
    kernel; every transmission is still bit-exact against the oracle"""
    st = _run_harq(fec, orc, 75376, 6, 86400, 0.48, ntb=8, seed=5, amp=250.0)
    assert st["rounds"] >= 2 and st["acked"] >= 1, st
    assert st["trk"] > 0, st
    st2 = _run_harq(fec, orc, 6200, 6, 14400, 0.7, ntb=24, seed=6, amp=250.0)
    assert st2["trk"] > 0, st2


def test_config3_harq_retransmissions_tm4_shape(fec, orc):
    """2-codeword TM4 shape: TBS 75376 in 86400 LLRs (rate 0.876), 13 x K=5824; noise chosen so that rv 0 fails for
    most processes and later rvs rescue them: soft combining, cb_crc skipping and the decoded-CB cache are exercised"""
    st = _run_harq(fec, orc, 75376, 6, 86400, 0.48, ntb=6, seed=3)
    assert st["rounds"] >= 2 and st["acked"] >= 1 and st["skipped_cbs"] >= 1, st


def test_config5_ulsch_shapes(fec, orc):
    """UL-SCH: 16QAM / QPSK grants as srsENB schedules them (nbits = 12*12*L_prb*Qm, ra.c:220-225), C from 1 to 6"""
    for tbs, Qm, prb, sigma in ((2216, 2, 10, 0.7), (15264, 4, 40, 0.7), (36696, 4, 96, 0.7)):
        nbits = 12 * 12 * prb * Qm
        st = _run_harq(fec, orc, tbs, Qm, nbits, sigma, ntb=3, seed=tbs)
        assert st["acked"] >= 1, (tbs, st)


def test_config4_all_k_with_crc_early_stop(fec, orc):
    """every LTE code-block size, CRC24B attached, AWGN at the waterfall so that the stopping half-iteration varies;
    one batched call; half-iteration count, CRC verdict and bytes against the oracle's per-half-iteration outputs"""
    rng = np.random.default_rng(44)
    Ks, llrs, want = [], [], []
    for idx, K in enumerate(ALL_K):
        info = rng.integers(0, 256, (1, (K - 24) // 8), dtype=np.uint8)
        crc = synth.crc24(synth.CRC24B, info)
        blk = np.concatenate([info, np.array([[(crc[0] >> 16) & 255, (crc[0] >> 8) & 255, crc[0] & 255]], np.uint8)], axis=1)
        coded = synth.turbo_encode(np.unpackbits(blk, axis=1))
        sigma = (0.75, 0.85, 0.95, 1.05)[idx % 4]
        llr = synth.awgn_llr(coded, sigma, rng)[0]
        ob = np.zeros((8, K // 8), np.uint8)
        assert orc.orc_tdec_halfits(K, P(llr), 8, P(ob), None) == 0
        n = next((i + 1 for i in range(8) if orc.orc_crc(1, P(np.ascontiguousarray(ob[i])), K) == 0), 8)
        Ks.append(K); llrs.append(llr); want.append((n, ob[n - 1], int(orc.orc_crc(1, P(np.ascontiguousarray(ob[n - 1])), K) == 0)))
    res = fec.decode_cbs(Ks, llrs, 8, crc_type=1, early_stop=True)
    bad = [(K, r["its"], w[0]) for K, r, w in zip(Ks, res, want) if r["its"] != w[0] or r["crc_ok"] != w[2] or not (r["out"] == w[1]).all()]
    assert not bad, bad[:10]
    its = [w[0] for w in want]
    assert min(its) <= 2 and max(its) >= 6, "the sweep should cover early and late stops"


def test_config2_full_shape_vs_oracle(fec, orc):
    """MCS28 / 100 PRB (TBS 75376, 90000 LLRs, 13 x K=5824) on 148 transport blocks, every one of them against the oracle's
    decode_tb (return value, per-CB CRC flags, per-CB half-iteration counts, all TBS/8+3 bytes); and: every CRC-passing
    block returns the transmitted payload, the call is deterministic, host and device buffers give identical results"""
    import torch
    tbs, Qm, nbits, n = 75376, 6, 90000, 148
    rng = np.random.default_rng(2)
    payload = rng.integers(0, 256, (n, tbs // 8), dtype=np.uint8)
    llr = synth.awgn_llr(synth.encode_tbs(payload, tbs, Qm, 0, nbits), 0.47, rng)
    sb0 = fec.softbuffer_alloc(n)

    crcs = []

    def run(dev):
        d = (pkg.TbDesc * n)()
        if dev:
            dl = torch.from_numpy(llr).cuda()
            do = torch.zeros((n, tbs // 8 + 3 + 768), dtype=torch.uint8, device="cuda")
        else:
            do = np.zeros((n, tbs // 8 + 3 + 768), np.uint8)
        for i in range(n):
            d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its, d[i].softbuffer, d[i].new_data = tbs, Qm, 0, nbits, 8, sb0 + i, 1
            d[i].e_bits = dl[i].data_ptr() if dev else llr[i].ctypes.data
            d[i].data = do[i].data_ptr() if dev else do[i].ctypes.data
        fec.decode_tbs(d, pkg.F_DEVICE_IO if dev else 0)
        out = do.cpu().numpy() if dev else do
        crcs.append([list(d[i].cb_crc) for i in range(n)])
        return [d[i].ret for i in range(n)], [list(d[i].cb_its) for i in range(n)], out

    r1, i1, o1 = run(False)
    r2, i2, o2 = run(False)
    r3, i3, o3 = run(True)
    assert r1 == r2 == r3 and i1 == i2 == i3
    ot = C.c_void_p(orc.orc_tb_new())
    bad = []
    for i in range(n):
        orc.orc_tb_reset(ot, tbs)
        ret, dd, crc, its = _oracle_tb(orc, ot, tbs, Qm, 0, nbits, llr[i], 8)
        if r1[i] != ret or i1[i] != list(its) or crcs[0][i] != list(crc) or not (o1[i, : tbs // 8 + 3] == dd[: tbs // 8 + 3]).all():
            bad.append((i, r1[i], ret, i1[i], list(its)))
    orc.orc_tb_free(ot)
    assert not bad, bad[:5]
    ok = [i for i in range(n) if r1[i] == 0]
    assert 0.5 * n < len(ok), "operating point should decode most blocks"
    for i in ok:
        assert (o1[i, : tbs // 8] == payload[i]).all()
    assert (o1[:, : tbs // 8 + 3] == o2[:, : tbs // 8 + 3]).all() and (o1[:, : tbs // 8 + 3] == o3[:, : tbs // 8 + 3]).all()
    spread = {v for row in i1 for v in row[:13]}
    assert len(spread) >= 3, "half-iteration counts should vary across code blocks (early stop)"


def test_device_call_cut_into_two_sub_batches_and_calls_in_flight(orc):
    """a device-resident call is cut into two sub-batches of at least 592 transport blocks (csrc/fec_cuda_api.cu submit_common):
    1301 small transport blocks in one call must give what the host-buffer call (sub-batches of 296) gives and what the
    oracle gives; then four calls in flight on a fresh context (every pipe of the ring is used before its first wait)"""
    import torch
    tbs, Qm, nbits, n = 1000, 2, 1440, 1301
    rng = np.random.default_rng(31)
    payload = rng.integers(0, 256, (n, tbs // 8), dtype=np.uint8)
    llr = synth.awgn_llr(synth.encode_tbs(payload, tbs, Qm, 0, nbits), 0.62, rng)
    f = pkg.FecCuda(0, max_tbs=64)                      # deliberately small: the staging grows with the first call
    try:
        sb0 = f.softbuffer_alloc(4 * n)
        dl = torch.from_numpy(llr).cuda()
        stride = tbs // 8 + 3 + 768

        def descs(dev, out, sb):
            d = (pkg.TbDesc * n)()
            for i in range(n):
                d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its, d[i].softbuffer, d[i].new_data = tbs, Qm, 0, nbits, 8, sb + i, 1
                d[i].e_bits = dl[i].data_ptr() if dev else llr[i].ctypes.data
                d[i].data = out[i].data_ptr() if dev else out[i].ctypes.data
            return d

        oh = np.zeros((n, stride), np.uint8)
        dh = descs(False, oh, sb0)
        f.decode_tbs(dh, 0)
        od = torch.zeros((n, stride), dtype=torch.uint8, device="cuda")
        dd = descs(True, od, sb0 + n)
        f.decode_tbs(dd, pkg.F_DEVICE_IO)
        o2 = od.cpu().numpy()
        assert [dh[i].ret for i in range(n)] == [dd[i].ret for i in range(n)]
        assert [list(dh[i].cb_its) for i in range(n)] == [list(dd[i].cb_its) for i in range(n)]
        assert (oh[:, : tbs // 8 + 3] == o2[:, : tbs // 8 + 3]).all()
        ot = C.c_void_p(orc.orc_tb_new())
        for i in list(range(0, 8)) + list(range(646, 656)) + [n - 1]:      # both sides of the cut
            orc.orc_tb_reset(ot, tbs)
            ret, data, crc, its = _oracle_tb(orc, ot, tbs, Qm, 0, nbits, llr[i], 8)
            assert dd[i].ret == ret and list(dd[i].cb_its) == list(its), i
            assert (o2[i, : tbs // 8 + 3] == data[: tbs // 8 + 3]).all(), i
        orc.orc_tb_free(ot)
        assert sum(1 for i in range(n) if dd[i].ret == 0) > n // 2
    finally:
        f.close()
    g = pkg.FecCuda(0, max_tbs=64)
    try:
        sb0 = g.softbuffer_alloc(4 * n)
        outs = [torch.zeros((n, stride), dtype=torch.uint8, device="cuda") for _ in range(4)]
        sets = []
        for k in range(4):
            d = (pkg.TbDesc * n)()
            for i in range(n):
                d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its, d[i].softbuffer, d[i].new_data = tbs, Qm, 0, nbits, 8, sb0 + k * n + i, 1
                d[i].e_bits, d[i].data = dl[i].data_ptr(), outs[k][i].data_ptr()
            sets.append(d)
        tickets = [g.submit_tbs(d, pkg.F_DEVICE_IO) for d in sets]
        for t in tickets:
            g.wait(t)
        for k in range(4):
            assert [sets[k][i].ret for i in range(n)] == [dd[i].ret for i in range(n)], k
            assert (outs[k].cpu().numpy()[:, : tbs // 8 + 3] == o2[:, : tbs // 8 + 3]).all(), k
    finally:
        g.close()


def test_tb_path_exact_kernel_deferred_then_eager(orc):
    """LLR amplitudes beyond the reach of both packed kernels (max|soft| > 2250): every code block ends up in the exact
    kernel. On a fresh context the exact kernel is not launched with the sub-batch (nothing needed it recently): the host
    launches it when the hand-over counts come back, then eagerly for the following calls. Both ways must give the oracle's
    bits, iteration counts and return codes; a last call with ordinary amplitudes takes the fast path again."""
    tbs, Qm, nbits, n = 6200, 2, 7200, 6
    rng = np.random.default_rng(77)
    payload = rng.integers(0, 256, (n, tbs // 8), dtype=np.uint8)
    e = synth.encode_tbs(payload, tbs, Qm, 0, nbits)
    f = pkg.FecCuda(0, max_tbs=16)
    try:
        sb0 = f.softbuffer_alloc(n)
        ot = C.c_void_p(orc.orc_tb_new())
        for call, amp in enumerate((3000.0, 3000.0, 2600.0, 100.0)):
            llr = synth.awgn_llr(e, 0.8, rng, amp=amp)
            out = np.zeros((n, tbs // 8 + 3 + 768), np.uint8)
            d = (pkg.TbDesc * n)()
            for i in range(n):
                d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its, d[i].softbuffer, d[i].new_data = tbs, Qm, 0, nbits, 8, sb0 + i, 1
                d[i].e_bits, d[i].data = llr[i].ctypes.data, out[i].ctypes.data
            f.decode_tbs(d, 0)
            paths = f.last_path_counts()
            if amp > 2500:
                assert paths[1] > 0, (call, paths)          # the exact kernel did the work
            else:
                assert paths == (0, 0) or list(paths) == [0, 0], (call, paths)
            for i in range(n):
                orc.orc_tb_reset(ot, tbs)
                ret, dd, crc, its = _oracle_tb(orc, ot, tbs, Qm, 0, nbits, llr[i], 8)
                assert d[i].ret == ret and list(d[i].cb_its) == list(its) and list(d[i].cb_crc) == list(crc), (call, i)
                assert (out[i, : tbs // 8 + 3] == dd[: tbs // 8 + 3]).all(), (call, i)
        orc.orc_tb_free(ot)
    finally:
        f.close()
