"""The reference's own entry points (include/srslte_fec_compat.h) on the GPU: srslte_tdec_*, srslte_rm_turbo_rx_lut[_],
srslte_softbuffer_rx_* and the decode_tb body, called the way the reference's tests call them
(turbodecoder_test.c:181-330, rm_turbo_test.c, pdsch_test.c) and compared bit for bit with the golden vectors
generated from the unmodified reference and with the CPU oracle."""
import ctypes as C
import importlib

import numpy as np
import pytest

import _golden as G
from _libs import ALL_K, P, int_gauss

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    pkg = importlib.import_module("empower-srslte_b200")
    l = pkg.load()
    l.srslte_tdec_init.argtypes = [C.c_void_p, C.c_uint32]
    l.srslte_tdec_free.argtypes = [C.c_void_p]
    l.srslte_tdec_force_not_sb.argtypes = [C.c_void_p]
    l.srslte_tdec_new_cb.argtypes = [C.c_void_p, C.c_uint32]
    l.srslte_tdec_get_nof_iterations.argtypes = [C.c_void_p]
    l.srslte_tdec_iteration.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    l.srslte_tdec_iteration.restype = None
    l.srslte_tdec_run_all.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    l.srslte_rm_turbo_rx_lut_.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_bool]
    l.srslte_rm_turbo_rx_lut.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
    l.srslte_softbuffer_rx_init.argtypes = [C.c_void_p, C.c_uint32]
    l.srslte_softbuffer_rx_free.argtypes = [C.c_void_p]
    l.srslte_softbuffer_rx_reset_tbs.argtypes = [C.c_void_p, C.c_uint32]
    l.srslte_sch_decode_tb_cuda.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                            C.c_uint32, C.c_void_p]
    l.srslte_tdec_iteration_8bit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    l.srslte_tdec_iteration_8bit.restype = None
    l.srslte_tdec_run_all_8bit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    l.srslte_sch_decode_tb_cuda_8bit.argtypes = l.srslte_sch_decode_tb_cuda.argtypes
    l.srslte_rm_turbo_rx_lut_8bit.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
    return l


def _tdec(lib, max_k=6144):
    h = C.create_string_buffer(18264)          # sizeof(srslte_tdec_t), caller-allocated as in the reference
    assert lib.srslte_tdec_init(h, max_k) == 0
    return h


@pytest.mark.parametrize("case", list(G.tdec_cases()), ids=lambda c: c["id"])
def test_tdec_run_all_and_iteration_golden(lib, case):
    """turbodecoder_test.c drives srslte_tdec_run_all on natural-layout input with force_not_sb; the golden bytes after
    every half-iteration come from the reference's srslte_tdec_iteration"""
    K = case["K"]
    h = _tdec(lib)
    lib.srslte_tdec_force_not_sb(h)
    inp = np.ascontiguousarray(case["inp"])
    out = np.zeros(K // 8, np.uint8)
    assert lib.srslte_tdec_run_all(h, P(inp), P(out), case["nof"], K) == 0
    assert lib.srslte_tdec_get_nof_iterations(h) == case["nof"]
    assert (out == case["bytes"][case["nof"] - 1]).all()
    assert lib.srslte_tdec_new_cb(h, K) == 0
    for n in range(case["nof"]):
        lib.srslte_tdec_iteration(h, P(inp), P(out))
        assert lib.srslte_tdec_get_nof_iterations(h) == n + 1
        assert (out == case["bytes"][n]).all(), (K, n)
    lib.srslte_tdec_free(h)


def test_tdec_sb_layout_input_vs_oracle(lib, orc):
    """without force_not_sb the decoder expects the sub-block layout srslte_rm_turbo_rx_lut_(.., true) produces
    (turbodecoder_iter.h:61-70): rate de-match into a soft buffer, decode from it"""
    for K in (408, 1024, 5824, 6144):
        E = 3 * K + 12 + 600
        e = (int_gauss(E, 31 * K) % 1201 - 600).astype(np.int16)
        soft = np.zeros(18600, np.int16)
        assert lib.srslte_rm_turbo_rx_lut_(P(e), P(soft), E, ALL_K.index(K), 0, True) == 0
        want_soft = np.zeros(18600, np.int16)
        assert orc.orc_rm_rx(P(e), E, K, 0, 1, P(want_soft)) == 0
        assert (soft == want_soft).all()
        want = np.zeros(K // 8, np.uint8)
        nits = C.c_int(0)
        orc.orc_decode_cb(P(want_soft), K, 4, 1, P(want), C.byref(nits))
        h = _tdec(lib)
        out = np.zeros(K // 8, np.uint8)
        assert lib.srslte_tdec_new_cb(h, K) == 0
        for _ in range(nits.value):
            lib.srslte_tdec_iteration(h, P(soft), P(out))
        assert (out == want).all(), K
        lib.srslte_tdec_free(h)


def test_tdec_argument_checks(lib):
    h = _tdec(lib, 1024)
    assert lib.srslte_tdec_new_cb(h, 2048) == -1        # larger than max_long_cb (turbodecoder.c:495-499)
    assert lib.srslte_tdec_new_cb(h, 1000) == -1        # not an LTE size
    assert lib.srslte_tdec_new_cb(h, 1024) == 0
    lib.srslte_tdec_free(h)


def test_rm_rx_lut_accumulates_like_reference(lib, orc):
    """two transmissions (rv 0 then rv 2) into one caller-owned soft buffer, natural and sub-block layouts"""
    for K, sb in ((40, False), (512, False), (512, True), (6144, True)):
        soft = np.zeros(18600, np.int16)
        want = np.zeros(18600, np.int16)
        for rv in (0, 2):
            E = 2 * K + 7
            e = (int_gauss(E, K + rv) % 30001 - 15000).astype(np.int16)        # large: wrapping adds are exercised
            assert lib.srslte_rm_turbo_rx_lut_(P(e), P(soft), E, ALL_K.index(K), rv, sb) == 0
            assert orc.orc_rm_rx(P(e), E, K, rv, int(sb), P(want)) == 0
        assert (soft == want).all(), (K, sb)
    assert lib.srslte_rm_turbo_rx_lut(P(e), P(soft), 10, 188, 0) == -2
    assert lib.srslte_rm_turbo_rx_lut(P(e), P(soft), 10, 0, 4) == -2


@pytest.mark.parametrize("case", list(G.tb_cases()), ids=lambda c: c["id"])
def test_decode_tb_with_reference_softbuffer_calls(lib, case):
    """pdsch_test.c shape: srslte_softbuffer_rx_init / reset_tbs, then one decode_tb per (re)transmission"""
    tbs = case["tbs"]
    nC = importlib.import_module("empower-srslte_b200").cbsegm(tbs)["C"]
    q = C.create_string_buffer(40)             # sizeof(srslte_softbuffer_rx_t)
    assert lib.srslte_softbuffer_rx_init(q, 100) == 0
    assert C.c_uint32.from_buffer(q, 0).value == 13
    lib.srslte_softbuffer_rx_reset_tbs(q, tbs)
    for tx in case["tx"]:
        data = np.zeros(tbs // 8 + 3 + 768, np.uint8)
        noi = C.c_uint32(0)
        llr = np.ascontiguousarray(tx["llr"])
        ret = lib.srslte_sch_decode_tb_cuda(q, tbs, case["Qm"], tx["rv"], case["nbits"], P(llr), P(data), case["max_its"], C.byref(noi))
        assert ret == tx["ret"]
        assert (data[: tbs // 8 + 3] == tx["data"][: tbs // 8 + 3]).all()
        crc = (C.c_bool * 13).from_address(C.c_void_p.from_buffer(q, 24).value)
        assert [bool(crc[i]) for i in range(nC)] == [bool(v) for v in tx["crc"][:nC]]
    lib.srslte_softbuffer_rx_free(q)
    assert bytes(q.raw) == bytes(40)


def test_encode_tb_wrapper_golden(lib):
    """the transmit-side body a patched sch.c calls: the packed bits srslte_dlsch_encode emitted (golden fixtures)"""
    import os
    z = np.load(os.path.join(G.GOLD, "tb.npz"))
    lib.srslte_sch_encode_tb_cuda.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
    for ci, c in enumerate(G.tb_cases()):
        data = np.ascontiguousarray(c["tbdata"][: c["tbs"] // 8])
        for tx, t in enumerate(c["tx"]):
            e = np.zeros((c["nbits"] + 7) // 8, np.uint8)
            assert lib.srslte_sch_encode_tb_cuda(c["tbs"], c["Qm"], t["rv"], c["nbits"], P(data), P(e)) == 0
            assert (np.unpackbits(e)[: c["nbits"]] == np.unpackbits(z[f"e_{ci}_{tx}"])[: c["nbits"]]).all(), (c["id"], tx)
    assert lib.srslte_sch_encode_tb_cuda(75376, 6, 0, 90000, None, P(np.zeros(8, np.uint8))) == -2


@pytest.mark.parametrize("case", list(G.tdec8_cases()), ids=lambda c: c["id"])
def test_tdec_8bit_run_all_and_iteration_golden(lib, case):
    """srslte_tdec_run_all_8bit / srslte_tdec_iteration_8bit (turbodecoder.c:536-563) on natural-layout int8 input with
    force_not_sb: the golden bytes after every half-iteration of the reference's 8-bit decoders"""
    K = case["K"]
    h = _tdec(lib)
    lib.srslte_tdec_force_not_sb(h)
    inp = np.ascontiguousarray(case["inp"])
    out = np.zeros(K // 8, np.uint8)
    assert lib.srslte_tdec_run_all_8bit(h, P(inp), P(out), case["nof"], K) == 0
    assert lib.srslte_tdec_get_nof_iterations(h) == case["nof"]
    assert (out == case["bytes"][case["nof"] - 1]).all()
    assert lib.srslte_tdec_new_cb(h, K) == 0
    for n in range(case["nof"]):
        lib.srslte_tdec_iteration_8bit(h, P(inp), P(out))
        assert (out == case["bytes"][n]).all(), (K, n)
    lib.srslte_tdec_free(h)


def test_tdec_8bit_sb_layout_input_vs_oracle(lib, orc):
    """without force_not_sb the int8 input is in the sub-block layout of srslte_rm_turbo_rx_lut_8bit for the 8-bit class of K"""
    for K in (408, 1024, 2112, 6144):
        E = 3 * K + 12 + 300
        e = (int_gauss(E, 37 * K) % 101 - 50).astype(np.int8)
        soft = np.zeros(2 * 18600, np.int8)
        assert orc.orc_rm_rx8(P(e), E, K, 0, 1, P(soft)) == 0
        want = np.zeros((3, K // 8), np.uint8)
        t = orc.orc_tb_new()
        orc.orc_tb_free(t)
        nat = np.zeros(3 * K + 12, np.int8)
        P8 = orc.orc_subblocks8(K)
        for tt in range(3 * K + 12):
            nat[tt] = soft[orc.orc_rm_soft_index(tt, K, P8)]
        assert orc.orc_tdec_halfits8(K, P(nat), 3, P(want), None) == 0
        h = _tdec(lib)
        out = np.zeros(K // 8, np.uint8)
        assert lib.srslte_tdec_run_all_8bit(h, P(soft), P(out), 3, K) == 0
        assert (out == want[2]).all(), K
        lib.srslte_tdec_free(h)


@pytest.mark.parametrize("case", list(G.tb8_cases()), ids=lambda c: c["id"])
def test_decode_tb_8bit_with_reference_softbuffer_calls(lib, case):
    """decode_tb with q->llr_is_8bit: one srslte_sch_decode_tb_cuda_8bit per (re)transmission of the golden sequences"""
    tbs = case["tbs"]
    nC = importlib.import_module("empower-srslte_b200").cbsegm(tbs)["C"]
    q = C.create_string_buffer(40)
    assert lib.srslte_softbuffer_rx_init(q, 100) == 0
    lib.srslte_softbuffer_rx_reset_tbs(q, tbs)
    for tx in case["tx"]:
        data = np.zeros(tbs // 8 + 3 + 768, np.uint8)
        noi = C.c_uint32(0)
        llr = np.ascontiguousarray(tx["llr"])
        ret = lib.srslte_sch_decode_tb_cuda_8bit(q, tbs, case["Qm"], tx["rv"], case["nbits"], P(llr), P(data), case["max_its"], C.byref(noi))
        assert ret == tx["ret"] and noi.value == tx["avg_its"]
        if ret == 0:
            assert (data[: tbs // 8 + 3] == tx["data"][: tbs // 8 + 3]).all()
        crc = (C.c_bool * 13).from_address(C.c_void_p.from_buffer(q, 24).value)
        assert [bool(crc[i]) for i in range(nC)] == [bool(v) for v in tx["crc"][:nC]]
    lib.srslte_softbuffer_rx_free(q)


def test_rm_rx_lut_8bit_accumulates_like_reference(lib, orc):
    """srslte_rm_turbo_rx_lut_8bit on host buffers: wrapping int8 accumulation into a non-empty soft buffer, every 8-bit
    decoder class, E below / above 3K+12, then the decoder entry point on the result"""
    rng = np.random.default_rng(77)
    for K in (40, 408, 1024, 2112, 6144):
        for rv, E in ((0, 3 * K + 12 + 333), (2, K + 7), (3, 2 * (3 * K + 12) + 5)):
            e = rng.integers(-128, 128, E).astype(np.int8)
            s0 = np.zeros(18600, np.int8)
            s0[: 3 * K + 108] = rng.integers(-128, 128, 3 * K + 108)
            got, want = s0.copy(), np.zeros(2 * 18600, np.int8)
            want[:18600] = s0
            assert lib.srslte_rm_turbo_rx_lut_8bit(P(e), P(got), E, ALL_K.index(K), rv) == 0
            assert orc.orc_rm_rx8(P(e), E, K, rv, 1, P(want)) == 0
            P8 = orc.orc_subblocks8(K)
            assert (G.soft8_view(got, K, P8) == G.soft8_view(want, K, P8)).all(), (K, rv, E)
    assert lib.srslte_rm_turbo_rx_lut_8bit(P(e), P(got), 10, 0, 4) == -2


def _tb_case(name_tbs):
    return [c for c in G.tb_cases() if c["tbs"] == name_tbs][0]


def test_reset_then_refused_call_then_valid_call(lib, orc):
    """a decode_tb the library refuses (-2: rv out of range) writes nothing to the device buffer, so it must not spend the
    deferred reset of srslte_softbuffer_rx_reset*: the next valid decode starts from zero like the reference's, whose
    reset zeroed the memory (softbuffer.c:138-151)"""
    case = _tb_case(6200)
    tbs, tx = case["tbs"], case["tx"][0]
    llr = np.ascontiguousarray(tx["llr"])
    q = C.create_string_buffer(40)
    assert lib.srslte_softbuffer_rx_init(q, 100) == 0
    data, noi = np.zeros(tbs // 8 + 3 + 768, np.uint8), C.c_uint32(0)
    # fill the device buffer with something that would break a later combine
    junk = (int_gauss(llr.size, 99) % 1601 - 800).astype(np.int16)
    lib.srslte_softbuffer_rx_reset_tbs(q, tbs)
    lib.srslte_sch_decode_tb_cuda(q, tbs, case["Qm"], 0, case["nbits"], P(junk), P(data), case["max_its"], C.byref(noi))
    lib.srslte_softbuffer_rx_reset_tbs(q, tbs)
    assert lib.srslte_sch_decode_tb_cuda(q, tbs, case["Qm"], 7, case["nbits"], P(llr), P(data), case["max_its"], C.byref(noi)) == -2
    ret = lib.srslte_sch_decode_tb_cuda(q, tbs, case["Qm"], tx["rv"], case["nbits"], P(llr), P(data), case["max_its"], C.byref(noi))
    assert ret == tx["ret"]
    if ret == 0:
        assert (data[: tbs // 8 + 3] == tx["data"][: tbs // 8 + 3]).all()
    lib.srslte_softbuffer_rx_free(q)


def test_softbuffer_free_recycles_device_buffers(lib):
    """srsENB creates and frees soft buffers per UE on attach / detach: freed buffers return to the pool (and come back
    zeroed), device memory does not grow"""
    import torch
    case = _tb_case(6200)
    tbs, tx = case["tbs"], case["tx"][0]
    llr = np.ascontiguousarray(tx["llr"])
    qs = [C.create_string_buffer(40) for _ in range(3)]
    for q in qs:                                     # make sure the pooled chunk exists before measuring
        assert lib.srslte_softbuffer_rx_init(q, 100) == 0
    for q in qs:
        lib.srslte_softbuffer_rx_free(q)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for it in range(200):
        q = C.create_string_buffer(40)
        assert lib.srslte_softbuffer_rx_init(q, 100) == 0
        if it % 50 == 0:
            data, noi = np.zeros(tbs // 8 + 3 + 768, np.uint8), C.c_uint32(0)
            lib.srslte_softbuffer_rx_reset_tbs(q, tbs)
            ret = lib.srslte_sch_decode_tb_cuda(q, tbs, case["Qm"], tx["rv"], case["nbits"], P(llr), P(data), case["max_its"], C.byref(noi))
            assert ret == tx["ret"]
        lib.srslte_softbuffer_rx_free(q)
    torch.cuda.synchronize()
    assert free0 - torch.cuda.mem_get_info()[0] < (8 << 20), "device memory grew across 200 init/free cycles"


def test_two_worker_threads_decode_concurrently(lib):
    """SURVEY 8b "Threading": one object per worker thread, workers run at once (pdsch.c:837-864), and a 16-bit and an 8-bit
    object may overlap. Each thread gets its own decode context; results must equal the golden ones of the single-threaded
    runs, in both modes, while the other thread is busy"""
    import threading
    c16 = _tb_case(6200)
    c8 = [c for c in G.tb8_cases()][0]
    errs = []

    def worker(case, fn, reps):
        try:
            tbs = case["tbs"]
            q = C.create_string_buffer(40)
            assert lib.srslte_softbuffer_rx_init(q, 100) == 0
            for rep in range(reps):
                lib.srslte_softbuffer_rx_reset_tbs(q, tbs)
                for txi, tx in enumerate(case["tx"]):
                    data, noi = np.zeros(tbs // 8 + 3 + 768, np.uint8), C.c_uint32(0)
                    llr = np.ascontiguousarray(tx["llr"])
                    ret = fn(q, tbs, case["Qm"], tx["rv"], case["nbits"], P(llr), P(data), case["max_its"], C.byref(noi))
                    assert ret == tx["ret"], (fn.__name__, rep, txi, ret, tx["ret"])
                    if ret == 0:
                        assert (data[: tbs // 8 + 3] == tx["data"][: tbs // 8 + 3]).all(), (fn.__name__, rep, txi)
            lib.srslte_softbuffer_rx_free(q)
        except Exception as e:                       # noqa: BLE001
            errs.append(repr(e))

    th = [threading.Thread(target=worker, args=(c16, lib.srslte_sch_decode_tb_cuda, 40)),
          threading.Thread(target=worker, args=(c8, lib.srslte_sch_decode_tb_cuda_8bit, 40)),
          threading.Thread(target=worker, args=(c16, lib.srslte_sch_decode_tb_cuda, 40))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
