"""CPU-only checks of the product's host side: the C-ABI library loads without a GPU, exports every
symbol include/*.h declares, fails loudly (no fallback) when no device exists, and its integer
bookkeeping (segmentation, size table, decoder class) matches the golden data from the reference."""
import ctypes as C
import importlib
import os
import re

import numpy as np

import _golden as G
from _libs import ALL_K, ROOT


def _pkg():
    pkg = importlib.import_module("empower-srslte_b200")
    pkg.build()
    return pkg


def test_exports_every_declared_symbol():
    pkg = _pkg()
    lib = pkg.load()
    names = set()
    for h in os.listdir(os.path.join(ROOT, "include")):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b((?:fec_cuda|srslte)_[a-z0-9_]+)\s*\(", src))
    assert len(names) >= 15
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, missing


def test_no_cpu_fallback_without_device():
    import torch
    pkg = _pkg()
    if torch.cuda.is_available():
        return
    h = C.c_void_p()
    assert pkg.load().fec_cuda_create(C.byref(h), 0, 16) == -100   # FEC_CUDA_ENODEV
    try:
        pkg.FecCuda(0)
        raise AssertionError("FecCuda must raise without a GPU")
    except RuntimeError:
        pass


def test_segmentation_and_size_table_golden():
    pkg = _pkg()
    lib = pkg.load()
    z = G.tables()
    for row in z["cbsegm"]:
        o = (C.c_uint32 * 9)()
        assert lib.fec_cuda_cbsegm(int(row[0]), o) == int(row[1])
        if row[1] == 0:
            assert [int(v) for v in o] == [int(v) for v in row[2:]]
    for i, K in enumerate(ALL_K):
        assert lib.fec_cuda_cb_size(i) == K and lib.fec_cuda_cb_index(K) == i
        assert lib.fec_cuda_cb_index(K - 1) == i
        assert lib.fec_cuda_subblocks(K) == (16 if K > 800 and K % 16 == 0 else 8 if K > 400 else 0)
    assert lib.fec_cuda_cb_index(6145) == -1 and lib.fec_cuda_cb_size(188) == -1


def test_product_never_touches_oracle():
    """the shipped sources must not reference oracle/ (SURVEY tier rule: the oracle is the checker only)"""
    csrc = os.path.join(ROOT, "empower-srslte_b200")
    for dp, _, fs in os.walk(csrc):
        for f in fs:
            if f.endswith((".cu", ".cuh", ".c", ".h", ".py")):
                s = open(os.path.join(dp, f)).read()
                assert "fec_oracle" not in s and "libsrslte_ref" not in s and "oracle/" not in s, f


def test_compat_entry_points_without_device():
    """the reference's own symbols (include/srslte_fec_compat.h): pure bookkeeping works anywhere, everything that
    computes refuses to run without a device (no CPU fallback), with the reference's error codes"""
    import torch
    pkg = _pkg()
    lib = pkg.load()
    for K in ALL_K:
        assert lib.srslte_tdec_autoimp_get_subblocks(K) == (16 if K > 800 and K % 16 == 0 else 8 if K > 400 else 0)
        assert lib.srslte_tdec_autoimp_get_subblocks_8bit(K) == (32 if K > 2048 and K % 32 == 0 else 16 if K > 800 and K % 16 == 0
                                                                 else 8 if K > 400 else 0)
    e, s = np.zeros(16, np.int16), np.zeros(18600, np.int16)
    lib.srslte_rm_turbo_rx_lut.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
    assert lib.srslte_rm_turbo_rx_lut(e.ctypes.data, s.ctypes.data, 16, 188, 0) == -2     # rm_turbo.c:426-429
    assert lib.srslte_rm_turbo_rx_lut(e.ctypes.data, s.ctypes.data, 16, 0, 4) == -2
    if torch.cuda.is_available():
        return
    h = C.create_string_buffer(18264)
    lib.srslte_tdec_init.argtypes = [C.c_void_p, C.c_uint32]
    assert lib.srslte_tdec_init(h, 6144) == -1
    q = C.create_string_buffer(40)
    lib.srslte_softbuffer_rx_init.argtypes = [C.c_void_p, C.c_uint32]
    assert lib.srslte_softbuffer_rx_init(q, 100) == -1
    assert lib.srslte_rm_turbo_rx_lut(e.ctypes.data, s.ctypes.data, 16, 0, 0) == -1
