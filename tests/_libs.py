"""ctypes loaders for the CHECKERS used by the tests (never by the product):

* ``oracle()``  -> oracle/_build/libfec_oracle.so, our scalar C restatement (built on demand with gcc);
* ``ref()``     -> oracle/_ref/libsrslte_ref.so, the unmodified reference objects + oracle/ref_shim.c
                   (prebuilt; rebuilt here only when /root/reference is mounted). ``None`` if unavailable.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_cache = {}


def P(a):
    """pointer to a numpy array's data"""
    return a.ctypes.data_as(C.c_void_p)


def oracle():
    if "orc" not in _cache:
        so = os.path.join(ORACLE_DIR, "_build", "libfec_oracle.so")
        srcs = [os.path.join(ORACLE_DIR, f) for f in ("fec_oracle.c", "fec_oracle.h", "lte_qpp_table.h")]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "oracle"])
        lib = C.CDLL(so)
        lib.orc_tb_new.restype = C.c_void_p
        lib.orc_tb_free.argtypes = [C.c_void_p]
        lib.orc_tb_reset.argtypes = [C.c_void_p, C.c_uint32]
        lib.orc_tb_decode.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                                      C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.orc_tb_decode8.argtypes = lib.orc_tb_decode.argtypes
        lib.orc_tb_get_soft.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.orc_crc.restype = C.c_uint32
        lib.orc_crc.argtypes = [C.c_int, C.c_void_p, C.c_int]
        lib.orc_bench_tb_mt.restype = C.c_double
        lib.orc_bench_tb_mt.argtypes = [C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                                        C.c_uint32, C.c_int, C.c_void_p]
        lib.orc_map.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 6
        lib.orc_demod_s.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        lib.orc_demod_b.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        lib.orc_descramble_b.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        lib.orc_descramble_b.restype = None
        lib.orc_csi_b.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        lib.orc_csi_b.restype = None
        lib.orc_sequence.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p]
        lib.orc_sequence.restype = None
        lib.orc_descramble_s.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        lib.orc_descramble_s.restype = None
        lib.orc_csi_s.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
        lib.orc_csi_s.restype = None
        lib.orc_cinit_pdsch.restype = C.c_uint32
        lib.orc_cinit_pusch.restype = C.c_uint32
        lib.orc_ulsch_deinterleave.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]
        lib.orc_ulsch_deinterleave.restype = None
        _cache["orc"] = lib
    return _cache["orc"]


def ref():
    if "ref" not in _cache:
        so = os.path.join(ORACLE_DIR, "_ref", "libsrslte_ref.so")
        if os.path.isdir("/root/reference/lib/src/phy/fec"):
            shims = [os.path.join(ORACLE_DIR, f) for f in ("ref_shim.c", "ref_shim_fe.c", "Makefile")]
            if not os.path.exists(so) or any(os.path.getmtime(f) > os.path.getmtime(so) for f in shims):
                subprocess.check_call(["make", "-s", "-j8", "-C", ORACLE_DIR, "ref"])
        if not os.path.exists(so):
            _cache["ref"] = None
            return None
        lib = C.CDLL(so)
        lib.ref_tb_new.restype = C.c_void_p
        lib.ref_tb_free.argtypes = [C.c_void_p]
        lib.ref_tb_reset.argtypes = [C.c_void_p, C.c_uint32]
        lib.ref_tb_max_cb.argtypes = [C.c_void_p]
        lib.ref_tb_encode.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        lib.ref_tb_decode.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                                      C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.ref_tb_decode_detail.argtypes = lib.ref_tb_decode.argtypes
        if hasattr(lib, "ref_tb_decode8"):
            lib.ref_tb_decode8.argtypes = lib.ref_tb_decode.argtypes
        lib.ref_tb_get_soft.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.ref_tb_set_soft.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.ref_crc.restype = C.c_uint32
        lib.ref_crc.argtypes = [C.c_int, C.c_void_p, C.c_int]
        lib.ref_bench_tb_mt.restype = C.c_double
        lib.ref_bench_tb_mt.argtypes = [C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p,
                                        C.c_uint32, C.c_int, C.c_int, C.c_void_p]
        if hasattr(lib, "ref_cw_new"):           # front-end shim (ref_shim_fe.c)
            lib.ref_cw_new.restype = C.c_void_p
            lib.ref_cw_new.argtypes = [C.c_int, C.c_uint32]
            lib.ref_cw_free.argtypes = [C.c_void_p]
            lib.ref_cw_reset.argtypes = [C.c_void_p, C.c_uint32]
            lib.ref_cw_decode.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                          C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
            lib.ref_ul_new.restype = C.c_void_p
            lib.ref_ul_new.argtypes = [C.c_int]
            lib.ref_ul_free.argtypes = [C.c_void_p]
            lib.ref_ul_reset.argtypes = [C.c_void_p, C.c_uint32]
            lib.ref_ul_decode.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_void_p] * 3 + \
                                         [C.c_uint32] + [C.c_void_p] * 9
            lib.ref_modulate.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
            lib.ref_demod_s.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
            lib.ref_sequence.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p]
            lib.ref_descramble_s.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32]
            lib.ref_csi_s.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
            if hasattr(lib, "ref_demod_b"):
                lib.ref_demod_b.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
                lib.ref_descramble_b.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32]
                lib.ref_csi_b.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]
            lib.ref_cinit_pdsch.restype = C.c_uint32
            lib.ref_cinit_pusch.restype = C.c_uint32
            lib.ref_ulsch_deinterleave.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]
        assert lib.ref_init() == 0
        _cache["ref"] = lib
    return _cache["ref"]


def aligned(n, dtype, align=64):
    """zeroed 1-D array whose data pointer is `align`-byte aligned (what srslte_vec_malloc hands the reference's SIMD loops)"""
    item = np.dtype(dtype).itemsize
    raw = np.zeros(n * item + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * item].view(dtype)


ALL_K = (list(range(40, 513, 8)) + list(range(528, 1025, 16)) + list(range(1056, 2049, 32))
         + list(range(2112, 6145, 64)))


def awgn_llr(bits, sigma, rng, amp=100.0):
    """LLR = rint(amp * (+-1 + sigma*n)), bit 1 -> +1 (turbodecoder_test.c:245-252 convention)"""
    x = 2.0 * bits.astype(np.float64) - 1.0
    return np.clip(np.rint(amp * (x + sigma * rng.standard_normal(bits.shape))), -32768, 32767).astype(np.int16)


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def int_gauss(n, seed):
    """Platform-independent pseudo-Gaussian integers: sum of twelve 16-bit uniforms minus the mean
    (Irwin-Hall), std ~= 65536. Pure uint64 arithmetic -> bit-identical on every box."""
    with np.errstate(over="ignore"):
        idx = np.arange(n, dtype=np.uint64) * np.uint64(3) + np.uint64(seed) * np.uint64(0x100000001B3)
        acc = np.zeros(n, np.int64)
        for j in range(3):
            r = _splitmix64(idx + np.uint64(j))
            for s in (0, 16, 32, 48):
                acc += ((r >> np.uint64(s)) & np.uint64(0xFFFF)).astype(np.int64)
    return acc - 6 * 65535


def int_awgn_llr(bits, sigma_milli, seed, amp=100):
    """LLR = amp*(2b-1) + amp*sigma*n with integer-only arithmetic (sigma in 1/1000), int16-clipped.
    Same convention as turbodecoder_test.c:245-252 (bit 1 -> +amp)."""
    g = int_gauss(bits.size, seed)
    noise = (amp * sigma_milli * g) // (1000 * 65536)
    v = amp * (2 * bits.astype(np.int64) - 1) + noise
    return np.clip(v, -32768, 32767).astype(np.int16)
