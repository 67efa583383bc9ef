"""empower-srslte_b200 — Python host mirror of libsrslte_fec_cuda (ctypes over the C ABI in include/).

The package holds only what the hot path needs: ``csrc/`` (CUDA kernels + C ABI) and this thin binding,
which mirrors the reference's operator interface for the path (``srslte_dlsch_decode`` ->
:meth:`FecCuda.decode_tbs`, ``srslte_tdec_run_all`` -> :meth:`FecCuda.decode_cbs`,
``srslte_rm_turbo_rx_lut_`` -> :meth:`FecCuda.rm_rx`). There is no CPU fallback: constructing
:class:`FecCuda` without the built extension or without a B200 raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FEC_CUDA_LIB") or os.path.join(HERE, "libsrslte_fec_cuda.so")   # FEC_CUDA_LIB: tuning builds (tools/variants.sh)
CSRC = os.path.join(HERE, "csrc")
SOURCES = ["fec_kernels.cu", "fec_tdec_packed.cu", "fec_frontend.cu", "fec_tx.cu", "fec_cuda_api.cu", "fec_compat.cu", "fec_noalloc.cu", "fec_tables.c"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-shared"]

MAX_CB = 13
F_DEVICE_IO = 1
F_NO_EARLY_STOP = 2
F_SERIAL = 4


class TxTbDesc(C.Structure):          # == fec_cuda_txtb_t
    _fields_ = [("tbs", C.c_uint32), ("Qm", C.c_uint32), ("rv", C.c_uint32), ("nof_e_bits", C.c_uint32),
                ("data", C.c_void_p), ("e_bits", C.c_void_p), ("ret", C.c_int32), ("reserved", C.c_uint32)]


class TbDesc(C.Structure):
    """fec_cuda_tb_t (include/srslte_fec_cuda.h)"""
    _fields_ = [("tbs", C.c_uint32), ("Qm", C.c_uint32), ("rv", C.c_uint32), ("nof_e_bits", C.c_uint32),
                ("max_its", C.c_uint32), ("softbuffer", C.c_uint32), ("new_data", C.c_uint32), ("reserved", C.c_uint32),
                ("e_bits", C.c_void_p), ("data", C.c_void_p),
                ("ret", C.c_int32), ("avg_its", C.c_uint32),
                ("cb_crc", C.c_uint8 * MAX_CB), ("cb_its", C.c_uint8 * MAX_CB), ("pad", C.c_uint8 * 6)]


class SymDesc(C.Structure):
    """fec_cuda_sym_t: one transport block given as equalised symbols (front end, include/srslte_fec_cuda.h)"""
    _fields_ = [("symbols", C.c_void_p), ("csi", C.c_void_p), ("nof_re", C.c_uint32), ("mod_bits", C.c_uint32),
                ("c_init", C.c_uint32), ("nof_symb", C.c_uint32), ("ack_pos", C.c_void_p), ("ri_pos", C.c_void_p),
                ("nof_ack_pos", C.c_uint32), ("nof_ri_pos", C.c_uint32), ("nof_cqi_bits", C.c_uint32), ("reserved", C.c_uint32),
                ("uci_out", C.c_void_p)]


class CbDesc(C.Structure):
    """fec_cuda_cb_t"""
    _fields_ = [("K", C.c_uint32), ("max_its", C.c_uint32), ("crc_type", C.c_uint32), ("reserved", C.c_uint32),
                ("llr", C.c_void_p), ("out", C.c_void_p), ("its", C.c_uint32), ("crc_ok", C.c_uint32)]


def build(force=False, extra_flags=()):
    """Compile every CUDA source for sm_100a into the in-tree shared library (nvcc cross-compiles on CPU).
    One object per source under csrc/_obj (stale ones only, in parallel), then one link."""
    from concurrent.futures import ThreadPoolExecutor
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hdrs += [os.path.join(os.path.dirname(HERE), "include", f) for f in ("srslte_fec_cuda.h", "srslte_fec_compat.h", "srslte_fec_abi.h")]
    hdr_t = max(os.path.getmtime(h) for h in hdrs if os.path.exists(h))
    objdir = os.path.join(CSRC, "_obj")
    os.makedirs(objdir, exist_ok=True)
    cflags = [f for f in NVCC_FLAGS if f != "-shared"] + list(extra_flags)

    def compile_one(src):
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        if force or extra_flags or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t):
            subprocess.check_call(["nvcc"] + cflags + ["-c", src, "-o", obj])
            return obj, True
        return obj, False

    with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
        res = list(ex.map(compile_one, srcs))
    if any(r[1] for r in res) or not os.path.exists(LIB_PATH) or any(os.path.getmtime(r[0]) > os.path.getmtime(LIB_PATH) for r in res):
        subprocess.check_call(["nvcc"] + NVCC_FLAGS + ["-o", LIB_PATH] + [r[0] for r in res])
    return LIB_PATH


_lib = None


def load():
    """dlopen the C-ABI library; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(the CUDA extension is mandatory, there is no CPU path)")
        lib = C.CDLL(LIB_PATH)
        lib.fec_cuda_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_uint32]
        lib.fec_cuda_destroy.argtypes = [C.c_void_p]
        lib.fec_cuda_last_error.argtypes = [C.c_void_p]
        lib.fec_cuda_last_error.restype = C.c_char_p
        lib.fec_cuda_stream.argtypes = [C.c_void_p]
        lib.fec_cuda_stream.restype = C.c_void_p
        lib.fec_cuda_launch_count.argtypes = [C.c_void_p]
        lib.fec_cuda_launch_count.restype = C.c_uint64
        lib.fec_cuda_softbuffer_alloc.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
        lib.fec_cuda_softbuffer_reset.argtypes = [C.c_void_p, C.c_uint32]
        lib.fec_cuda_softbuffer_free.argtypes = [C.c_void_p, C.c_uint32]
        lib.fec_cuda_set_decoder_class.argtypes = [C.c_void_p, C.c_int]
        lib.fec_cuda_softbuffer_read_ref_layout.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        lib.fec_cuda_decode_tbs.argtypes = [C.c_void_p, C.POINTER(TbDesc), C.c_uint32, C.c_uint32]
        lib.fec_cuda_submit_tbs.argtypes = [C.c_void_p, C.POINTER(TbDesc), C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
        lib.fec_cuda_wait.argtypes = [C.c_void_p, C.c_uint64]
        lib.fec_cuda_decode_cbs.argtypes = [C.c_void_p, C.POINTER(CbDesc), C.c_uint32, C.c_uint32]
        lib.fec_cuda_rm_rx.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
        lib.fec_cuda_rm_rx_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]
        lib.fec_cuda_cbsegm.argtypes = [C.c_uint32, C.POINTER(C.c_uint32)]
        lib.fec_cuda_last_phase_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        lib.fec_cuda_last_path_counts.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        lib.fec_cuda_set_llr_8bit.argtypes = [C.c_void_p, C.c_int]
        lib.fec_cuda_encode_tbs.argtypes = [C.c_void_p, C.POINTER(TxTbDesc), C.c_uint32, C.c_uint32]
        lib.fec_cuda_decode_syms.argtypes = [C.c_void_p, C.POINTER(TbDesc), C.POINTER(SymDesc), C.c_uint32, C.c_uint32]
        lib.fec_cuda_submit_syms.argtypes = [C.c_void_p, C.POINTER(TbDesc), C.POINTER(SymDesc), C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
        lib.fec_cuda_last_frontend_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        lib.fec_cuda_frontend.argtypes = [C.c_void_p, C.POINTER(SymDesc), C.c_void_p]
        lib.fec_cuda_cinit_pdsch.restype = C.c_uint32
        lib.fec_cuda_cinit_pdsch.argtypes = [C.c_uint32] * 4
        lib.fec_cuda_cinit_pusch.restype = C.c_uint32
        lib.fec_cuda_cinit_pusch.argtypes = [C.c_uint32] * 3
        lib.fec_cuda_host_alloc.argtypes = [C.c_size_t]
        lib.fec_cuda_host_alloc.restype = C.c_void_p
        lib.fec_cuda_host_free.argtypes = [C.c_void_p]
        _lib = lib
    return _lib


def cbsegm(tbs):
    """srslte_cbsegm: dict(F,C,K1,K2,K1_idx,K2_idx,C1,C2,tbs) or None on error"""
    o = (C.c_uint32 * 9)()
    if load().fec_cuda_cbsegm(tbs, o):
        return None
    return dict(zip(("F", "C", "K1", "K2", "K1_idx", "K2_idx", "C1", "C2", "tbs"), [int(v) for v in o]))


class PinnedArray:
    """numpy view over page-locked host memory from fec_cuda_host_alloc (DMA'd without a staging copy)"""

    def __init__(self, shape, dtype):
        self.lib = load()
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = self.lib.fec_cuda_host_alloc(nbytes)
        if not self.ptr:
            raise RuntimeError("fec_cuda_host_alloc failed (no CUDA device?)")
        buf = (C.c_uint8 * nbytes).from_address(self.ptr)
        self.a = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self):
        if self.ptr:
            self.a = None
            self.lib.fec_cuda_host_free(self.ptr)
            self.ptr = None


class FecCuda:
    """One decode context = one CUDA device + one stream (mirrors one srslte_sch_t per worker thread)."""

    def __init__(self, device=0, max_tbs=1024):
        self.lib = load()
        self.h = C.c_void_p()
        rc = self.lib.fec_cuda_create(C.byref(self.h), device, max_tbs)
        if rc != 0:
            raise RuntimeError(f"fec_cuda_create failed ({rc}): no usable sm_100a device; this library has no CPU path")

    def close(self):
        if self.h:
            self.lib.fec_cuda_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def stream(self):
        return self.lib.fec_cuda_stream(self.h)

    @property
    def launches(self):
        return int(self.lib.fec_cuda_launch_count(self.h))

    def last_phase_ms(self):
        """device ms of (rate de-matching, turbo decode, TB assembly+CRC24A) in the last decode_tbs call"""
        o = (C.c_float * 3)()
        rc = self.lib.fec_cuda_last_phase_ms(self.h, o)
        if rc:
            raise RuntimeError(f"fec_cuda_last_phase_ms: {rc}")
        return [float(v) for v in o]

    def encode_tbs(self, descs, flags=0):
        """fec_cuda_encode_tbs on a ctypes array of TxTbDesc"""
        rc = self.lib.fec_cuda_encode_tbs(self.h, descs, len(descs), flags)
        if rc:
            raise RuntimeError(f"fec_cuda_encode_tbs: {rc} {self.error()}")

    def encode_tb(self, tbs, Qm, rv, nof_e_bits, data):
        """one transport block: data[tbs/8] uint8 -> (ret, packed e_bits[(nof_e_bits+7)/8]) as srslte_dlsch_encode emits them"""
        d = (TxTbDesc * 1)()
        a = np.ascontiguousarray(data, dtype=np.uint8)
        o = np.zeros((nof_e_bits + 7) // 8, np.uint8)
        d[0].tbs, d[0].Qm, d[0].rv, d[0].nof_e_bits = tbs, Qm, rv, nof_e_bits
        d[0].data, d[0].e_bits = a.ctypes.data, o.ctypes.data
        self.encode_tbs(d)
        return int(d[0].ret), o

    def set_llr_8bit(self, on):
        """srslte_sch_t.llr_is_8bit: e_bits of the following decode calls are int8, the 8-bit rate de-matcher and decoders run"""
        rc = self.lib.fec_cuda_set_llr_8bit(self.h, 1 if on else 0)
        if rc:
            raise RuntimeError(f"fec_cuda_set_llr_8bit: {rc}")

    def last_path_counts(self):
        """code blocks of the last decode call handed (to the certificate variant of the packed kernel, on to the exact kernel)"""
        o = (C.c_uint64 * 2)()
        rc = self.lib.fec_cuda_last_path_counts(self.h, o)
        if rc:
            raise RuntimeError(f"fec_cuda_last_path_counts: {rc}")
        return int(o[0]), int(o[1])

    def error(self):
        return self.lib.fec_cuda_last_error(self.h).decode()

    def softbuffer_alloc(self, n):
        first = C.c_uint32()
        rc = self.lib.fec_cuda_softbuffer_alloc(self.h, n, C.byref(first))
        if rc:
            raise RuntimeError(f"softbuffer_alloc: {rc} {self.error()}")
        return int(first.value)

    def softbuffer_free(self, sb):
        return self.lib.fec_cuda_softbuffer_free(self.h, sb)

    def set_decoder_class(self, P):
        """srslte_tdec_init_manual: force the non-windowed (0) / 8 / 16 sub-block 16-bit decoder for every K; -1 = AUTO"""
        rc = self.lib.fec_cuda_set_decoder_class(self.h, P)
        if rc:
            raise RuntimeError(f"fec_cuda_set_decoder_class: {rc}")

    def softbuffer_reset(self, sb):
        return self.lib.fec_cuda_softbuffer_reset(self.h, sb)

    def softbuffer_read(self, sb, cb, K):
        out = np.zeros(18600, np.int16)
        rc = self.lib.fec_cuda_softbuffer_read_ref_layout(self.h, sb, cb, K, out.ctypes.data_as(C.c_void_p))
        if rc:
            raise RuntimeError(f"softbuffer_read: {rc} {self.error()}")
        return out

    def decode_tbs(self, descs, flags=0):
        """descs: ctypes array of TbDesc (filled in place). Raises on a batch-level failure."""
        rc = self.lib.fec_cuda_decode_tbs(self.h, descs, len(descs), flags)
        if rc:
            raise RuntimeError(f"fec_cuda_decode_tbs: {rc} {self.error()}")
        return descs

    def submit_tbs(self, descs, flags=0):
        """asynchronous half of decode_tbs: enqueue the batch, return its ticket (results are valid after wait)"""
        t = C.c_uint64(0)
        rc = self.lib.fec_cuda_submit_tbs(self.h, descs, len(descs), flags, C.byref(t))
        if rc:
            raise RuntimeError(f"fec_cuda_submit_tbs: {rc} {self.error()}")
        return t.value

    def wait(self, ticket):
        rc = self.lib.fec_cuda_wait(self.h, ticket)
        if rc:
            raise RuntimeError(f"fec_cuda_wait: {rc} {self.error()}")

    def decode_syms(self, descs, syms, flags=0):
        """transport blocks given as equalised symbols (TbDesc array + SymDesc array, filled in place)"""
        rc = self.lib.fec_cuda_decode_syms(self.h, descs, syms, len(descs), flags)
        if rc:
            raise RuntimeError(f"fec_cuda_decode_syms: {rc} {self.error()}")
        return descs

    def submit_syms(self, descs, syms, flags=0):
        t = C.c_uint64(0)
        rc = self.lib.fec_cuda_submit_syms(self.h, descs, syms, len(descs), flags, C.byref(t))
        if rc:
            raise RuntimeError(f"fec_cuda_submit_syms: {rc} {self.error()}")
        return t.value

    def last_frontend_ms(self):
        o = C.c_float(0)
        rc = self.lib.fec_cuda_last_frontend_ms(self.h, C.byref(o))
        if rc:
            raise RuntimeError(f"fec_cuda_last_frontend_ms: {rc}")
        return float(o.value)

    def frontend(self, symbols, mod_bits, c_init, csi=None, nof_symb=0, ack_pos=(), ri_pos=(), nof_cqi_bits=0, llr8=False):
        """symbols -> LLRs as decode_tb receives them (demodulation, descrambling, CSI weights / UL de-interleaver).
        symbols: float32 pairs (re, im). Returns (llr int16[nof_re*mod_bits], uci int16[nack+nri+ncqi])"""
        sym = np.ascontiguousarray(symbols, dtype=np.float32).reshape(-1)
        n = sym.size // 2
        buf = np.zeros(2 * n + 8, np.float32)
        off = ((-buf.ctypes.data) % 16) // 4
        buf[off:off + 2 * n] = sym                               # 16-byte aligned like srslte_vec_malloc memory
        d = SymDesc()
        d.symbols, d.nof_re, d.mod_bits, d.c_init, d.nof_symb = buf.ctypes.data + 4 * off, n, mod_bits, c_init, nof_symb
        keep = [buf]
        if csi is not None:
            c = np.ascontiguousarray(csi, dtype=np.float32)
            keep.append(c)
            d.csi = c.ctypes.data
        ack = np.ascontiguousarray(ack_pos, dtype=np.uint32)
        ri = np.ascontiguousarray(ri_pos, dtype=np.uint32)
        uci = np.zeros(ack.size + ri.size + nof_cqi_bits + 1, np.int16)
        d.ack_pos, d.ri_pos, d.nof_ack_pos, d.nof_ri_pos, d.nof_cqi_bits = ack.ctypes.data, ri.ctypes.data, ack.size, ri.size, nof_cqi_bits
        d.uci_out = uci.ctypes.data
        out = np.zeros(n * mod_bits, np.int8 if llr8 else np.int16)      # llr8: the context is in 8-bit LLR mode (set_llr_8bit)
        rc = self.lib.fec_cuda_frontend(self.h, C.byref(d), out.ctypes.data_as(C.c_void_p))
        if rc:
            raise RuntimeError(f"fec_cuda_frontend: {rc} {self.error()}")
        return out, uci[:-1]

    def decode_tb(self, tbs, Qm, rv, e_bits, softbuffer, new_data, max_its=8, flags=0):
        """single transport block convenience (the srslte_dlsch_decode call shape). e_bits: int16 numpy."""
        e = np.ascontiguousarray(e_bits, dtype=np.int16)
        data = np.zeros(tbs // 8 + 3 + 768, np.uint8)
        d = (TbDesc * 1)()
        d[0].tbs, d[0].Qm, d[0].rv, d[0].nof_e_bits, d[0].max_its = tbs, Qm, rv, e.size, max_its
        d[0].softbuffer, d[0].new_data = softbuffer, int(new_data)
        d[0].e_bits, d[0].data = e.ctypes.data, data.ctypes.data
        self.decode_tbs(d, flags)
        Cn = cbsegm(tbs)["C"] if tbs else 0
        return dict(ret=int(d[0].ret), data=data[: tbs // 8 + 3], avg_its=int(d[0].avg_its),
                    cb_crc=np.array(d[0].cb_crc[:Cn], np.uint8), cb_its=np.array(d[0].cb_its[:Cn], np.uint32))

    def decode_cbs(self, K_list, llr_list, max_its, crc_type=1, early_stop=True):
        """bare code blocks from natural-layout LLRs; returns list of dict(out, its, crc_ok)"""
        n = len(K_list)
        d = (CbDesc * n)()
        keep = []
        for i, (K, llr) in enumerate(zip(K_list, llr_list)):
            a = np.ascontiguousarray(llr, dtype=np.int16)
            o = np.zeros(K // 8, np.uint8)
            keep.append((a, o))
            d[i].K, d[i].max_its, d[i].crc_type = K, max_its, crc_type
            d[i].llr, d[i].out = a.ctypes.data, o.ctypes.data
        rc = self.lib.fec_cuda_decode_cbs(self.h, d, n, 0 if early_stop else F_NO_EARLY_STOP)
        if rc:
            raise RuntimeError(f"fec_cuda_decode_cbs: {rc} {self.error()}")
        return [dict(out=keep[i][1], its=int(d[i].its), crc_ok=int(d[i].crc_ok)) for i in range(n)]

    def rm_rx(self, e, cb_idx, rv, soft, sb_layout=True):
        """srslte_rm_turbo_rx_lut_(in, soft, E, cb_idx, rv, enable_input_tdec): soft (18600 int16) updated in place"""
        e = np.ascontiguousarray(e, dtype=np.int16)
        assert soft.dtype == np.int16 and soft.size >= 18600 and soft.flags.c_contiguous
        return self.lib.fec_cuda_rm_rx(self.h, e.ctypes.data, e.size, cb_idx, rv, int(sb_layout), soft.ctypes.data)
