"""fec_cuda_rm_rx_noalloc (the allocation-free entry point, static __device__ scratch) through the C ABI against the
existing fec_cuda_rm_rx on the same inputs, bit-exact, on calls in a row that share the scratch array. Collected last on
purpose (zz)."""
import ctypes as C
import importlib

import numpy as np
import pytest

from _libs import ALL_K, P, int_gauss

pytestmark = pytest.mark.gpu


def _extent(K, sb):
    return 3 * (K + 32) + 12 if sb and K > 400 else 3 * K + 12


def test_rm_rx_noalloc_matches_rm_rx_twice_in_a_row():
    pkg = importlib.import_module("empower-srslte_b200")
    lib = pkg.load()
    fec = pkg.FecCuda(0)
    try:
        # (K, rv, sb_layout, E): second call changes K, rv, layout and length (E > 3K+12: repeats accumulate), third goes
        # back to the first K with another rv (rank table rewritten in both directions), fourth keeps K (table reused)
        calls = [(6144, 0, 1, 20000), (1024, 2, 0, 5000), (6144, 3, 1, 9999), (6144, 1, 1, 7)]
        for n, (K, rv, sb, E) in enumerate(calls):
            e = (int_gauss(E, 100 + n) % 2001 - 1000).astype(np.int16)
            s0 = (int_gauss(18600, 200 + n) % 20001 - 10000).astype(np.int16)
            want = s0.copy()
            assert fec.rm_rx(e, ALL_K.index(K), rv, want, bool(sb)) == 0
            # the new entry point gets a buffer of exactly the extent the layout uses, between canaries
            ext = _extent(K, sb)
            buf = np.full(ext + 64, 0x5A5A, np.int16)
            buf[32:32 + ext] = s0[:ext]
            rc = lib.fec_cuda_rm_rx_noalloc(0, P(e), E, ALL_K.index(K), rv, sb, C.c_void_p(buf.ctypes.data + 64))
            assert rc == 0, (n, rc)
            assert (buf[32:32 + ext] == want[:ext]).all(), (n, K, rv, sb, E)
            assert (buf[:32] == 0x5A5A).all() and (buf[32 + ext:] == 0x5A5A).all(), "wrote outside the extent"
            assert not (want[:ext] == s0[:ext]).all()
    finally:
        fec.close()


def test_rm_rx_noalloc_rejects_beyond_its_maximum():
    pkg = importlib.import_module("empower-srslte_b200")
    lib = pkg.load()
    s = np.zeros(18600, np.int16)
    e = np.zeros(32769, np.int16)
    assert lib.fec_cuda_rm_rx_noalloc(0, P(e), 32769, 187, 0, 1, P(s)) == -2      # E > FEC_CUDA_NOALLOC_MAX_E
    assert lib.fec_cuda_rm_rx_noalloc(0, P(e), 10, 188, 0, 1, P(s)) == -2         # no such code block size
    assert lib.fec_cuda_rm_rx_noalloc(0, P(e), 10, 0, 4, 1, P(s)) == -2           # rv
    assert lib.fec_cuda_rm_rx_noalloc(0, P(e), 32768, 187, 0, 1, P(s)) == 0       # the maximum itself is fine
