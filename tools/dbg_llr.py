"""debug helper: compare per-bit LLRs of the CUDA decoder with the CPU oracle after nof half-iterations"""
import ctypes as C, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from _libs import P, oracle, int_gauss
pkg = importlib.import_module("empower-srslte_b200")
fec = pkg.FecCuda(0); lib = pkg.load(); orc = oracle()
lib.fec_cuda_debug_llr_enable.argtypes = [C.c_void_p, C.c_uint32]
lib.fec_cuda_debug_llr_read.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
lib.fec_cuda_debug_llr_enable(fec.h, 4)
for K in [int(a) for a in sys.argv[1:]] or [408, 816]:
    inp = (int_gauss(3 * K + 12, K) % 601 - 300).astype(np.int16)
    for nof in (1, 2, 3):
        ol = np.zeros((nof, K), np.int16); ob = np.zeros((nof, K // 8), np.uint8)
        orc.orc_tdec_halfits(K, P(inp), nof, P(ob), P(ol))
        fec.decode_cbs([K], [inp], nof, early_stop=False)
        g = np.zeros(6144, np.int16); lib.fec_cuda_debug_llr_read(fec.h, P(g), 1)
        g = g[:K]; bad = np.nonzero(g != ol[nof - 1])[0]
        Pn = lib.fec_cuda_subblocks(K); L = K // max(Pn, 1)
        print(f"K={K} nof={nof} mismatches={bad.size}/{K}", "first:", bad[:12], "lanes:", sorted(set((bad // L).tolist()))[:16], "k in lane:", (bad % L)[:12])
        if bad.size:
            print("  gpu", g[bad[:8]], "ref", ol[nof - 1][bad[:8]])
