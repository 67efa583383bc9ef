"""diagnostic: 8-bit mode at the bench shape, failures per noise level and comparison of a sample with the oracle"""
import importlib, sys, numpy as np
sys.path.insert(0, "tests")
from _libs import P, oracle
orc = oracle()
pkg = importlib.import_module("empower-srslte_b200")
f = pkg.FecCuda(0); f.set_llr_8bit(True)
TBS, QM, NB = 75376, 6, 90000
n = int(sys.argv[1]) if len(sys.argv) > 1 else 296
rng = np.random.default_rng(43)
data = rng.integers(0, 256, (n, TBS // 8)).astype(np.uint8)
sb = f.softbuffer_alloc(n)
out = np.zeros((n, TBS // 8 + 3 + 768), np.uint8)
e = np.zeros((n, NB // 8), np.uint8)
d = (pkg.TxTbDesc * n)()
for i in range(n):
    d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits = TBS, QM, 0, NB
    d[i].data, d[i].e_bits = data[i].ctypes.data, e[i].ctypes.data
f.encode_tbs(d)
bits = np.unpackbits(e, axis=1)[:, :NB].astype(np.float32)
for sigma in (0.2, 0.3):
    llr = np.clip(np.rint(24 * ((2 * bits - 1) + sigma * rng.standard_normal(bits.shape, dtype=np.float32))), -128, 127).astype(np.int8)
    r = (pkg.TbDesc * n)()
    for i in range(n):
        r[i].tbs, r[i].Qm, r[i].rv, r[i].nof_e_bits, r[i].max_its = TBS, QM, 0, NB, 8
        r[i].softbuffer, r[i].new_data = sb + i, 1
        r[i].e_bits, r[i].data = llr[i].ctypes.data, out[i].ctypes.data
    f.decode_tbs(r)
    bad = [i for i in range(n) if r[i].ret]
    print("sigma", sigma, "fails", len(bad), bad[:20])
    mism = 0
    for i in (bad[:6] + [0, 1]):
        t = orc.orc_tb_new(); orc.orc_tb_reset(t, TBS)
        do = np.zeros(TBS // 8 + 3 + 1024, np.uint8); co = np.zeros(16, np.uint8); io = np.zeros(16, np.uint32)
        ro = orc.orc_tb_decode8(t, TBS, QM, 0, NB, P(np.ascontiguousarray(llr[i])), 8, P(do), P(co), P(io))
        orc.orc_tb_free(t)
        g = [int(r[i].cb_its[k]) for k in range(13)]
        if ro != r[i].ret or [int(x) for x in io[:13]] != g: mism += 1
        print(i, "oracle", ro, list(io[:13]), "gpu", r[i].ret, g)
    print("oracle mismatches", mism)
