import importlib, numpy as np, time, sys, os
sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("empower-srslte_b200"); synth = importlib.import_module("empower-srslte_b200.synth")
TBS, QM, NBITS = 75376, 6, 90000
rng = np.random.default_rng(1)
pay = rng.integers(0, 256, (1, TBS // 8), dtype=np.uint8)
llr = synth.awgn_llr(synth.encode_tbs(pay, TBS, QM, 0, NBITS), 0.45, rng)
f = pkg.FecCuda(0, max_tbs=4)
h = pkg.PinnedArray((1, NBITS), np.int16); h.a[:] = llr
out = pkg.PinnedArray((1, TBS // 8 + 3 + 768), np.uint8)
sb = f.softbuffer_alloc(1)
d = (pkg.TbDesc * 1)()
d[0].tbs, d[0].Qm, d[0].rv, d[0].nof_e_bits, d[0].max_its, d[0].softbuffer, d[0].new_data = TBS, QM, 0, NBITS, 8, sb, 1
d[0].e_bits, d[0].data = h.a.ctypes.data, out.a.ctypes.data
for _ in range(20): f.decode_tbs(d, 0)
ts = []
for _ in range(50):
    t0 = time.perf_counter(); f.decode_tbs(d, 0); ts.append((time.perf_counter() - t0) * 1e6)
print("median us", sorted(ts)[25], "its", list(d[0].cb_its), "phases", f.last_phase_ms())
os.environ["X"] = "1"
