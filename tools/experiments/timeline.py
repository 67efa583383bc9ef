#!/usr/bin/env python
"""CTA timeline of the decoder launches of a pipelined run (library built with -DPK_TIMELINE=1):
   FEC_CUDA_LIB=build/variants/timeline.so python tools/experiments/timeline.py
prints, per launch in order of start: first CTA start, time the first / last CTA found the queue empty, last CTA end, and how
many CTAs of the NEXT launch had started before this launch's last CTA ended (tail overlap)."""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("empower-srslte_b200")
synth = importlib.import_module("empower-srslte_b200.synth")
B, TBS, QM, NBITS = int(os.environ.get("TL_BATCH", "4736")), 75376, 6, 90000
rng = np.random.default_rng(1)
payload = rng.integers(0, 256, (16, TBS // 8), dtype=np.uint8)
e = synth.encode_tbs(payload, TBS, QM, 0, NBITS)
llr = np.zeros((B, NBITS), np.int16)
for i0 in range(0, B, 64):
    i1 = min(B, i0 + 64)
    synth.awgn_llr(e[np.arange(i0, i1) % 16], 0.45, rng, out=llr[i0:i1])
f = pkg.FecCuda(0, max_tbs=B)
dl = torch.from_numpy(llr).cuda()
out = [torch.zeros((B, TBS // 8 + 3 + 768), dtype=torch.uint8, device="cuda") for _ in range(2)]
sb = f.softbuffer_alloc(2 * B)
sets = []
for s in range(2):
    d = (pkg.TbDesc * B)()
    for i in range(B):
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its, d[i].softbuffer, d[i].new_data = TBS, QM, 0, NBITS, 8, sb + s * B + i, 1
        d[i].e_bits, d[i].data = dl[i].data_ptr(), out[s][i].data_ptr()
    sets.append(d)
for i in range(8):
    f.decode_tbs(sets[i & 1], pkg.F_DEVICE_IO)
buf = np.zeros(6 * 16384, np.uint64)
f.lib.fec_cuda_debug_timeline(buf.ctypes.data_as(C.c_void_p), 16384)          # drop the warm-up records
prev = None
for i in range(6):
    t = f.submit_tbs(sets[i & 1], pkg.F_DEVICE_IO)
    if prev is not None:
        f.wait(prev)
    prev = t
f.wait(prev)
n = f.lib.fec_cuda_debug_timeline(buf.ctypes.data_as(C.c_void_p), 16384)
r = buf[: 6 * n].reshape(n, 6).astype(np.int64)
t0 = r[:, 2].min()
launches = {}
for row in r:
    launches.setdefault((row[0]), []).append(row)
groups = []
# the same queue address recurs (pipes are reused): split a queue's records into launches by start-time gaps
for q, rows in launches.items():
    rows = sorted(rows, key=lambda x: x[2])
    cur = [rows[0]]
    for x in rows[1:]:
        if x[2] - cur[-1][2] > 1_500_000:
            groups.append(cur); cur = [x]
        else:
            cur.append(x)
    groups.append(cur)
groups.sort(key=lambda g: min(x[2] for x in g))
print("launch  CTAs  first start   last start   first empty   last empty   first end   last end  (ms)   next launch's CTAs started before last end")
for k, g in enumerate(groups):
    a = np.array(g)
    st, em, en = (a[:, 2] - t0) / 1e6, (a[:, 3] - t0) / 1e6, (a[:, 4] - t0) / 1e6
    nxt = 0
    if k + 1 < len(groups):
        b = np.array(groups[k + 1])
        nxt = int(((b[:, 2] - t0) / 1e6 < en.max()).sum())
    print("%4d  %5d  %10.3f  %10.3f  %11.3f  %11.3f  %10.3f  %9.3f     %d" % (k, len(g), st.min(), st.max(), em[em > 0].min() if (em > 0).any() else -1, em.max(), en.min(), en.max(), nxt))
f.close()
