#!/usr/bin/env python
"""How far is real traffic from the packed kernel's proven range?  CPU only, instrumented oracle.
For the bench's 64QAM symbol workload at several noise levels: per code block and half-iteration, G = max(bx, M) + M of the
range condition, the oracle's saturation-clip count and the peak |intermediate| of the reference arithmetic."""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from _libs import P, oracle  # noqa: E402

synth = importlib.import_module("empower-srslte_b200.synth")
orc = oracle()
orc.orc_dbg_sat_count.restype = C.c_long


class Tdec(C.Structure):
    K_ = 6144
    _fields_ = [("K", C.c_int), ("P", C.c_int), ("n_iter", C.c_int),
                ("syst", C.c_int16 * (K_ + 4)), ("p0", C.c_int16 * (K_ + 4)), ("p1", C.c_int16 * (K_ + 4)),
                ("t_syst", C.c_int16 * 3), ("t_p0", C.c_int16 * 3), ("t_app2", C.c_int16 * 3), ("t_p1", C.c_int16 * 3),
                ("app1", C.c_int16 * K_), ("app2", C.c_int16 * K_), ("ext1", C.c_int16 * K_), ("ext2", C.c_int16 * K_),
                ("fwd", C.c_uint16 * K_), ("rev", C.c_uint16 * K_)]


def run(sigma_sym, ntb=2, mode="sym", amp=100.0, sigma=0.45, max_its=8, seed=1):
    TBS, QM, NB = 75376, 6, 90000
    rng = np.random.default_rng(seed)
    payload = rng.integers(0, 256, (ntb, TBS // 8), dtype=np.uint8)
    e = synth.encode_tbs(payload, TBS, QM, 0, NB)
    seg = synth.cbsegm(TBS); Cn, K = seg["C"], seg["K1"]
    rows = []
    for t in range(ntb):
        if mode == "sym":
            ci = synth.cinit_pdsch(0x1000 + t, 0, 2, 1)
            c = synth.gold_sequence(ci, NB)
            sym = synth.awgn_symbols(synth.modulate(e[t] ^ c, QM), sigma_sym, rng).astype(np.float32)
            llr = np.zeros(NB, np.int16)
            orc.orc_demod_s(QM, P(sym), P(llr), NB // QM)
            cc = np.ascontiguousarray(c.astype(np.uint8))
            orc.orc_descramble_s(P(cc), P(llr), NB)
        else:
            llr = synth.awgn_llr(e[t], sigma, rng, amp)
        Gp = NB // QM; gamma = Gp % Cn; n_e = QM * (Gp // Cn)
        for cb in range(Cn):
            rp, E = cb * n_e, n_e
            if cb > Cn - gamma:
                E = n_e + QM; rp = (Cn - gamma) * n_e + (cb - (Cn - gamma)) * E
            soft = np.zeros(18600, np.int16)
            orc.orc_rm_rx(P(np.ascontiguousarray(llr[rp:rp + E])), E, K, 0, 1, P(soft))
            M = int(np.abs(soft.astype(np.int32)).max())
            h = Tdec()
            orc.orc_tdec_new_cb_soft(C.byref(h), K, P(soft))
            out = np.zeros(K // 8, np.uint8)
            its, sat_first, Gs, peaks, spr = 0, None, [], [], []
            emax = 0
            for n in range(max_its):
                bx = emax + M if n % 2 == 0 else emax
                if n == 0: bx = M
                Gs.append(max(bx, M) + M)
                orc.orc_dbg_reset()
                orc.orc_tdec_iteration(C.byref(h), P(out), None)
                its += 1
                sc = orc.orc_dbg_sat_count()
                peaks.append(orc.orc_dbg_peak_val())
                spr.append([orc.orc_dbg_spread(i) for i in range(4)])
                if sc and sat_first is None: sat_first = n
                # next apriori = LLR - given: decoder 1 leaves ext1 (= LLR), app of decoder 2 = ext1 - app1
                if n % 2 == 0:
                    a = np.frombuffer(h.ext1, np.int16)[:K].astype(np.int32) - (np.frombuffer(h.app1, np.int16)[:K].astype(np.int32) if n else 0)
                else:
                    a = np.frombuffer(h.app1, np.int16)[:K].astype(np.int32) - np.frombuffer(h.ext1, np.int16)[:K].astype(np.int32)
                emax = int(np.abs(a).max())
                if orc.orc_crc(1, P(out), K) == 0:
                    break
            rows.append(dict(M=M, its=its, sat_first=sat_first, G=Gs, peak=peaks, spr=spr))
    return rows


def summarise(tag, rows):
    n = len(rows)
    its = np.array([r["its"] for r in rows])
    fast = sum(1 for r in rows if max(r["G"]) <= 2000)
    nosat = sum(1 for r in rows if r["sat_first"] is None)
    pk = np.array([max(r["peak"]) for r in rows])
    gm = np.array([max(r["G"]) for r in rows])
    print(f"{tag}: {n} CBs, mean half-its {its.mean():.2f}, M median {int(np.median([r['M'] for r in rows]))}, "
          f"inside G<=2000 for all half-its: {fast / n:.2f}, reference never saturates: {nosat / n:.2f}, "
          f"max G median {int(np.median(gm))} p90 {int(np.percentile(gm, 90))}, peak|intermediate| median {int(np.median(pk))} max {int(pk.max())}")
    # run-time certificate candidates: Sa + Sb + c*G <= 32767 with the spreads measured at the reference's normalisation points
    for c in (4, 7):
        okc = sum(1 for r in rows if all(max(s[1], s[3]) + max(s[0], s[1]) + 0 * g + max(s[2], s[3]) * 0 + c * g <= 32767 and s[3] + s[1] + c * g <= 32767
                                         for s, g in zip(r["spr"], r["G"])))
        print(f"    certificate Sa+Sb+{c}G<=32767 holds for all half-its: {okc / n:.2f}")
    allsp = np.array([[s + [g] for s, g in zip(r["spr"], r["G"])][-1] for r in rows])
    print(f"    last half-it: spreads (bw, b, aw, a) median {np.median(allsp[:, :4], axis=0).astype(int)} max {allsp[:, :4].max(axis=0)}; G median {int(np.median(allsp[:, 4]))} max {allsp[:, 4].max()}")
    for k in (2, 3, 4, 6, 8):
        sel = [r for r in rows if r["its"] >= k]
        if sel:
            print(f"    blocks reaching half-it {k}: {len(sel)}; G at its start median {int(np.median([r['G'][k - 1] for r in sel]))}; "
                  f"peak there median {int(np.median([r['peak'][k - 1] for r in sel]))} max {max(r['peak'][k - 1] for r in sel)}")


if __name__ == "__main__":
    for s in (0.062, 0.075, 0.082, 0.088):
        summarise(f"64QAM symbols sigma_sym={s}", run(s))
    for amp in (100.0, 250.0):
        summarise(f"int16 LLR amp={amp} sigma=0.45", run(0, mode="llr", amp=amp))
