#!/usr/bin/env python
"""bench.py — throughput of the DL-SCH transport-block decode path on B200 (contract: see the task brief).

One "step" = one pass of the hot path (rate de-matching + HARQ combine -> turbo decode with CRC24B early
stop -> TB assembly + CRC24A) over one batch of synthetic transport blocks of BASELINE.json configs[1]:
20 MHz SISO MCS28, TBS 75376, 13 code blocks of K=5824, 90000 int16 LLRs per TB, rv 0, max 8 half-iterations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch TBs/GPU] [--sigma S]

`value`   : decoded information Gbit/s (TBS bits of CRC-passing TBs / device time), inputs resident in HBM.
`e2e`     : same metric through the C ABI with HOST (pinned) buffers, H2D/D2H inside the timed region.
`roofline`: MAP-decoder kernel, algorithmic int16 ops / its CUDA-event duration vs ALU peak,
            plus the HBM figure for the same launch.
`cpu_baseline`: the reference's own AVX2 path (oracle/_ref) on the box's host cores, bounded sample.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TBS, QM, NBITS, RV, MAX_ITS = 75376, 6, 90000, 0, 8
K_CB, C_CB, P_CB = 5824, 13, 16
DATA_STRIDE = 10240            # >= TBS/8 + 3 + 768 bytes of room per TB output
METRIC = "decoded DL-SCH Gbps & code blocks/s per B200 (20MHz MCS28, 8 it)"
UNIT = "Gbit/s"


NCU_DRAM_BYTES_PER_CB = 47705          # measured, see roofline.traffic_source


def ops_per_halfit(K, P):
    """algorithmic int16 lane-ops of one half-iteration"""
    return 84 * K + 2040 * P + 63 if P else 90 * K


def algorithmic_bytes_per_cb(K, E):
    """HBM bytes the unfused pipeline must move per code block with new data: LLRs in, soft-buffer write,
    decoder read of the soft buffer, decoded bytes out"""
    return 2 * E + 2 * (3 * K + 12) + 2 * (3 * K + 12) + K // 8


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(hbm_gbs=float(j["hbm_gbs"]), sm_max_mhz=float(j.get("sm_max_mhz", 1965.0)), src="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, src="fallback")


def alu_peak_tops(n_sm, sm_mhz):
    """int16 ALU peak in Tops/s: SMs x int16 lane-ops/clk/SM x clock. The per-SM rate is the measured
    register-only packed s16x2 rate from tools/alu_peak.cu when profiles/alu_peak.json holds it,
    else the nominal 128 lanes x 2."""
    per_clk, src = 256.0, "nominal 128 lanes x s16x2"
    p = os.path.join(ROOT, "profiles", "alu_peak.json")
    if os.path.exists(p):
        j = json.load(open(p))
        per_clk, src = float(j["int16_ops_per_clk_per_sm"]), j.get("how", "measured microbenchmark")
    return n_sm * per_clk * sm_mhz * 1e6 / 1e12, src


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md clocks line)"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for r in self.rows)]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=max(mx) if mx else None,
                    power_w_max=max(pw) if pw else None, samples=len(self.rows), reasons=reasons)


def make_llrs(batch, sigma, seed, out):
    """batch AWGN-corrupted MCS28 transport blocks -> out[batch, NBITS] int16; returns payload[ncw, TBS/8].
    16 distinct codewords (numpy encoder, empower-srslte_b200/synth.py), independent noise per TB."""
    synth = importlib.import_module("empower-srslte_b200.synth")
    ncw = min(16, batch)
    rng = np.random.default_rng(seed)
    payload = rng.integers(0, 256, size=(ncw, TBS // 8), dtype=np.uint8)
    e = synth.encode_tbs(payload, TBS, QM, RV, NBITS)
    for i0 in range(0, batch, 64):
        i1 = min(batch, i0 + 64)
        synth.awgn_llr(e[np.arange(i0, i1) % ncw], sigma, rng, out=out[i0:i1])
    return payload


def fnv1a_rows(a):
    """FNV-1a (32 bit) of every row of a uint8 matrix — the checksum the CPU arm computes over its decoded bytes"""
    h = np.full(a.shape[0], 2166136261, np.uint64)
    for j in range(a.shape[1]):
        h = ((h ^ a[:, j].astype(np.uint64)) * np.uint64(16777619)) & np.uint64(0xFFFFFFFF)
    return h.astype(np.uint32)


def bind_to_gpu_cpus(index):
    """sched_setaffinity to the CPUs NVML reports as local to GPU `index` (nvmlDeviceGetCpuAffinity); returns what was done"""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, v in enumerate(words) for b in range(64) if (int(v) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        use = sorted(cpus & allowed)
        if not use:
            return dict(bound=False, why="no NVML-local CPU in the allowed set", allowed=len(allowed))
        os.sched_setaffinity(0, use)
        return dict(bound=True, cpus=f"{use[0]}-{use[-1]}", n_cpus=len(use), allowed_before=len(allowed))
    except Exception as e:                       # noqa: BLE001
        return dict(bound=False, why=repr(e)[:120])


def h2d_link_peak(torch, device, barrier, mb=256, reps=8):
    """what the host delivers to THIS GPU while every rank does the same: pinned 256 MB buffers, cudaMemcpyAsync H2D back to back
    (tools/h2d_peak.cu is the standalone version). GB/s; the denominator of the end-to-end figures."""
    h = torch.empty(mb << 20, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(mb << 20, dtype=torch.uint8, device=device)
    d.copy_(h, non_blocking=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    e1.record()
    e1.synchronize()
    barrier()
    return (mb << 20) * reps / (e0.elapsed_time(e1) / 1e3) / 1e9


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_arm(llr, ntb, budget_s):
    """the reference's CPU implementation of the same step on this box's host cores (bounded sample).
    oracle/_ref (unmodified reference objects, AVX2 AUTO decoders) when present, else the scalar oracle port."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _libs import P, oracle, ref
    cores = cpu_cores()
    lib = ref()
    kind = "reference" if lib is not None else "port"
    if lib is None:
        lib = oracle()
    ntb = min(ntb, llr.shape[0])
    sample = np.ascontiguousarray(llr[:ntb])
    out = np.zeros(3, np.uint64)
    res = np.zeros((ntb, 3), np.uint32)       # per TB of the last pass: return value, mean half-iterations (sch.c:420), FNV-1a of the output
    (lib.ref_bench_set_results if kind == "reference" else lib.orc_bench_set_results)(P(res))

    def run(reps):
        if kind == "reference":
            return lib.ref_bench_tb_mt(cores, ntb, TBS, QM, RV, NBITS, P(sample), MAX_ITS, reps, 100, P(out))
        return lib.orc_bench_tb_mt(cores, ntb, TBS, QM, RV, NBITS, P(sample), MAX_ITS, reps, P(out))

    t1 = run(1)
    reps = max(1, min(2000, int(budget_s / max(t1, 1e-3))))
    t = run(reps)
    n_tb, n_ok, sum_its = int(out[0]), int(out[1]), int(out[2])
    (lib.ref_bench_set_results if kind == "reference" else lib.orc_bench_set_results)(None)
    return dict(results=res, value=n_ok * TBS / t / 1e9, unit=UNIT, cores=cores, kind=kind, seconds=t,
                tbs_per_s=n_tb / t, cbs_per_s=n_tb * C_CB / t, tb_ok=n_ok, tb_total=n_tb,
                mean_halfits=sum_its / max(n_tb, 1),
                sample=f"{ntb} of the step's TBs x {reps} passes, {cores} threads (one srslte_sch_t + soft buffers each), "
                       f"srslte_dlsch_decode AUTO AVX2" if kind == "reference" else
                       f"{ntb} of the step's TBs x {reps} passes, {cores} threads, scalar C port of the reference path")


def reference_main(args, rank, world):
    """--impl reference: the reference's own CPU path, rank 0 only"""
    if rank != 0:
        return
    ntb = max(cpu_cores() * 4, 32)
    llr = np.empty((ntb, NBITS), np.int16)
    make_llrs(ntb, args.sigma, 1234, llr)
    per_step_budget = min(20.0, args.ref_seconds / max(args.steps + args.warmup, 1))
    vals = []
    last = None
    for s in range(args.warmup + args.steps):
        last = cpu_arm(llr, ntb, per_step_budget)
        if s >= args.warmup:
            vals.append(last)
    tot_bits = sum(v["tb_ok"] for v in vals) * TBS
    tot_s = sum(v["seconds"] for v in vals)
    val = tot_bits / tot_s / 1e9
    line = dict(impl="reference", metric=METRIC, value=val, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * tot_s / max(len(vals), 1), higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="int16", data="synthetic",
                config=dict(workload="pdsch_mcs28_100prb_tbs75376_13xK5824", sigma=args.sigma, max_half_iterations=MAX_ITS,
                            tbs_per_step=last["tb_total"], early_stop=True, host_only=True),
                cbs_per_s=sum(v["tb_total"] for v in vals) * C_CB / tot_s,
                mean_halfits=last["mean_halfits"],
                cpu_baseline=dict(value=val, unit=UNIT, cores=last["cores"], kind=last["kind"], sample=last["sample"]),
                e2e=dict(value=val, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def fill_descs(pkg, d, n, e_ptr0, data_ptr0, sb0, new_data=1):
    for i in range(n):
        d[i].tbs, d[i].Qm, d[i].rv, d[i].nof_e_bits, d[i].max_its = TBS, QM, RV, NBITS, MAX_ITS
        d[i].softbuffer, d[i].new_data = sb0 + i, new_data
        d[i].e_bits, d[i].data = e_ptr0 + i * NBITS * 2, data_ptr0 + i * DATA_STRIDE


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps (default 40; 10 for --impl reference, whose step is ~4.6 s of CPU work)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=4736, help="transport blocks per GPU per step (default 32 per SM = 61 568 code blocks)")
    ap.add_argument("--sigma", type=float, default=0.45, help="AWGN std of the synthetic LLRs (0.45 = operating point)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-bind", action="store_true", help="do not bind the rank to the CPUs local to its GPU")
    ap.add_argument("--no-worst-case", action="store_true")
    ap.add_argument("--no-uplink", action="store_true", help="skip the 64-cell PUSCH leg (BASELINE configs[4])")
    ap.add_argument("--ul-subframes", type=int, default=37, help="subframes per cell and step of the 64-cell PUSCH leg (64 x 37 = 2368 TBs)")
    ap.add_argument("--no-harq", action="store_true", help="skip the HARQ leg (BASELINE configs[2])")
    ap.add_argument("--sigma-harq", type=float, default=0.47, help="AWGN std of the HARQ leg (rv 0 mostly fails)")
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of the sustained leg (0 = skip)")
    ap.add_argument("--ref-seconds", type=float, default=150.0, help="--impl reference: CPU time budget of the whole run (bounded samples per step)")
    ap.add_argument("--no-llr8", action="store_true", help="skip the 8-bit LLR leg")
    ap.add_argument("--no-grid", action="store_true", help="skip the sigma grid (int16 and int8 paths at sigma 0.30 / 0.40 / 0.45)")
    ap.add_argument("--sigma8", type=float, default=0.30, help="AWGN std of the 8-bit LLR leg")
    ap.add_argument("--sigma-ul", type=float, default=0.17, help="AWGN std per real dimension of the synthetic 16QAM PUSCH symbols")
    ap.add_argument("--no-tx", action="store_true", help="skip the transmit-side leg (turbo encoder + rate matching on the device)")
    ap.add_argument("--no-symbols", action="store_true", help="skip the leg that starts from equalised symbols (front end)")
    ap.add_argument("--sigma-sym", type=float, default=0.062, help="AWGN std per real dimension of the synthetic 64QAM symbols")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.steps is None:
        args.steps = 10 if args.impl == "reference" else 40
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return reference_main(args, rank, world)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pkg = importlib.import_module("empower-srslte_b200")
    shard = importlib.import_module("empower-srslte_b200.shard")
    # run this rank (and first-touch its pinned staging) on the CPUs next to its GPU: with 8 ranks pulling 50 GB/s each out of
    # host memory the NUMA placement of the pinned buffers is part of the end-to-end figure
    affinity = bind_to_gpu_cpus(local_rank) if not args.no_bind else None
    fec = pkg.FecCuda(local_rank, max_tbs=args.batch)
    B = args.batch

    # ---- synthetic workload: host pinned LLRs (e2e arm) and a device-resident copy (kernel arm) ----
    h_llr = pkg.PinnedArray((B, NBITS), np.int16)
    h_out = pkg.PinnedArray((B, DATA_STRIDE), np.uint8)
    payload = make_llrs(B, args.sigma, 1000 + rank, h_llr.a)
    d_llr = torch.from_numpy(h_llr.a).to(f"cuda:{local_rank}")
    # two batch sets (A/B): same read-only LLRs, own soft buffers and outputs, so that two batches can be in flight
    d_out = [torch.zeros((B, DATA_STRIDE), dtype=torch.uint8, device=f"cuda:{local_rank}") for _ in range(2)]
    h_out2 = pkg.PinnedArray((B, DATA_STRIDE), np.uint8)
    sb0 = fec.softbuffer_alloc(2 * B)
    dd_ab, hd_ab = [], []
    for s_ in range(2):
        d_ = (pkg.TbDesc * B)()
        fill_descs(pkg, d_, B, d_llr.data_ptr(), d_out[s_].data_ptr(), sb0 + s_ * B)
        dd_ab.append(d_)
        h_ = (pkg.TbDesc * B)()
        fill_descs(pkg, h_, B, h_llr.a.ctypes.data, (h_out if s_ == 0 else h_out2).a.ctypes.data, sb0 + s_ * B)
        hd_ab.append(h_)
    dd, hd = dd_ab[0], hd_ab[0]
    stream = torch.cuda.ExternalStream(fec.stream, device=f"cuda:{local_rank}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(descs, flags, steps, phases=None, f=None, st=None):
        """time exactly `steps` SYNCHRONOUS passes (one fec_cuda_decode_tbs call each); max over ranks"""
        f, st = f or fec, st or stream
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(steps):
            f.decode_tbs(descs, flags)
            if phases is not None:
                phases.append(f.last_phase_ms())
        e1.record(st)
        barrier()
        ms, _ = shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", e0.elapsed_time(e1))
        return ms

    def timed_pipelined(sets, flags, steps, f=None, st=None):
        """time exactly `steps` passes with two batches in flight (fec_cuda_submit_tbs / fec_cuda_wait, alternating
        batch sets A/B): step i+1 is submitted before step i is collected, every step is collected inside the timed
        region; max over ranks"""
        f, st = f or fec, st or stream
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        prev = None
        for i in range(steps):
            t = f.submit_tbs(sets[i & 1], flags)
            if prev is not None:
                f.wait(prev)
            prev = t
        f.wait(prev)
        e1.record(st)
        barrier()
        ms, _ = shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", e0.elapsed_time(e1))
        return ms

    def tally(descs):
        ok = sum(1 for i in range(B) if descs[i].ret == 0)
        its = sum(sum(descs[i].cb_its[:C_CB]) for i in range(B))
        return ok, its

    def allsum(v):
        return shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", 0.0, v=v)[1]["v"]

    # ---- device-resident arm (value) ----
    for i in range(max(args.warmup, 6)):
        fec.decode_tbs(dd_ab[i & 1], pkg.F_DEVICE_IO)
    l0 = fec.launches
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    ms = timed_pipelined(dd_ab, pkg.F_DEVICE_IO, args.steps)
    launches = fec.launches - l0
    ms_sync = timed(dd, pkg.F_DEVICE_IO, args.steps)
    t0 = time.perf_counter()                     # host side of one step: argument checks, segmentation, descriptors, launches
    tk_ = fec.submit_tbs(dd, pkg.F_DEVICE_IO)
    host_submit_ms = (time.perf_counter() - t0) * 1e3
    fec.wait(tk_)
    ok, its = tally(dd)
    # parity spot check of what was decoded: payload bytes of every CRC-passing TB
    out_host = d_out[0].cpu().numpy()
    bad = sum(1 for i in range(B) if dd[i].ret == 0 and not (out_host[i, :TBS // 8] == payload[i % payload.shape[0]]).all())
    if bad:
        raise SystemExit(f"bench.py: {bad} CRC-passing transport blocks differ from the transmitted payload")
    ok_all, its_all = allsum(ok), allsum(its)
    sec = ms / 1e3
    value = ok_all * TBS * args.steps / sec / 1e9
    cbs_per_s = world * B * C_CB * args.steps / sec

    # ---- end-to-end arm: host pinned buffers through the C ABI, H2D + D2H inside the timed region ----
    for i in range(6):
        fec.decode_tbs(hd_ab[i & 1], 0)
    ms_e = timed_pipelined(hd_ab, 0, args.steps)
    ms_e_sync = timed(hd, 0, args.steps)
    clocks = sampler.stop()
    ok_e, _ = tally(hd)
    n_par = min(max(cpu_cores() * 4, 32), B)     # the transport blocks the CPU arm decodes too (parity_in_bench)
    gpu_par = np.stack([np.array([hd[i].ret & 0xFFFFFFFF for i in range(n_par)], np.uint32),
                        np.array([hd[i].avg_its for i in range(n_par)], np.uint32),
                        fnv1a_rows(h_out.a[:n_par, :TBS // 8 + 3])], axis=1)
    e2e_val = allsum(ok_e) * TBS * args.steps / (ms_e / 1e3) / 1e9
    link = h2d_link_peak(torch, f"cuda:{local_rank}", barrier)
    link_min = -shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", -link)[0]      # slowest rank's link (MAX of the negated rates)
    link_sum = allsum(link)
    h2d = B * NBITS * 2 + B * C_CB * (32 + 48) + B * 520
    d2h = B * (((TBS // 8 + 3 + 768 + 15) // 16) * 16) + B * C_CB * 8 + B * 4

    # ---- roofline of the dominant kernel (turbo MAP decoder), per launch: the same step un-pipelined (one launch per
    #      kernel, FEC_CUDA_F_SERIAL), so that the CUDA events around each kernel do not overlap with a neighbour ----
    pk = peaks()
    n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
    fec.decode_tbs(dd, pkg.F_DEVICE_IO | pkg.F_SERIAL)
    phases = []
    for _ in range(max(3, min(args.steps, 10))):
        fec.decode_tbs(dd, pkg.F_DEVICE_IO | pkg.F_SERIAL)
        phases.append(fec.last_phase_ms())
    ok, its = tally(dd)
    dec_ms = sum(p[1] for p in phases) / len(phases)
    rm_ms = sum(p[0] for p in phases) / len(phases)
    asm_ms = sum(p[2] for p in phases) / len(phases)
    ops = its * ops_per_halfit(K_CB, P_CB)                     # this rank, one launch
    achieved = ops / (dec_ms / 1e3) / 1e12
    sm_mhz = clocks["sm_mhz"] or pk["sm_max_mhz"]
    peak_nom, peak_src = alu_peak_tops(n_sm, pk["sm_max_mhz"])
    peak_run, _ = alu_peak_tops(n_sm, sm_mhz)
    hbm_bytes = B * C_CB * algorithmic_bytes_per_cb(K_CB, NBITS // C_CB)
    roof = dict(bound="alu_int16", kernel="fec_tdec_pk_kernel<16,true>", achieved=achieved, peak=peak_nom, unit="Tops/s",
                frac=achieved / peak_nom, frac_at_sustained_clock=achieved / peak_run, peak_source=peak_src + f" @ {pk['sm_max_mhz']:.0f} MHz",
                traffic=NCU_DRAM_BYTES_PER_CB * B * C_CB, traffic_source="ncu --set full of one 30 784-CB launch (profiles/r2_map_kernel_ncu.txt): "
                "dram__bytes_read.sum + dram__bytes_write.sum = 1443.8 + 24.7 MB = 47.7 KB per code block, scaled to this launch",
                ops_per_launch=ops, ms_per_launch=dec_ms,
                # the packed add / add-max instructions only exist on the ALU pipe (half rate: 64 lanes/clk/SM); the peak above also
                # counts the FMA pipe (IMAD), which this instruction mix cannot use. Against the ALU pipe alone:
                alu_pipe_only=dict(peak=peak_nom * 64.0 / 124.3, frac=achieved / (peak_nom * 64.0 / 124.3),
                                   ncu="profiles/r2_map_kernel_ncu.txt (30 784-CB launch): sm__inst_executed_pipe_alu 41 % of peak, issue slots 47.5 % busy; "
                                       "stall samples: 47.6 % issuing, 25.6 % fixed-latency wait (dependent issue, second cycle of the half-rate packed "
                                       "instructions), 8.8 % global loads, 5.5 % instruction fetch, 4.8 % dispatch, 3.9 % shared-memory loads, 1.6 % final barrier"),
                share_of_step=dec_ms / (rm_ms + dec_ms + asm_ms), serial_step_ms=rm_ms + dec_ms + asm_ms,
                phases_ms=dict(rate_dematch=rm_ms, turbo_decode=dec_ms, tb_assemble_crc24a=asm_ms),
                hbm=dict(bound="hbm", algorithmic_bytes_per_step=hbm_bytes, achieved=hbm_bytes / (ms / args.steps / 1e3) / 1e9,
                         peak=pk["hbm_gbs"], unit="GB/s", frac=hbm_bytes / (ms / args.steps / 1e3) / 1e9 / pk["hbm_gbs"],
                         peak_source=pk["src"],
                         # the rate de-matcher alone, new data: E LLRs in, the slot (3*SS+16 int16, SS = P*Lp) written once
                         rate_dematch_kernel=dict(ms_per_launch=rm_ms, bytes_per_cb=2 * (NBITS // C_CB) + 2 * (3 * P_CB * ((K_CB // P_CB + 7) // 8 * 8) + 16),
                                                  achieved=B * C_CB * (2 * (NBITS // C_CB) + 2 * (3 * P_CB * ((K_CB // P_CB + 7) // 8 * 8) + 16)) / (rm_ms / 1e3) / 1e9,
                                                  unit="GB/s", peak=pk["hbm_gbs"],
                                                  frac=B * C_CB * (2 * (NBITS // C_CB) + 2 * (3 * P_CB * ((K_CB // P_CB + 7) // 8 * 8) + 16)) / (rm_ms / 1e3) / 1e9 / pk["hbm_gbs"])))

    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        c = cpu_arm(h_llr.a, max(cpu_cores() * 4, 32), 15.0)
        cpu = dict(value=c["value"], unit=UNIT, cores=c["cores"], kind=c["kind"], sample=c["sample"],
                   cbs_per_s=c["cbs_per_s"], mean_halfits=c["mean_halfits"], tb_ok=c["tb_ok"], tb_total=c["tb_total"])
        # the CPU arm decoded the first n_par transport blocks of the step: the GPU's results for the same blocks (e2e arm, host
        # buffers) must be the same — return value, mean half-iterations as srslte_sch_last_noi reports them, every output byte
        n_cmp = min(n_par, c["results"].shape[0])
        same = (c["results"][:n_cmp] == gpu_par[:n_cmp]).all(axis=1)
        parity = dict(tbs=int(n_cmp), equal=int(same.sum()), checker=c["kind"], compared="decode_tb return value, mean half-iterations (sch.c:420), FNV-1a of TBS/8+3 output bytes")
        if not same.all():
            raise SystemExit(f"bench.py: parity_in_bench failed: {parity}, first mismatch TB {int(np.argmin(same))}: "
                             f"cpu {c['results'][int(np.argmin(same))].tolist()} gpu {gpu_par[int(np.argmin(same))].tolist()}")

    # ---- the other legs. A failure in one of them (allocation, a driver hiccup) must not cost the run its headline: the error is
    #      reported in the line (`aux_error`); a payload or parity mismatch is a SystemExit and still ends the run ----
    front = ulsch5 = harq = sustained = tx = llr8 = lat = worst = None
    aux_error = None
    try:
        # ---- front-end leg: the same transport blocks given as equalised 64QAM symbols (8 B per resource element
        #      instead of 12 B of LLRs); soft demodulation + descrambling run in the fused front-end kernel on the device ----
        front, ulsch5 = None, None
        if not args.no_symbols:
            synth = importlib.import_module("empower-srslte_b200.synth")
            nre = NBITS // QM
            h_sym = pkg.PinnedArray((B, nre, 2), np.float32)
            rng = np.random.default_rng(5000 + rank)
            ncw = payload.shape[0]
            c_inits = [synth.cinit_pdsch(0x1000 + k, 0, 2 * (k % 10), 1) for k in range(ncw)]
            e = synth.encode_tbs(payload, TBS, QM, RV, NBITS)
            cw_sym = np.stack([synth.modulate(e[k] ^ synth.gold_sequence(c_inits[k], NBITS), QM) for k in range(ncw)])
            for i0 in range(0, B, 64):
                i1 = min(B, i0 + 64)
                h_sym.a[i0:i1] = synth.awgn_symbols(cw_sym[np.arange(i0, i1) % ncw], args.sigma_sym, rng)
            d_sym = torch.from_numpy(h_sym.a).to(f"cuda:{local_rank}")
            sy_ab = []
            for dev in (True, False):
                sets = []
                for s_ in range(2):
                    sd = (pkg.SymDesc * B)()
                    base = d_sym.data_ptr() if dev else h_sym.a.ctypes.data
                    for i in range(B):
                        sd[i].symbols, sd[i].nof_re, sd[i].mod_bits, sd[i].c_init = base + i * nre * 8, nre, QM, c_inits[i % ncw]
                    sets.append(sd)
                sy_ab.append(sets)

            def timed_syms(tb_sets, sym_sets, flags, steps, f=None, st=None):
                f, st = f or fec, st or stream
                barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                prev = None
                for i in range(steps):
                    t = f.submit_syms(tb_sets[i & 1], sym_sets[i & 1], flags)
                    if prev is not None:
                        f.wait(prev)
                    prev = t
                f.wait(prev)
                e1.record(st)
                barrier()
                return shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", e0.elapsed_time(e1))[0]

            for i in range(6):
                fec.decode_syms(dd_ab[i & 1], sy_ab[0][i & 1], pkg.F_DEVICE_IO)
            ms_sd = timed_syms(dd_ab, sy_ab[0], pkg.F_DEVICE_IO, args.steps)
            ok_s, its_s = tally(dd)
            fec.decode_syms(dd, sy_ab[0][0], pkg.F_DEVICE_IO | pkg.F_SERIAL)
            fe_ms = fec.last_frontend_ms()
            ph_s = fec.last_phase_ms()
            path_s = fec.last_path_counts()
            out_host = d_out[0].cpu().numpy()
            bad = sum(1 for i in range(B) if dd[i].ret == 0 and not (out_host[i, :TBS // 8] == payload[i % ncw]).all())
            if bad:
                raise SystemExit(f"bench.py: {bad} CRC-passing transport blocks of the symbol leg differ from the transmitted payload")
            for i in range(6):
                fec.decode_syms(hd_ab[i & 1], sy_ab[1][i & 1], 0)
            ms_sh = timed_syms(hd_ab, sy_ab[1], 0, args.steps)
            ok_sh, _ = tally(hd)
            fe_bytes = B * (nre * 8 + NBITS * 2)
            front = dict(workload="same TBs as 64QAM symbols (15000 RE), soft demodulation + descrambling on the device",
                         sigma_sym=args.sigma_sym, value=allsum(ok_s) * TBS * args.steps / (ms_sd / 1e3) / 1e9, unit=UNIT,
                         ms_per_step=ms_sd / args.steps, tb_ok_frac=allsum(ok_s) / (world * B), halfits_per_cb=allsum(its_s) / (world * B * C_CB),
                         e2e=dict(value=allsum(ok_sh) * TBS * args.steps / (ms_sh / 1e3) / 1e9, unit=UNIT, ms_per_step=ms_sh / args.steps,
                                  h2d_bytes_per_step=B * nre * 8 + B * C_CB * (32 + 48) + B * (520 + 80), d2h_bytes_per_step=d2h,
                                  frac_of_h2d_link=(B * nre * 8 + B * C_CB * (32 + 48) + B * (520 + 80)) / (ms_sh / args.steps / 1e3) / 1e9 / link_min),
                         code_blocks_by_kernel=dict(total=B * C_CB, certificate_variant=path_s[0], exact=path_s[1]),
                         serial_phases_ms=dict(frontend=fe_ms, rate_dematch=ph_s[0], turbo_decode=ph_s[1], tb_assemble_crc24a=ph_s[2]),
                         frontend_kernel=dict(ms_per_launch=fe_ms, algorithmic_bytes=fe_bytes, achieved=fe_bytes / (fe_ms / 1e3) / 1e9,
                                              peak=peaks()["hbm_gbs"], unit="GB/s", frac=fe_bytes / (fe_ms / 1e3) / 1e9 / peaks()["hbm_gbs"], bound="hbm"))
            # ---- uplink leg (BASELINE configs[4] shape per cell-subframe): PUSCH from the demodulator on — 16QAM (what
            #      srsENB grants, phch_worker.cc:598-601), 96 PRB, TBS 36696 = 6 x K=6144, HARQ-ACK / RI / CQI multiplexed: soft
            #      demodulation, ACK zeroing, descrambling, channel de-interleaver with RI removal, then decode_tb on g_bits ----
            if not args.no_uplink:
                # BASELINE configs[4]: "multi-cell eNB UL-SCH decode: 64 cells x 20 MHz ... sharded across 1/2/4/8 B200". The 64 cells
                # are placed with shard.owner (cell-affine: every HARQ process of a cell lives on ONE GPU, no soft buffer ever moves,
                # no collective); this rank decodes args.ul_subframes subframes of each of ITS cells per step. Strong scaling: the
                # job (64 cells) is fixed, the per-rank share shrinks with N.
                ACK_COLS, RI_COLS, uci_positions = synth.ACK_COLS, synth.RI_COLS, synth.uci_positions
                U_TBS, U_MOD, U_PRB, U_COLS, U_CQI, U_C, N_CELLS = 36696, 4, 96, 12, 32, 6, 64
                u_rows, u_nre = 12 * U_PRB, 12 * U_PRB * U_COLS
                ack_pos = np.array(uci_positions(8, U_MOD, u_rows, ACK_COLS), np.uint32)
                ri_pos = np.array(uci_positions(4, U_MOD, u_rows, RI_COLS), np.uint32)
                U_G = u_nre * U_MOD - ri_pos.size - U_CQI
                keys = [(c, 0) for c in range(N_CELLS)]
                my_cells = [keys[i][0] for i in shard.partition(keys, rank, world)]
                nsf = args.ul_subframes
                nU = len(my_cells) * nsf                                  # transport blocks of this rank per step
                u_pay = np.random.default_rng(7000).integers(0, 256, (N_CELLS, U_TBS // 8), dtype=np.uint8)     # one codeword per cell
                u_e = synth.encode_tbs(u_pay[my_cells], U_TBS, U_MOD, 0, U_G)
                u_cq = np.random.default_rng(7001).integers(0, 2, (N_CELLS, U_CQI)).astype(u_e.dtype)[my_cells]
                u_q = synth.ulsch_interleave(np.concatenate([u_cq, u_e], axis=1), U_MOD, U_COLS, list(ri_pos))
                u_ci = [synth.cinit_pusch(0x46 + c, 2 * (c % 10), 1 + c) for c in my_cells]                     # cell id = c + 1
                u_clean = np.stack([synth.modulate(u_q[k] ^ synth.gold_sequence(u_ci[k], u_nre * U_MOD), U_MOD) for k in range(len(my_cells))])
                hu_sym = pkg.PinnedArray((nU, u_nre, 2), np.float32)
                rng_u = np.random.default_rng(7100 + rank)
                cell_of = np.repeat(np.arange(len(my_cells)), nsf)        # TB i belongs to this rank's cell cell_of[i]
                for i0 in range(0, nU, 64):
                    i1 = min(nU, i0 + 64)
                    hu_sym.a[i0:i1] = synth.awgn_symbols(u_clean[cell_of[i0:i1]], args.sigma_ul, rng_u)
                du_sym = torch.from_numpy(hu_sym.a).to(f"cuda:{local_rank}")
                n_uci = ack_pos.size + ri_pos.size + U_CQI
                du_uci = torch.zeros((2, nU, (n_uci + 7) & ~7), dtype=torch.int16, device=f"cuda:{local_rank}")
                du_out = torch.zeros((2, nU, DATA_STRIDE), dtype=torch.uint8, device=f"cuda:{local_rank}")
                sbu = fec.softbuffer_alloc(2 * nU)                        # per (cell, subframe) HARQ entity, two sets in flight
                u_tb, u_sy = [], []
                for s_ in range(2):
                    t_ = (pkg.TbDesc * nU)()
                    y_ = (pkg.SymDesc * nU)()
                    for i in range(nU):
                        t_[i].tbs, t_[i].Qm, t_[i].rv, t_[i].nof_e_bits, t_[i].max_its = U_TBS, U_MOD, 0, U_G, 8
                        t_[i].softbuffer, t_[i].new_data = sbu + s_ * nU + i, 1
                        t_[i].data = du_out[s_, i].data_ptr()
                        y_[i].symbols, y_[i].nof_re, y_[i].mod_bits, y_[i].c_init = du_sym.data_ptr() + i * u_nre * 8, u_nre, U_MOD, u_ci[cell_of[i]]
                        y_[i].nof_symb = U_COLS
                        y_[i].ack_pos, y_[i].ri_pos, y_[i].nof_ack_pos, y_[i].nof_ri_pos = ack_pos.ctypes.data, ri_pos.ctypes.data, ack_pos.size, ri_pos.size
                        y_[i].nof_cqi_bits = U_CQI
                        y_[i].uci_out = du_uci[s_, i].data_ptr()
                    u_tb.append(t_); u_sy.append(y_)
                for i in range(6):
                    fec.decode_syms(u_tb[i & 1], u_sy[i & 1], pkg.F_DEVICE_IO)
                ms_ul = timed_syms(u_tb, u_sy, pkg.F_DEVICE_IO, args.steps)
                t0 = time.perf_counter()                                  # host side of one step: descriptors, position lists, launches
                tk_ = fec.submit_syms(u_tb[0], u_sy[0], pkg.F_DEVICE_IO)
                host_submit_ul_ms = (time.perf_counter() - t0) * 1e3
                fec.wait(tk_)
                ok_u = sum(1 for i in range(nU) if u_tb[0][i].ret == 0)
                its_u = sum(sum(u_tb[0][i].cb_its[:U_C]) for i in range(nU))
                out_host = du_out[0].cpu().numpy()
                bad = sum(1 for i in range(nU) if u_tb[0][i].ret == 0 and not (out_host[i, :U_TBS // 8] == u_pay[my_cells[cell_of[i]]]).all())
                if bad:
                    raise SystemExit(f"bench.py: {bad} CRC-passing transport blocks of the uplink leg differ from the transmitted payload")
                fec.decode_syms(u_tb[0], u_sy[0], pkg.F_DEVICE_IO | pkg.F_SERIAL)
                path_u = fec.last_path_counts()
                tot_u = allsum(nU)
                ulsch5 = dict(workload="BASELINE configs[4]: 64 cells x PUSCH 16QAM 96 PRB, TBS 36696 (6 x K=6144), 32 HARQ-ACK + 16 RI positions + 32 CQI bits "
                                       "multiplexed; equalised symbols in (device-resident), UL front end + decode_tb", scaling="strong",
                              placement="shard.partition: cell c on rank c mod N with all of its HARQ soft buffers; no collective on the data path",
                              cells=N_CELLS, cells_this_rank=len(my_cells), subframes_per_cell_per_step=nsf, tbs_per_step=tot_u, sigma_sym=args.sigma_ul,
                              value=allsum(ok_u) * U_TBS * args.steps / (ms_ul / 1e3) / 1e9, unit=UNIT, ms_per_step=ms_ul / args.steps,
                              tb_ok_frac=allsum(ok_u) / tot_u, halfits_per_cb=allsum(its_u) / (tot_u * U_C),
                              cbs_per_s=tot_u * U_C * args.steps / (ms_ul / 1e3), payload_check="every CRC-passing TB of set 0 equals its cell's transmitted payload",
                              code_blocks_by_kernel_rank0=dict(total=nU * U_C, certificate_variant=path_u[0], exact=path_u[1]),
                              host_submit_ms_rank0=host_submit_ul_ms,
                              serial_phases_ms_rank0=dict(zip(("rate_dematch", "turbo_decode", "tb_assemble_crc24a"), fec.last_phase_ms()),
                                                          frontend=fec.last_frontend_ms()))
                del du_sym, du_uci, du_out
                hu_sym.free()
            # the LLR legs below reuse dd / hd: put the soft buffers back into the state of the LLR workload
            fec.decode_tbs(dd, pkg.F_DEVICE_IO)

        # ---- HARQ leg (BASELINE configs[2]: TM4 2-codeword shape, 8-process HARQ soft combining): TBS 75376 in 86400 LLRs (rate 0.876),
        #      13 x K=5824, rv sequence 0,2,3,1; noise so that rv 0 mostly fails. One step = one whole HARQ cycle of B/2 processes:
        #      round 0 decodes all of them (new data), every later round only the NACKed ones with new_data = 0 — the read-modify-write
        #      combine (softbuffer.c:128-151 state, rm_turbo.c:411-420), the cb_crc skip and the cached-CB copy (sch.c:396-418) are
        #      inside the timed region. Rounds of one process are sequential by nature (same soft buffer), so the calls are blocking.
        harq = None
        if not args.no_harq:
            synth_h = importlib.import_module("empower-srslte_b200.synth")
            H_NBITS, Bh = 86400, max(8, B // 2)
            rng_h = np.random.default_rng(9000 + rank)
            ncw = payload.shape[0]
            d_h = []
            tmp_h = np.empty((64, H_NBITS), np.int16)
            for rv in (0, 2, 3, 1):
                e_h = synth_h.encode_tbs(payload, TBS, QM, rv, H_NBITS)
                dev = torch.empty((Bh, H_NBITS), dtype=torch.int16, device=f"cuda:{local_rank}")
                for i0 in range(0, Bh, 64):
                    i1 = min(Bh, i0 + 64)
                    synth_h.awgn_llr(e_h[np.arange(i0, i1) % ncw], args.sigma_harq, rng_h, out=tmp_h[: i1 - i0])
                    dev[i0:i1] = torch.from_numpy(tmp_h[: i1 - i0]).to(dev.device)
                d_h.append(dev)

            def harq_round(r, live):
                d_ = (pkg.TbDesc * len(live))()
                for s_, i in enumerate(live):
                    d_[s_].tbs, d_[s_].Qm, d_[s_].rv, d_[s_].nof_e_bits, d_[s_].max_its = TBS, QM, (0, 2, 3, 1)[r], H_NBITS, MAX_ITS
                    d_[s_].softbuffer, d_[s_].new_data = sb0 + i, int(r == 0)
                    d_[s_].e_bits, d_[s_].data = d_h[r].data_ptr() + i * H_NBITS * 2, d_out[0].data_ptr() + i * DATA_STRIDE
                return d_

            # discovery pass (untimed): which processes are still NACKed after each round — deterministic, the timed passes repeat it.
            # The processes form two groups (a process never changes group): the rounds of ONE group are sequential by nature, the
            # other group's round is in flight meanwhile, as the PHY workers of an eNB overlap the HARQ processes of different UEs
            groups = []
            for g_ in range(2):
                rounds_g, live = [], list(range(g_ * (Bh // 2), Bh if g_ else Bh // 2))
                for r in range(4):
                    if not live:
                        break
                    d_ = harq_round(r, live)
                    fec.decode_tbs(d_, pkg.F_DEVICE_IO)
                    rounds_g.append((d_, list(live), [d_[s_].ret for s_ in range(len(live))]))
                    live = [i for s_, i in enumerate(live) if d_[s_].ret != 0]
                groups.append(rounds_g)
            seq_g = [(g_, groups[g_][r]) for r in range(4) for g_ in range(2) if r < len(groups[g_])]      # G0 r0, G1 r0, G0 r1, G1 r1, ...
            seq = [it for _, it in seq_g]
            rounds = seq
            for d_, _, _ in seq[:2]:
                fec.decode_tbs(d_, pkg.F_DEVICE_IO)                # warm
            # blocking reference figure: one call after the other
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(max(2, args.steps // 4)):
                for d_, _, _ in seq:
                    fec.decode_tbs(d_, pkg.F_DEVICE_IO)
            e1.record(stream)
            barrier()
            ms_h_block = shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", e0.elapsed_time(e1))[0] / max(2, args.steps // 4)
            # timed: round r + 1 of a group is submitted only after its round r has been collected; the other group's call is in
            # flight meanwhile
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            pend = {}
            for _ in range(args.steps):
                for g_, (d_, _, _) in seq_g:
                    if g_ in pend:
                        fec.wait(pend.pop(g_))                     # this group's previous round: its NACK list is what this call decodes
                    pend[g_] = fec.submit_tbs(d_, pkg.F_DEVICE_IO)
            for t_ in pend.values():
                fec.wait(t_)
            e1.record(stream)
            barrier()
            ms_h = shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", e0.elapsed_time(e1))[0]
            out_host = d_out[0].cpu().numpy()
            acked, ntx, skipped, per_round = 0, 0, 0, []
            for (g_, (d_, lv, ret0)), r in zip(seq_g, [r_ for r_ in range(4) for g2 in range(2) if r_ < len(groups[g2])]):
                if [d_[s_].ret for s_ in range(len(lv))] != ret0:
                    raise SystemExit("bench.py: the HARQ leg is not deterministic across steps")
                ok_r = [i for s_, i in enumerate(lv) if d_[s_].ret == 0]
                bad = sum(1 for i in ok_r if not (out_host[i, :TBS // 8] == payload[i % ncw]).all())
                if bad:
                    raise SystemExit(f"bench.py: {bad} ACKed transport blocks of the HARQ leg differ from the transmitted payload")
                sk = sum(sum(1 for c in range(C_CB) if d_[s_].cb_its[c] == 0) for s_ in range(len(lv))) if r else 0
                acked += len(ok_r); ntx += len(lv); skipped += sk
                per_round.append(dict(group=g_, rv=(0, 2, 3, 1)[r], tbs=len(lv), acked=len(ok_r), code_blocks_skipped_already_ok=sk))
            # the rate de-matcher in its read-modify-write case: the first retransmission round, serialised
            rmw = None
            if len(groups[0]) > 1:
                for d_, _, _ in groups[0][:2]:
                    fec.decode_tbs(d_, pkg.F_DEVICE_IO | pkg.F_SERIAL)
                ph = fec.last_phase_ms()
                d1, lv1 = groups[0][1][0], groups[0][1][1]
                ncb1 = sum(sum(1 for c in range(C_CB) if d1[s_].cb_its[c] > 0) for s_ in range(len(lv1)))
                Eh = H_NBITS // C_CB
                byt = ncb1 * (2 * Eh + 2 * 2 * (3 * 5888 + 16))       # LLRs in, slot read + slot written (3*SS+16 int16, SS = 16*368)
                rmw = dict(code_blocks=ncb1, ms=ph[0], algorithmic_bytes=byt, achieved=byt / (ph[0] / 1e3) / 1e9, unit="GB/s", peak=peaks()["hbm_gbs"],
                           frac=byt / (ph[0] / 1e3) / 1e9 / peaks()["hbm_gbs"], bytes_per_cb="2E + 2 x 2 x (3*SS+16)")
            harq = dict(workload="BASELINE configs[2] shape: TBS 75376 in 86400 LLRs (13 x K=5824, E = 6642/6648), rv 0,2,3,1, "
                                 "new_data only in round 0; later rounds combine into the soft buffers, skip code blocks that are already OK and copy their cached bytes",
                        sigma=args.sigma_harq, harq_processes=allsum(Bh), value=allsum(acked) * TBS * args.steps / (ms_h / 1e3) / 1e9, unit=UNIT + " of ACKed bits",
                        ms_per_cycle=ms_h / args.steps, blocking_ms_per_cycle=ms_h_block,
                        pipelining="two groups of processes: a group's round r + 1 is submitted after its round r was collected, the other group's call is in flight meanwhile",
                        mean_transmissions_per_tb=allsum(ntx) / allsum(Bh), acked_frac=allsum(acked) / allsum(Bh),
                        rounds_rank0=per_round, rate_dematch_read_modify_write_rank0=rmw,
                        payload_check="every ACKed transport block equals the transmitted payload; return codes identical in every cycle")
            del d_h
            fec.decode_tbs(dd, pkg.F_DEVICE_IO)

        # ---- sustained leg: the headline step looped for args.sustain seconds (the timed region of `value` is tens of milliseconds) ----
        sustained = None
        if args.sustain > 0:
            n_s = max(args.steps, int(args.sustain * 1e3 / (ms / args.steps)) + 1)
            smp = ClockSampler(local_rank)
            smp.start()
            time.sleep(0.2)
            ms_s = timed_pipelined(dd_ab, pkg.F_DEVICE_IO, n_s)
            ck = smp.stop()
            ok_s2, _ = tally(dd)
            sustained = dict(seconds=ms_s / 1e3, steps=n_s, value=allsum(ok_s2) * TBS * n_s / (ms_s / 1e3) / 1e9, unit=UNIT, ms_per_step=ms_s / n_s,
                             sm_mhz_median=ck["sm_mhz"], power_w_max=ck["power_w_max"], reasons=ck["reasons"])

        # ---- transmit side: the same transport blocks through CRC + segmentation + turbo encoding + rate matching
        #      on the device (fec_cuda_encode_tbs, device-resident payloads and packed output), checked against the oracle ----
        tx = None
        if not args.no_tx:
            ncw = payload.shape[0]
            d_pay = torch.from_numpy(np.ascontiguousarray(payload[np.arange(B) % ncw])).to(f"cuda:{local_rank}")
            EST = (NBITS // 8 + 15) & ~15
            d_e = torch.zeros((B, EST), dtype=torch.uint8, device=f"cuda:{local_rank}")
            td = (pkg.TxTbDesc * B)()
            for i in range(B):
                td[i].tbs, td[i].Qm, td[i].rv, td[i].nof_e_bits = TBS, QM, RV, NBITS
                td[i].data, td[i].e_bits = d_pay.data_ptr() + i * (TBS // 8), d_e.data_ptr() + i * EST
            for _ in range(3):
                fec.encode_tbs(td, pkg.F_DEVICE_IO)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.steps):
                fec.encode_tbs(td, pkg.F_DEVICE_IO)
            e1.record(stream)
            barrier()
            ms_tx = shard.reduce_job(dist if world > 1 else None, f"cuda:{local_rank}", e0.elapsed_time(e1))[0]
            if rank == 0:
                synth_ = importlib.import_module("empower-srslte_b200.synth")
                want = np.packbits(synth_.encode_tbs(payload[:2], TBS, QM, RV, NBITS), axis=1)
                got = d_e[:2].cpu().numpy()
                if not (got[:, : NBITS // 8] == want).all():
                    raise SystemExit("bench.py: the device encoder's bits differ from the host synthesiser's")
            tx = dict(workload="the same transport blocks encoded: CRC-24A, segmentation, CRC-24B, turbo encoder, rate matching",
                      value=world * B * TBS * args.steps / (ms_tx / 1e3) / 1e9, unit="Gbit/s of transport-block bits", ms_per_step=ms_tx / args.steps,
                      cbs_per_s=world * B * C_CB * args.steps / (ms_tx / 1e3), checked="first 2 transport blocks bit-exact vs the numpy synthesiser (itself pinned to srslte_dlsch_encode)")

        # ---- 8-bit LLR leg: the same transport blocks as int8 LLRs (amplitude 24, the scale of the reference's
        #      8-bit demodulator output after its int8 clipping) through fec_cuda_set_llr_8bit: half the ingest bytes, int8
        #      combining, the int8 decoders (P = 32 for K = 5824). The int8 decoders lose ~0.7 dB: quoted at sigma 0.3 ----
        llr8 = None
        if not args.no_llr8:
            synth8 = importlib.import_module("empower-srslte_b200.synth")
            f8 = pkg.FecCuda(local_rank, max_tbs=B)
            f8.set_llr_8bit(True)
            st8 = torch.cuda.ExternalStream(f8.stream, device=f"cuda:{local_rank}")
            h8 = pkg.PinnedArray((B, NBITS), np.int8)
            rng8 = np.random.default_rng(3000 + rank)
            e8 = synth8.encode_tbs(payload, TBS, QM, RV, NBITS)
            tmp = np.zeros((64, NBITS), np.int16)
            for i0 in range(0, B, 64):
                i1 = min(B, i0 + 64)
                synth8.awgn_llr(e8[np.arange(i0, i1) % payload.shape[0]], args.sigma8, rng8, out=tmp[: i1 - i0])
                h8.a[i0:i1] = np.clip(np.rint(tmp[: i1 - i0] * 0.24), -128, 127)
            d8 = torch.from_numpy(h8.a).to(f"cuda:{local_rank}")
            sb8 = f8.softbuffer_alloc(2 * B)
            dd8, hd8, dd8b, hd8b = (pkg.TbDesc * B)(), (pkg.TbDesc * B)(), (pkg.TbDesc * B)(), (pkg.TbDesc * B)()
            for i in range(B):
                for dsc, ep, op, sbx in ((dd8, d8.data_ptr(), d_out[1].data_ptr(), sb8), (hd8, h8.a.ctypes.data, h_out2.a.ctypes.data, sb8),
                                         (dd8b, d8.data_ptr(), d_out[0].data_ptr(), sb8 + B), (hd8b, h8.a.ctypes.data, h_out.a.ctypes.data, sb8 + B)):
                    dsc[i].tbs, dsc[i].Qm, dsc[i].rv, dsc[i].nof_e_bits, dsc[i].max_its = TBS, QM, RV, NBITS, MAX_ITS
                    dsc[i].softbuffer, dsc[i].new_data = sbx + i, 1
                    dsc[i].e_bits, dsc[i].data = ep + i * NBITS, op + i * DATA_STRIDE
            for _ in range(8):                       # every pipe of the ring sizes its scratch on first use
                f8.decode_tbs(dd8, pkg.F_DEVICE_IO)
            n8 = max(2, args.steps // 2)
            f8.decode_tbs(dd8b, pkg.F_DEVICE_IO)
            ms8 = timed_pipelined([dd8, dd8b], pkg.F_DEVICE_IO, n8, f=f8, st=st8)
            ms8s = timed(dd8, pkg.F_DEVICE_IO, n8, f=f8, st=st8)
            ok8 = sum(1 for i in range(B) if dd8[i].ret == 0)
            its8 = sum(sum(dd8[i].cb_its[:C_CB]) for i in range(B))
            o8 = d_out[1].cpu().numpy()
            bad8 = sum(1 for i in range(B) if dd8[i].ret == 0 and not (o8[i, :TBS // 8] == payload[i % payload.shape[0]]).all())
            if bad8:
                raise SystemExit(f"bench.py: {bad8} CRC-passing transport blocks of the 8-bit leg differ from the transmitted payload")
            ph8 = []
            for _ in range(3):
                f8.decode_tbs(dd8, pkg.F_DEVICE_IO | pkg.F_SERIAL)
                ph8.append(f8.last_phase_ms())
            for _ in range(6):                       # the pipes of the ring size their pinned staging on their first use
                f8.decode_tbs(hd8, 0)
                f8.decode_tbs(hd8b, 0)
            ms8e = timed_pipelined([hd8, hd8b], 0, n8, f=f8, st=st8)
            ms8es = timed(hd8, 0, n8, f=f8, st=st8)
            ok8e = sum(1 for i in range(B) if hd8[i].ret == 0)
            llr8 = dict(workload="the same transport blocks as int8 LLRs (amplitude 24, sigma %.2f), fec_cuda_set_llr_8bit" % args.sigma8,
                        value=allsum(ok8) * TBS * n8 / (ms8 / 1e3) / 1e9, unit=UNIT, ms_per_step=ms8 / n8, sync_ms_per_step=ms8s / n8, tb_ok_frac=allsum(ok8) / (world * B),
                        halfits_per_cb=allsum(its8) / (world * B * C_CB),
                        phases_ms=dict(rate_dematch=sum(p[0] for p in ph8) / 3, turbo_decode=sum(p[1] for p in ph8) / 3, tb_assemble_crc24a=sum(p[2] for p in ph8) / 3),
                        e2e=dict(value=allsum(ok8e) * TBS * n8 / (ms8e / 1e3) / 1e9, unit=UNIT, ms_per_step=ms8e / n8, sync_ms_per_step=ms8es / n8,
                                 h2d_bytes_per_step=B * NBITS + B * C_CB * (32 + 48) + B * 520,
                                 frac_of_h2d_link=(B * NBITS + B * C_CB * (32 + 48) + B * 520) / (ms8e / n8 / 1e3) / 1e9 / link_min),
                        kernel="fec_tdec_win8p_kernel<32,true>: two sub-blocks per thread in u16x2 lanes (VIADDMNMX.S16x2.RELU = saturating int8 add), beta checkpoints in tensor memory (profiles/r2_win8p_kernel_ncu.txt)")
            # ---- the 8-bit path fed with equalised symbols (8 B per resource element): int8 soft demodulation
            #      (srslte_demod_soft_demodulate_b), int8 descrambling, int8 rate de-matching, int8 decoders ----
            if front is not None:
                for i in range(6):
                    f8.decode_syms([dd8, dd8b][i & 1], sy_ab[0][i & 1], pkg.F_DEVICE_IO)
                ms8y = timed_syms([dd8, dd8b], sy_ab[0], pkg.F_DEVICE_IO, n8, f=f8, st=st8)
                ok8y = sum(1 for i in range(B) if dd8[i].ret == 0)
                its8y = sum(sum(dd8[i].cb_its[:C_CB]) for i in range(B))
                o8 = d_out[1].cpu().numpy()
                if any(dd8[i].ret == 0 and not (o8[i, :TBS // 8] == payload[i % payload.shape[0]]).all() for i in range(B)):
                    raise SystemExit("bench.py: a CRC-passing transport block of the 8-bit symbol leg differs from the transmitted payload")
                f8.decode_syms(dd8, sy_ab[0][0], pkg.F_DEVICE_IO | pkg.F_SERIAL)
                fe8_ms = f8.last_frontend_ms()
                for i in range(6):
                    f8.decode_syms([hd8, hd8b][i & 1], sy_ab[1][i & 1], 0)
                ms8yh = timed_syms([hd8, hd8b], sy_ab[1], 0, n8, f=f8, st=st8)
                ok8yh = sum(1 for i in range(B) if hd8[i].ret == 0)
                hb = B * (NBITS // QM) * 8 + B * C_CB * (32 + 48) + B * (520 + 80)
                llr8["symbols_in"] = dict(workload="same TBs as 64QAM symbols (sigma_sym %.3f), int8 front end + 8-bit path" % args.sigma_sym,
                                          value=allsum(ok8y) * TBS * n8 / (ms8y / 1e3) / 1e9, unit=UNIT, ms_per_step=ms8y / n8,
                                          tb_ok_frac=allsum(ok8y) / (world * B), halfits_per_cb=allsum(its8y) / (world * B * C_CB), frontend_ms=fe8_ms,
                                          e2e=dict(value=allsum(ok8yh) * TBS * n8 / (ms8yh / 1e3) / 1e9, unit=UNIT, ms_per_step=ms8yh / n8,
                                                   h2d_bytes_per_step=hb, frac_of_h2d_link=hb / (ms8yh / n8 / 1e3) / 1e9 / link_min))
            # ---- sigma grid: the int16 path and the int8 path on the SAME transport blocks at the same noise levels, with the
            #      half-iterations per code block beside each value (296 distinct noisy TBs per point, tiled to the batch) ----
            if not args.no_grid:
                G = min(B, 296)
                g16 = np.zeros((G, NBITS), np.int16)
                d_g16 = torch.empty((B, NBITS), dtype=torch.int16, device=f"cuda:{local_rank}")
                gsets = []
                for s_ in range(2):
                    d_ = (pkg.TbDesc * B)()
                    fill_descs(pkg, d_, B, d_g16.data_ptr(), d_out[s_].data_ptr(), sb0 + s_ * B)
                    gsets.append(d_)
                rows = np.arange(B) % G
                idx = torch.from_numpy(rows).to(f"cuda:{local_rank}")
                grid = []
                for sg in (0.30, 0.40, 0.45):
                    rng_g = np.random.default_rng(7000 + rank)
                    for i0 in range(0, G, 64):
                        i1 = min(G, i0 + 64)
                        synth8.awgn_llr(e8[np.arange(i0, i1) % payload.shape[0]], sg, rng_g, out=g16[i0:i1])
                    t16 = torch.from_numpy(g16).to(f"cuda:{local_rank}")
                    d_g16.copy_(t16[idx])
                    d8.copy_(torch.from_numpy(np.clip(np.rint(g16 * 0.24), -128, 127).astype(np.int8)).to(f"cuda:{local_rank}")[idx])
                    torch.cuda.synchronize()
                    row = dict(sigma=sg)
                    for name, ff, sets, stx in (("int16", fec, gsets, stream), ("int8", f8, [dd8, dd8b], st8)):
                        for d_ in sets:
                            ff.decode_tbs(d_, pkg.F_DEVICE_IO)
                        ms_g = timed_pipelined(sets, pkg.F_DEVICE_IO, n8, f=ff, st=stx)
                        okg = sum(1 for i in range(B) if sets[0][i].ret == 0)
                        itg = sum(sum(sets[0][i].cb_its[:C_CB]) for i in range(B))
                        og = d_out[1 if name == "int8" else 0].cpu().numpy()
                        if any(sets[0][i].ret == 0 and not (og[i, :TBS // 8] == payload[rows[i] % payload.shape[0]]).all() for i in range(B)):
                            raise SystemExit(f"bench.py: sigma grid {name} sigma {sg}: a CRC-passing transport block differs from the transmitted payload")
                        row[name] = dict(value=allsum(okg) * TBS * n8 / (ms_g / 1e3) / 1e9, unit=UNIT, ms_per_step=ms_g / n8,
                                         tb_ok_frac=allsum(okg) / (world * B), halfits_per_cb=allsum(itg) / (world * B * C_CB))
                    grid.append(row)
                llr8["sigma_grid"] = dict(note="device-resident, 2 batches in flight; value counts ACKed bits only; int8 LLRs = clip(rint(0.24 * int16 LLRs))", points=grid)
                del d_g16
            f8.close()

        # ---- latency of the reference's call shape: ONE transport block per blocking call (srslte_dlsch_decode as pdsch.c:816
        #      makes it, level 2 of INTEGRATION.md), host buffers, everything inside the timed call ----
        lat = None
        if rank == 0:
            one = (pkg.TbDesc * 1)()
            ts = []
            for i in range(64):
                fill_descs(pkg, one, 1, h_llr.a.ctypes.data + (i % B) * NBITS * 2, h_out.a.ctypes.data, sb0)
                t0 = time.perf_counter()
                fec.decode_tbs(one, 0)
                ts.append((time.perf_counter() - t0) * 1e6)
            ts = sorted(ts[8:])
            lat = dict(call="fec_cuda_decode_tbs(n=1), 13 code blocks, host LLRs in, payload out", median_us=ts[len(ts) // 2], p90_us=ts[int(len(ts) * 0.9)],
                       min_us=ts[0])

        # ---- worst case: all 8 half-iterations for every code block (no early stop), same inputs ----
        worst = None
        if not args.no_worst_case:
            fec.decode_tbs(dd, pkg.F_DEVICE_IO | pkg.F_NO_EARLY_STOP)
            ms_w = timed(dd, pkg.F_DEVICE_IO | pkg.F_NO_EARLY_STOP, max(2, args.steps // 3))
            ok_w, its_w = tally(dd)
            nst = max(2, args.steps // 3)
            worst = dict(gbps=allsum(ok_w) * TBS * nst / (ms_w / 1e3) / 1e9, cbs_per_s=world * B * C_CB * nst / (ms_w / 1e3),
                         halfits_per_cb=allsum(its_w) / (world * B * C_CB), ms_per_step=ms_w / nst)

    except Exception as ex:            # noqa: BLE001 — reported, not swallowed
        if world > 1:
            raise                      # the other ranks wait in the legs' reductions: fail the job rather than hang it
        import traceback
        aux_error = repr(ex) + " | " + traceback.format_exc(limit=3).replace("\n", " / ")

    if rank == 0:
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="int16",
                    data="synthetic",
                    config=dict(workload="pdsch_mcs28_100prb_tbs75376_13xK5824", tbs_per_gpu_per_step=B, sigma=args.sigma,
                                llr="rint(100*(+-1+sigma*n)) int16", max_half_iterations=MAX_ITS, early_stop=True, rv=RV, new_data=True,
                                sharding="independent TB batches per GPU, no collective",
                                pipelining="2 batches in flight (fec_cuda_submit_tbs / fec_cuda_wait), alternating soft-buffer sets; sync_ms_per_step = one blocking call per step",
                                l2="inputs larger than L2: %.0f MB LLRs + %.0f MB soft buffers per step" % (B * NBITS * 2 / 1e6, B * C_CB * 18464 * 2 / 1e6)),
                    cbs_per_s=cbs_per_s, tb_ok_frac=ok_all / (world * B), halfits_per_cb=its_all / (world * B * C_CB),
                    e2e=dict(value=e2e_val, unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h, ms_per_step=ms_e / args.steps,
                             sync_ms_per_step=ms_e_sync / args.steps,
                             # the ingest is link-bound by construction: the same ranks copying pinned 256 MB buffers at the same time
                             h2d_link=dict(gbs_slowest_rank=link_min, gbs_all_ranks=link_sum, how="pinned 256 MB x 8 cudaMemcpyAsync per rank, all ranks at once (tools/h2d_peak.cu standalone)",
                                           achieved_gbs_per_rank=h2d / (ms_e / args.steps / 1e3) / 1e9, frac=h2d / (ms_e / args.steps / 1e3) / 1e9 / link_min),
                             cpu_binding=affinity),
                    sync_ms_per_step=ms_sync / args.steps, host_submit_ms_per_step=host_submit_ms,
                    gpu_launches=launches, clocks=clocks, roofline=roof, worst_case_8_halfits=worst, frontend_symbols=front, ulsch_config5=ulsch5, harq_config3=harq, sustained=sustained, transmit_side=tx, llr_8bit=llr8, single_tb_latency=lat, cpu_baseline=cpu, parity_in_bench=parity, aux_error=aux_error)
        print(json.dumps(line), flush=True)
    fec.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
