import json, sys
j = json.loads([l for l in open(sys.argv[1] if len(sys.argv) > 1 else "/root/repo/gpurun_out/bench.json").read().splitlines() if l.startswith("{")][-1])
r = j["roofline"]
print("value %.2f Gbps  step %.3f ms | serial phases rm %.3f dec %.3f asm %.3f | frac %.3f | e2e %.2f Gbps (%.2f ms) | worst %s" % (
    j["value"], j["ms_per_step"], r["phases_ms"]["rate_dematch"], r["phases_ms"]["turbo_decode"], r["phases_ms"]["tb_assemble_crc24a"],
    r["frac"], j["e2e"]["value"], j["e2e"]["ms_per_step"],
    ("%.2f Gbps %.2f ms" % (j["worst_case_8_halfits"]["gbps"], j["worst_case_8_halfits"]["ms_per_step"])) if j.get("worst_case_8_halfits") else "-"))
f = j.get("frontend_symbols")
if f:
    k = f["frontend_kernel"]
    print("symbols: value %.2f Gbps step %.3f ms ok %.2f its %.2f | e2e %.2f Gbps (%.2f ms) | fe kernel %.3f ms %.0f GB/s (%.2f of HBM)" % (
        f["value"], f["ms_per_step"], f["tb_ok_frac"], f["halfits_per_cb"], f["e2e"]["value"], f["e2e"]["ms_per_step"],
        k["ms_per_launch"], k["achieved"], k["frac"]))
    if "serial_phases_ms" in f:
        print("symbols serial phases:", {k: round(v, 3) for k, v in f["serial_phases_ms"].items()})
u = (f or {}).get("uplink")
if u:
    print("uplink: value %.2f Gbps step %.3f ms ok %.2f its %.2f | phases %s | kernels %s" % (u["value"], u["ms_per_step"], u["tb_ok_frac"], u["halfits_per_cb"],
          {k: round(v, 3) for k, v in u["serial_phases_ms"].items()}, u.get("code_blocks_by_kernel")))
t = j.get("transmit_side")
if t:
    print("transmit side: %.1f Gbit/s, %.3f ms per step, %.2f M code blocks/s" % (t["value"], t["ms_per_step"], t["cbs_per_s"] / 1e6))
l = j.get("single_tb_latency")
if l:
    print("single-TB call latency: median %.0f us, p90 %.0f us, min %.0f us" % (l["median_us"], l["p90_us"], l["min_us"]))
q = j.get("llr_8bit")
if q:
    print("8-bit LLR leg: value %.2f Gbps step %.3f ms ok %.3f its %.2f | e2e %.2f Gbps (%.2f ms) | phases %s" % (
        q["value"], q["ms_per_step"], q["tb_ok_frac"], q["halfits_per_cb"], q["e2e"]["value"], q["e2e"]["ms_per_step"],
        {k: round(v, 3) for k, v in q["phases_ms"].items()}))
