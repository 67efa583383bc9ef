import json, sys
j = json.load(open(sys.argv[1] if len(sys.argv) > 1 else "/root/repo/gpurun_out/bench.json"))
r = j["roofline"]
print("value %.2f Gbps  step %.3f ms | serial phases rm %.3f dec %.3f asm %.3f | frac %.3f | e2e %.2f Gbps (%.2f ms) | worst %s" % (
    j["value"], j["ms_per_step"], r["phases_ms"]["rate_dematch"], r["phases_ms"]["turbo_decode"], r["phases_ms"]["tb_assemble_crc24a"],
    r["frac"], j["e2e"]["value"], j["e2e"]["ms_per_step"],
    ("%.2f Gbps %.2f ms" % (j["worst_case_8_halfits"]["gbps"], j["worst_case_8_halfits"]["ms_per_step"])) if j.get("worst_case_8_halfits") else "-"))
