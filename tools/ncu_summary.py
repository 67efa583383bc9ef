"""print the metrics of an ncu report that the profiles/ summaries quote: python tools/ncu_summary.py file.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, vals = rows[0], rows[-1]
d = dict(zip(hdr, vals))
print("kernel:", d.get("Kernel Name"))
keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct", "sm__inst_executed_pipe_alu",
        "sm__inst_executed_pipe_fma", "sm__inst_executed_pipe_lsu", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct",
        "launch__shared_mem_per_block", "smsp__warps_eligible.avg.per_cycle_active", "lts__t_sector_hit_rate"]
for k in hdr:
    if any(k.startswith(x) for x in keys):
        print(k, d[k])
    elif "smsp__average_warp" in k and "issue_stalled" in k and "not_issued" not in k and k.endswith(".ratio"):
        try:
            if float(d[k].replace(",", "")) >= 0.1:
                print(k, d[k])
        except ValueError:
            pass
