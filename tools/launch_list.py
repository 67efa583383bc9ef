"""aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel: python tools/launch_list.py launches.csv"""
import csv, sys, collections
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, mi, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[mi] != "gpu__time_duration.sum":
        continue
    v = float(r[vi].replace(",", ""))
    us = v / 1e3 if r[ui] in ("ns", "nsecond") else v * 1e3 if r[ui] in ("ms", "msecond") else v
    name = r[ki].split("(")[0][:60]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print("%-60s %8s %12s %10s %7s" % ("kernel", "launches", "total_us", "mean_us", "share"))
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-60s %8d %12.1f %10.1f %7.3f" % (k, a[0], a[1], a[1] / a[0], a[1] / tot))
