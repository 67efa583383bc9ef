#!/usr/bin/env python
"""Per-instruction stall samples of an ncu source page (ncu -i X.ncu-rep --page source --csv > f.csv):
   python tools/ncu_src.py f.csv [stall column, default all] [top N]
prints the totals per stall reason and the top-N instructions for the chosen reason with their SASS text."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr]
body = [r for r in rows[hdr + 1:] if len(r) == len(h)]
col = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]


def num(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


tot = {s: sum(num(r[col[s]]) for r in body) for s in stalls}
allsmp = sum(num(r[col["# Samples"]]) for r in body)
print("samples", allsmp, " instructions", len(body))
for s, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print("  %-24s %8.0f  %5.1f %%" % (s, v, 100 * v / allsmp))
which = sys.argv[2] if len(sys.argv) > 2 else None
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
if which:
    key = col[which] if which in col else col["# Samples"]
    for r in sorted(body, key=lambda r: -num(r[key]))[:topn]:
        print("%8s %7.0f  exec %8.0f  %s" % (r[col["Address"]][-5:], num(r[key]), num(r[col["Instructions Executed"]]), r[col["Source"]][:110]))
