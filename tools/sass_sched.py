#!/usr/bin/env python
"""Where did ptxas put the global loads, and who waits on their scoreboard?  (no GPU needed)
   python tools/sass_sched.py [-D...]   compiles csrc/fec_tdec_packed.cu to a cubin and, for fec_tdec_pk_kernel<16,true>, prints
   every LDG / LDTM / STTM / branch plus every instruction whose wait mask names a scoreboard that an LDG set."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "empower-srslte_b200", "csrc", "fec_tdec_packed.cu")


def decode(path, fun):
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout.split("\n")
    ins, on, i = [], False, 0
    while i < len(txt):
        if "Function :" in txt[i]:
            on = fun in txt[i]
        m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/", txt[i])
        if on and m and i + 1 < len(txt):
            m2 = re.match(r"\s+/\* (0x[0-9a-f]+) \*/", txt[i + 1])
            if m2:
                c = (int(m2.group(1), 16) >> 41) & 0x7FFFFF
                ins.append(dict(addr=int(m.group(1), 16), s=m.group(2).strip(), stall=c & 15, wb=(c >> 5) & 7, rb=(c >> 8) & 7, wait=(c >> 11) & 63))
                i += 2
                continue
        i += 1
    return ins


def main():
    flags = [a for a in sys.argv[1:] if a.startswith("-D")]
    cubin = "/tmp/sx/pk.cubin"
    os.makedirs("/tmp/sx", exist_ok=True)
    subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-cubin", "-o", cubin, SRC] + flags, check=True)
    ins = decode(cubin, "fec_tdec_pk_kernelILi16ELb1EE")
    ldg_sb = set(x["wb"] for x in ins if x["s"].find("LDG") >= 0 and x["wb"] != 7)
    print("instructions", len(ins), "LDG scoreboards", sorted(ldg_sb))
    last = {}
    print("-- waits that follow an LDG on the same scoreboard by < 150 instructions (exposed latency):")
    for k, x in enumerate(ins):
        for b in range(6):
            if (x["wait"] >> b) & 1 and b in last and k - last[b] < 150:
                print("   %5d %-60s waits SB%d set %d instructions earlier by %s" % (k, x["s"][:60], b, k - last[b], ins[last[b]]["s"][:50]))
        if "LDG" in x["s"] and x["wb"] != 7:
            last[x["wb"]] = k
        elif x["wb"] != 7 and x["wb"] in last:
            del last[x["wb"]]
    if "-v" not in sys.argv:
        return
    for k, x in enumerate(ins):
        s = x["s"]
        w = [b for b in range(6) if (x["wait"] >> b) & 1 and b in ldg_sb]
        if re.search(r"LDG|LDTM|STTM|BRA|WARPSYNC|BAR", s) or w:
            print("%5d %5x wb%s wait%s %s" % (k, x["addr"], x["wb"] if x["wb"] != 7 else "-", format(x["wait"], "06b"), s[:100]))


if __name__ == "__main__":
    main()
