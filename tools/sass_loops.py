#!/usr/bin/env python
"""Static schedule of the loops of one kernel in a cubin (no GPU needed): per loop (backward branch) the instruction count, the
sum of the stall counts ptxas encoded (= issue cycles of ONE warp if no scoreboard wait triggers) and the opcode mix.
   python tools/sass_loops.py file.cubin 'fec_tdec_pk_kernelILi16ELb1ELb0' [min_len]"""
import collections
import re
import subprocess
import sys


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
                ins.append(dict(addr=int(m.group(1), 16), s=m.group(2).strip(), stall=c & 15, yld=(c >> 4) & 1, wb=(c >> 5) & 7, rb=(c >> 8) & 7, wait=(c >> 11) & 63))
                i += 2
                continue
        i += 1
    return ins


def main():
    ins = decode(sys.argv[1], sys.argv[2])
    minlen = int(sys.argv[3]) if len(sys.argv) > 3 else 100
    byaddr = {x["addr"]: k for k, x in enumerate(ins)}
    for k, x in enumerate(ins):
        m = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", x["s"])
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt >= x["addr"] or tgt not in byaddr:
            continue
        body = ins[byaddr[tgt]:k + 1]
        if len(body) < minlen or len(body) > 1500:
            continue
        ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", b["s"]).split()[0].split(".")[0] + ("." + re.sub(r"^@!?U?P\d+\s+", "", b["s"]).split()[0].split(".")[1] if re.sub(r"^@!?U?P\d+\s+", "", b["s"]).split()[0].startswith(("VI", "IMAD")) and "." in b["s"].split()[0] else "") for b in body)
        st = collections.Counter(b["stall"] for b in body)
        waits = sum(1 for b in body if b["wait"])
        print("loop %5x..%5x  instrs %4d  sum(stall) %5d  (%.2f per instr)  instrs with a scoreboard wait: %d" % (tgt, x["addr"], len(body), sum(b["stall"] for b in body), sum(b["stall"] for b in body) / len(body), waits))
        print("    stall histogram:", dict(sorted(st.items())))
        print("    ops:", ", ".join("%s:%d" % kv for kv in ops.most_common(14)))


if __name__ == "__main__":
    main()
