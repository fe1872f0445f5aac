#!/usr/bin/env python
"""Registers / spills per conv_umma_kernel instantiation from build/obj/ptxas.log: python profiles/ptxas_summary.py [EF ...]"""
import re, sys
log = open(sys.argv[1] if len(sys.argv) > 1 and sys.argv[1].endswith(".log") else "build/obj/ptxas.log").read()
want = set(a for a in sys.argv[1:] if not a.endswith(".log"))
cur = None
rows = {}
for line in log.splitlines():
    m = re.search(r"Compiling entry function '.*conv_umma_kernelI([^E]*(?:E[^vN])*)E*vNS", line)
    m = re.search(r"Compiling entry function '.*?conv_umma_kernelI(.*?)EEvNS", line)
    if m:
        cur = tuple(x.replace("Li", "").replace("Lb", "").replace("n", "-") for x in m.group(1).split("E") if x)
        continue
    if "Compiling entry function" in line:
        cur = None
    if cur:
        m = re.search(r"(\d+) bytes spill stores", line)
        if m: rows.setdefault(cur, {})["spill"] = int(m.group(1))
        m = re.search(r"Used (\d+) registers", line)
        if m: rows.setdefault(cur, {})["regs"] = int(m.group(1)); cur = None
for k in sorted(rows, key=lambda k: (int(k[-1]), k)):
    if want and k[-1] not in want: continue
    print("<%s> regs %s spill %s" % (",".join(k), rows[k].get("regs"), rows[k].get("spill")))
