#!/usr/bin/env python
"""SASS evidence of the Blackwell path: every line of a kernel that touches an mbarrier (SYNCS.*), a TMA bulk copy
(UBLKCP = cp.async.bulk ... mbarrier::complete_tx::bytes), a CTA / cluster barrier (BAR.*, UCGABAR*), distributed shared
memory (ST with a cluster address: STS/ST.E via MAPA) -- +-1 line of context, and instruction counts.
    python tools/sass_excerpt.py <kernel-name-substring> <out.txt>"""
import re
import subprocess
import sys

ROOT = __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__)))
name, out = sys.argv[1], sys.argv[2]
lib = ROOT + "/nusiprop_b200/lib/libnusiprop_b200.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.splitlines()
# cut out the first function whose mangled name contains every part of `name` (parts separated by '+')
start = next(i for i, l in enumerate(sass) if "Function :" in l and all(p in l for p in name.split("+")))
end = next((i for i in range(start + 1, len(sass)) if "Function :" in sass[i]), len(sass))
L = sass[start:end]
pat = re.compile(r"UBLKCP|SYNCS|BAR\.|UCGABAR|MAPA|CCTL|FENCE|MEMBAR|ERRBAR")
keep = set()
for i, l in enumerate(L):
    if pat.search(l):
        keep.update((i - 1, i, i + 1))
ops = ("UBLKCP", "SYNCS.EXCH", "SYNCS.ARRIVE", "SYNCS.PHASECHK", "BAR.SYNC", "UCGABAR_ARV", "UCGABAR_WAIT", "MAPA", "DFMA", "DMUL", "DADD", "LDS.128", "LDS.64", "SHFL", "STS", "LDL", "STL")
res = [L[0].strip(), "cuobjdump -sass nusiprop_b200/lib/libnusiprop_b200.so (sm_100a); counts of static instructions in this kernel:",
       "  " + ", ".join("%s %d" % (o, sum(o in l for l in L)) for o in ops), ""]
prev = None
for k in sorted(keep):
    if 0 <= k < len(L):
        if prev is not None and k != prev + 1:
            res.append("    ...")
        res.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", L[k]).rstrip())
        prev = k
open(out, "w").write("\n".join(res) + "\n")
print("\n".join(res[:4]))
