#!/usr/bin/env python
"""Hot source lines of one kernel from an ncu report:
   ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > x.csv ; python tools/ncu_source_lines.py x.csv [N]
Prints, per source line, the share of executed warp instructions, of FP64 instructions and of stall samples."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1], errors="replace")))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fname = "?"; hdr = None; cur = None
agg = collections.OrderedDict()
for r in rows:
    if not r: continue
    if r[0] == "File Name": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = {h: i for i, h in enumerate(r)}; iS = [i for i, h in enumerate(r) if h == "Source"]; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0]:
        cur = (fname, int(r[0])); agg.setdefault(cur, [0, 0, 0, r[iS[0]].strip()])
        continue
    if cur is None: continue
    ins = r[iS[1]].strip().split()
    op = ins[1] if ins and ins[0].startswith("@") and len(ins) > 1 else (ins[0] if ins else "")
    num = lambda v: int(float(v)) if v not in ("", "-") else 0
    ex = num(r[hdr["Instructions Executed"]]); sm = num(r[hdr["# Samples"]])
    a = agg[cur]; a[0] += ex; a[1] += sm
    if op.startswith(("DFMA", "DMUL", "DADD", "DSETP")): a[2] += ex
te = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values()); tf = sum(a[2] for a in agg.values())
print("total warp-inst %d  fp64 %d (%.1f%%)  samples %d" % (te, tf, 100.0 * tf / te, ts))
byfile = collections.Counter()
for (f, l), a in agg.items(): byfile[f] += a[1]
print("samples by file:", ", ".join("%s %.1f%%" % (f, 100.0 * v / ts) for f, v in byfile.most_common()))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-16s %5d  inst %5.2f%%  fp64 %5.2f%%  samples %5.2f%%  | %s" % (f, l, 100.0 * a[0] / te, 100.0 * a[2] / max(tf, 1), 100.0 * a[1] / ts, a[3][:90]))
