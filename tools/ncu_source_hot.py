#!/usr/bin/env python
"""Aggregate an ncu source page (--page source --csv --print-source cuda,sass) by source file:line."""
import csv, sys, collections, re
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
# In cuda,sass mode each SASS row carries the CUDA source line in col "Line No"/"Source"? inspect
iexec = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples")
agg = collections.defaultdict(lambda: [0, 0, 0]); src = {}
cur_file = None
tot = [0, 0]
for r in rows:
    if r and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if len(r) <= iexec or r[0] in ("Line No", "Function Name"): continue
    try:
        ln = r[0]; ex = int(float(r[iexec] or 0)); sm = int(float(r[isamp] or 0))
    except ValueError:
        continue
    key = (cur_file, ln)
    agg[key][0] += ex; agg[key][1] += sm; agg[key][2] += 1
    src.setdefault(key, r[1][:110])
    tot[0] += ex; tot[1] += sm
print("total warp-inst %d, samples %d" % tuple(tot))
byfile = collections.defaultdict(lambda: [0, 0])
for (f, ln), v in agg.items():
    byfile[f][0] += v[0]; byfile[f][1] += v[1]
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("%-28s inst %5.1f%%  samples %5.1f%%" % (f, 100.0 * v[0] / max(tot[0], 1), 100.0 * v[1] / max(tot[1], 1)))
print()
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print("%5.1f%% inst %5.1f%% smp  n_sass=%4d  %s:%s  %s" % (100.0 * v[0] / max(tot[0], 1), 100.0 * v[1] / max(tot[1], 1), v[2], f, ln, src[(f, ln)]))
