#!/usr/bin/env python
"""Instruction mix and stall samples by SASS opcode from an ncu source page:
   ncu -i X.ncu-rep --page source --csv --print-source sass > x.csv ; python tools/ncu_sass_mix.py x.csv"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; c = {h: i for i, h in enumerate(hdr)}
ops = collections.defaultdict(lambda: [0, 0, collections.Counter()])
tot = [0, 0]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
for r in rows[hi + 1:]:
    if len(r) < len(hdr): continue
    src = r[c["Source"]].strip()
    tok = src.split()
    if not tok: continue
    op = tok[1] if tok[0].startswith("@") and len(tok) > 1 else tok[0]
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LD", "ST")) and "." in op else "")
    ex = int(float(r[c["Instructions Executed"]] or 0)); sm = int(float(r[c["# Samples"]] or 0))
    ops[op][0] += ex; ops[op][1] += sm
    for s in stalls:
        v = int(float(r[c[s]] or 0))
        if v: ops[op][2][s] += v
    tot[0] += ex; tot[1] += sm
print("total warp-inst %d  samples %d" % tuple(tot))
for op, (ex, sm, st) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    top = ", ".join("%s %.0f%%" % (k[6:], 100.0 * v / max(sm, 1)) for k, v in st.most_common(3))
    print("%-10s inst %5.1f%%  samples %5.1f%%   %s" % (op, 100.0 * ex / tot[0], 100.0 * sm / max(tot[1], 1), top))
