#!/usr/bin/env python
"""Aggregate an ncu source page by CUDA source line.
    ncu -i rep --page source --csv --print-source cuda,sass [-k regex:name] > src.csv ; python tools/ncu_source_hot.py src.csv [topn]
Per line: share of warp-stall samples, warp instructions executed, and the dominant stall reasons."""
import csv, sys, collections
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
agg = collections.defaultdict(lambda: collections.defaultdict(float)); src = {}
hdr = None
for r in rows:
    if r and r[0] == "Line No" and len(r) > 5:
        hdr = r; ci = {h: i for i, h in enumerate(hdr)}; stalls = [h for h in hdr if h.startswith("stall_")]
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    a = agg[ln]
    for key, col in (("smp", "# Samples"), ("inst", "Instructions Executed")):
        try: a[key] += float(r[ci[col]] or 0)
        except ValueError: pass
    for s in stalls:
        try: a[s] += float(r[ci[s]] or 0)
        except ValueError: pass
    a["nsass"] += 1
    src.setdefault(ln, r[1][:100])
ts = sum(a["smp"] for a in agg.values()); ti = sum(a["inst"] for a in agg.values())
print("total samples %d, warp instructions %d" % (ts, ti))
tot_st = collections.defaultdict(float)
for a in agg.values():
    for k, v in a.items():
        if k.startswith("stall_"): tot_st[k] += v
print("stall mix: " + ", ".join("%s %.1f%%" % (k[6:], 100 * v / max(ts, 1)) for k, v in sorted(tot_st.items(), key=lambda kv: -kv[1])[:8]))
for ln, a in sorted(agg.items(), key=lambda kv: -kv[1]["smp"])[:topn]:
    st = sorted(((k[6:], v) for k, v in a.items() if k.startswith("stall_") and v > 0), key=lambda kv: -kv[1])[:3]
    print("%5.1f%% smp %5.1f%% inst n=%3d L%-4d %-70s | %s" % (100 * a["smp"] / max(ts, 1), 100 * a["inst"] / max(ti, 1), a["nsass"], ln, src[ln][:70],
                                                          " ".join("%s:%.0f%%" % (k, 100 * v / max(a["smp"], 1)) for k, v in st)))
