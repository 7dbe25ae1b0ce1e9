#!/usr/bin/env python
"""Row kernels against opts.rows_split (CTAs that share one scan-row group: more CTAs = more parallelism, but each repeats the
g-independent part of its rows): 4 is the optimum at the benchmarked launch shape (B200, 8192-point shard)."""
import sys
import numpy as np
sys.path.insert(0, "/root/repo")
import nusiprop_b200 as nb, bench
m, g = bench.shard(0)
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
ref = None
for sp in (0, 8, 12, 16, 24, 32):
    ctx = nb.Context(rows_split=sp, **cfg)
    for _ in range(3):
        out = ctx.evolve(m, g, 0.1)
        tm = ctx.last_timing()
    ctx.close()
    ref = out if ref is None else ref
    print("rows_split %2d: rows %.3f ms (alpha %.3f, integrate %.3f); same bits %s" % (sp, tm["ms_rows"], tm["ms_alpha"], tm["ms_integrate"], np.array_equal(out[0], ref[0])), flush=True)
