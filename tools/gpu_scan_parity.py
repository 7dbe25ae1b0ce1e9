#!/usr/bin/env python
"""GPU side of the scan-parity study: evolve the 256 x 256 scan shard by shard in scan order with the default
builder (so that the grouped scan kernels produce the numbers), pick the seeded 1024-point subsample, and
store its mass-basis fluxes per builder mode in gpurun_out/scan_parity_cuda.npz for analysis against
tests/golden/oracle_scan1024.npz (tools/scan_parity_report.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nusiprop_b200 as nb  # noqa: E402
import bench  # noqa: E402

idx = np.random.default_rng(20261104).choice(65536, 1024, replace=False)
mphi, g = bench.scan_points()
out = {}
for mode in (3, 2, 0):
    ctx = nb.Context(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2,
                     elementary_mode=mode)
    res = np.zeros((1024, 3, 200))
    stat = np.zeros(1024, dtype=np.int32)
    if mode == 3:
        for r in range(8):
            m, gg = bench.shard(r)
            fm, ff, st = ctx.evolve(m, gg, 0.1)
            sel = np.nonzero(idx % 8 == r)[0]
            res[sel] = fm[idx[sel] // 8]
            stat[sel] = st[idx[sel] // 8]
    else:
        fm, ff, st = ctx.evolve(mphi[idx], g[idx], 0.1)
        res, stat = fm, st
    out["mass_mode%d" % mode] = res
    out["status_mode%d" % mode] = stat
    ctx.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", "scan_parity_cuda.npz"), index=idx, **out)
print("saved; finite:", all(np.isfinite(out["mass_mode%d" % m]).all() for m in (0, 2, 3)))
