#!/usr/bin/env python
"""The role-split integrator at small binnings (4 bulk warps, three CTAs per SM for N_E <= 128) against the diagonal-owned kernel
and the 7-bulk-warp configuration, 8192 points."""
import sys
import numpy as np
sys.path.insert(0, "/root/repo")
import nusiprop_b200 as nb
for N in (40, 60, 100, 128, 129, 160):
    cfg = dict(N_bins_E=N, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    npts = 8192
    mphi, g = np.repeat(10 ** np.linspace(1.5, 3.5, npts // 32), 32), np.tile(10 ** np.linspace(-4, -1, 32), npts // 32)
    t, out = {}, {}
    for integ, var in ((3, 0), (0, 0), (0, 11)):
        ctx = nb.Context(integrator=integ, fine_variant=var, **cfg)
        for _ in range(2):
            out[(integ, var)] = ctx.evolve(mphi, g, 0.1)[0]
            t[(integ, var)] = ctx.last_timing()["ms_integrate"]
        ctx.close()
    ok = np.isfinite(out[(3, 0)]) & (out[(3, 0)] != 0)
    print("N=%d: 8192 pts diag %.2f split(default) %.2f split<4,7,2> %.2f (rel %.1e)" % (N, t[(3, 0)], t[(0, 0)], t[(0, 11)], np.max(np.abs(out[(0, 0)][ok] - out[(3, 0)][ok]) / np.abs(out[(3, 0)][ok]))), flush=True)
