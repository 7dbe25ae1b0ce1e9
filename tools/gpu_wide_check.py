#!/usr/bin/env python
"""The one-CTA role-split integrator (default, N_E <= 576) against the diagonal-owned kernel (3) and the cluster kernel (4) in the
binnings 224 < N_E <= 576, over batch sizes:  python tools/gpu_wide_check.py [N_E ...]"""
import sys, os
import numpy as np
sys.path.insert(0, "/root/repo")
import nusiprop_b200 as nb
for N in [int(a) for a in sys.argv[1:]] or (240, 300, 384):
    cfg = dict(N_bins_E=N, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    line = "N=%d:" % N
    for npts in (8, 148, 592, 4096):
        mphi, g = np.repeat(10 ** np.linspace(1.5, 3.5, (npts + 31) // 32), 32)[:npts], np.tile(10 ** np.linspace(-4, -1, 32), (npts + 31) // 32)[:npts]
        t, out = {}, {}
        for integ in (3, 0, 4):
            ctx = nb.Context(integrator=integ, **cfg)
            for _ in range(2):
                out[integ] = ctx.evolve(mphi, g, 0.1)[0]
                t[integ] = ctx.last_timing()["ms_integrate"]
            ctx.close()
        ok = np.isfinite(out[3]) & (out[3] != 0)
        line += "  %d pts diag %.2f split-wide %.2f cluster %.2f (rel %.1e)" % (npts, t[3], t[0], t[4], np.max(np.abs(out[0][ok] - out[3][ok]) / np.abs(out[3][ok])))
    print(line, flush=True)
