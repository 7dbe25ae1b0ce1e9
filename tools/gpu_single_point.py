#!/usr/bin/env python
"""Latency of the drop-in use: one calculate_flux / pyprop object, evolve() after evolve() (a GPU batch of one point each)."""
import sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import nusiprop_b200 as nb
for N, kw in ((300, dict(lEmin=12.0, lEmax=17.0)), (200, dict(lEmin=4.0, lEmax=9.0)), (2000, dict(lEmin=4.0, lEmax=9.0))):
    p = nb.pyprop(100.0, 1e-2, 0.1, 2.0, phiphi=False, N_bins_E=N, **kw)
    p.evolve()
    t0 = time.perf_counter()
    n = 30
    for i in range(n):
        p.g = 1e-2 * (1 + 1e-3 * i)
        p.evolve()
    dt = (time.perf_counter() - t0) / n
    ctx = p._ctx
    tm = ctx.last_timing()
    print("N_E=%d: %.3f ms per evolve() of one object (kernels of the last %d calls: rows %.3f alpha %.3f integrate %.3f ms each)" % (
        N, 1e3 * dt, n, tm["ms_rows"] / n, tm["ms_alpha"] / n, tm["ms_integrate"] / n), flush=True)
