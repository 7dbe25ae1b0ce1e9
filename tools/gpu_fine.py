"""BASELINE configs[3]: N_E = 2000 (NT = 2311, 312 z-steps).  Times a small batch and reports us per z-step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nusiprop_b200 as nb
ctx = nb.Context(N_bins_E=2000, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
mphi = 10 ** np.linspace(1.5, 3.0, n); g = 10 ** np.linspace(-3, -1, n)
for rep in range(2):
    t0 = time.time(); fm, ff, st = ctx.evolve(mphi, g, 0.1); dt = time.time() - t0
    tm = ctx.last_timing()
    print("N_E=2000 n=%d wall %.3f s (%.1f ms/point) %s  us/z-step/point = %.2f  finite=%s" % (n, dt, 1e3 * dt / n, tm, 1e3 * tm["ms_integrate"] / n / (ctx.N_steps_z - 1), bool(np.isfinite(fm).all())))
