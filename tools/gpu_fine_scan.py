import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nusiprop_b200 as nb
ctx = nb.Context(N_bins_E=2000, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
# 4 mediator masses x 8 couplings (grouped) at N_E = 2000
M, G = np.meshgrid(10 ** np.linspace(1.5, 3.0, 4), 10 ** np.linspace(-3, -1, 8), indexing="ij")
for rep in range(2):
    t0 = time.time(); fm, ff, st = ctx.evolve(M.ravel(), G.ravel(), 0.1); dt = time.time() - t0
    print("N_E=2000 scan 4x8 wall %.3f s %s finite=%s" % (dt, ctx.last_timing(), bool(np.isfinite(fm).all())))
