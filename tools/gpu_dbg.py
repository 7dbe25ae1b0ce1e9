import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import nusiprop_b200 as nb, oracle_ref as R
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
ctx = nb.Context(**cfg)
mp = 10 ** np.linspace(1, 4, 16)
for g in (1e-12, 1e-9, 1e-7):
    fm, ff, st = ctx.evolve(mp, g, 0.1)
    d = R.relerr(fm, fm[0:1]).reshape(16, -1).max(axis=1)
    print("g=%g status %s maxrel vs first per point:" % (g, st.tolist()), np.array2string(d, precision=1))
    rm, rf, _ = R.evolve(mp[5], g, 0.1, **cfg)
    print("   vs oracle point 5: %.2e ; oracle p5 vs oracle p0: %.2e" % (R.relerr(fm[5], rm).max(), R.relerr(rm, R.evolve(mp[0], g, 0.1, **cfg)[0]).max()))
