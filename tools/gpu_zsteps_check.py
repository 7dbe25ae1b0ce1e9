#!/usr/bin/env python
"""Default integrators (role-split, one CTA or a cluster per point) against the diagonal-owned kernel for bands of every length:
few z-steps (one short band), many, and binnings on either side of the kernels' switch points."""
import sys
import numpy as np
sys.path.insert(0, "/root/repo")
import nusiprop_b200 as nb
rng = np.random.default_rng(4)
mphi, g = 10 ** rng.uniform(1, 4, 5), 10 ** rng.uniform(-4, 0, 5)
for N, lo, hi in ((700, 4.0, 9.0), (1200, 4.0, 9.0), (300, 12.0, 17.0), (450, 4.0, 6.0), (2000, 4.0, 9.0)):
    for zmax in (0.004, 0.03, 0.3, 1.5):
        cfg = dict(N_bins_E=N, lEmin=lo, lEmax=hi, zmax=zmax, flav=1)
        try:
            a = nb.Context(**cfg); b = nb.Context(integrator=3, **cfg)
        except Exception as ex:
            print("N=%d zmax=%g: %s" % (N, zmax, str(ex)[:80])); continue
        ra, rb = a.evolve(mphi, g, 0.1), b.evolve(mphi, g, 0.1)
        S = a.N_steps_z - 1
        a.close(); b.close()
        ok = np.isfinite(rb[0]) & (rb[0] != 0)
        rel = np.max(np.abs(ra[0][ok] - rb[0][ok]) / np.abs(rb[0][ok])) if ok.any() else 0.0
        print("N=%d zmax=%g S=%d: default vs diag max rel %.2e, same zero/finite masks %s" % (N, zmax, S, rel, np.array_equal(ra[0] == 0, rb[0] == 0) and np.array_equal(np.isfinite(ra[0]), np.isfinite(rb[0]))), flush=True)
