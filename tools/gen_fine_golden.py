#!/usr/bin/env python
"""Golden fixture for BASELINE configs[3] at FULL size: N_E = 2000, lE 4 -> 9 (N_z = 313, NT = 2311, 2.67 M alpha() calls),
(mphi, g) = (100 eV, 1e-2), mntot = 0.1 -- SURVEY 8(d) "S-fine" -- from the ORACLE (unmodified reference headers + refshim,
~35 s per evolve() on one core) and three members of its noise-floor ensemble.

    make -C oracle && python tools/gen_fine_golden.py        ->  tests/golden/oracle_fine2000.npz
      mass [3][2000], fla [3][2000], floor [2000] (largest relative deviation of a member per bin, no widening), members
"""
import os, sys, time
from multiprocessing import Pool
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
CFG = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=2000, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
POINT = (100.0, 1e-2, 0.1)
MEMBERS = ["", "_fast", "_ulp", "_ulp2"]


def _one(v):
    import oracle_ref as R
    fm, ff, en = R.evolve(*POINT, variant=v, **CFG)
    return fm, ff


if __name__ == "__main__":
    import oracle_ref as R
    R.build_oracle()
    t0 = time.time()
    with Pool(len(MEMBERS)) as pool:
        res = pool.map(_one, MEMBERS)
    fm, ff = res[0]
    floor = np.zeros(2000)
    for m, f in res[1:]:
        floor = np.maximum(floor, np.maximum(R.relerr(m, fm).max(axis=0), R.relerr(f, ff).max(axis=0)))
    out = os.path.join(ROOT, "tests", "golden", "oracle_fine2000.npz")
    np.savez_compressed(out, mass=fm, fla=ff, floor=floor, members=np.array(MEMBERS[1:]), point=np.array(POINT))
    print("wrote %s in %.0f s; bins within 1e-9 for every member: %.4f; largest member deviation %.2e" % (out, time.time() - t0, np.mean(floor <= 1e-9), floor.max()))
