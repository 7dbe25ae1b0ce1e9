#!/usr/bin/env python
"""Golden fixture for the HEADLINE workload (BASELINE configs[4]): the seeded 1024-point subsample of the
256 x 256 (mphi, g) scan (default_rng(20261104).choice(65536, 1024, replace=False)), evolved by the
ORACLE (unmodified reference headers + refshim) and by the members of its noise-floor ensemble:

    fast      -O2 -mfma -ffp-contract=fast rebuild of the same sources
    ulp, ulp2, ulp3, ulp4   +-1 ulp on a (seeded) quarter of the reference's libm results
    fastulp   both together

Run where /root/reference exists:   make -C oracle && python tools/gen_scan_golden.py [--points 1024] [--jobs 8]

tests/golden/oracle_scan1024.npz:
    index       [P]           indices into the row-major scan (mphi index * 256 + g index)
    mphi, g     [P]
    mass        [P][3][200]   flux of the mass eigenstates from the strict oracle (float64)
    floor       [P][200]      per energy bin: the largest relative deviation of any ensemble member from the strict
                              oracle, over the three mass eigenstates (float32; NO widening over neighbouring bins)
    members     names of the ensemble members
    member_strict_frac [M]    per member: fraction of (point, bin) pairs within 1e-9 of the strict oracle -- how often the
                              reference agrees with ITSELF to north_star's tolerance on this workload
    member_max  [M]           per member: its largest relative deviation
    dev_x87     [P][200]      NOT part of the floor: deviation of the same sources compiled for the x87 unit (-mfpmath=387: every
                              intermediate carries a 64-bit significand) -- how far the reference's own double rounding is
                              from a more exact evaluation of its formulas
The flavour fluxes are the |U|^2 projection of `mass` (nuSIprop.hpp:334-336) and are not stored.
"""
import argparse
import os
import sys
import time
from multiprocessing import Pool

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

MEMBERS = ["_fast", "_ulp", "_ulp2", "_ulp3", "_ulp4", "_fastulp"]
EXTRA = "_x87"  # not a member of the floor: the reference on the x87 unit (64-bit significands in every expression), reported beside it
CFG = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
MNTOT = 0.1


def scan_points():
    i, j = np.meshgrid(np.arange(256), np.arange(256), indexing="ij")
    return (10.0 ** (1.0 + 3.0 * i / 255.0)).ravel(), (10.0 ** (-5.0 + 5.0 * j / 255.0)).ravel()


def sample_indices(count):
    return np.random.default_rng(20261104).choice(65536, 1024, replace=False)[:count]


def _one(args):
    import oracle_ref as R
    mphi, g = args
    fm = R.evolve(mphi, g, MNTOT, **CFG)[0]
    dev = []
    for v in MEMBERS:
        fv = R.evolve(mphi, g, MNTOT, variant=v, **CFG)[0]
        dev.append(R.relerr(fv, fm).max(axis=0))
    dx = R.relerr(R.evolve(mphi, g, MNTOT, variant=EXTRA, **CFG)[0], fm).max(axis=0)
    return fm, np.array(dev), dx


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=1024)
    ap.add_argument("--jobs", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "oracle_scan1024.npz"))
    a = ap.parse_args()
    import oracle_ref as R
    R.build_oracle()
    mphi, g = scan_points()
    idx = sample_indices(a.points)
    t0 = time.time()
    with Pool(a.jobs) as pool:
        res = pool.map(_one, [(float(mphi[i]), float(g[i])) for i in idx], chunksize=4)
    mass = np.array([r[0] for r in res])
    dev = np.array([r[1] for r in res]).transpose(1, 0, 2)  # [M][P][N]
    dev = np.where(np.isfinite(dev), dev, 3e38)
    dx = np.array([r[2] for r in res])
    dx = np.where(np.isfinite(dx), dx, 3e38)
    np.savez_compressed(a.out, index=idx.astype(np.int32), mphi=mphi[idx], g=g[idx], mass=mass,
                        floor=dev.max(axis=0).astype(np.float32), members=np.array(MEMBERS),
                        member_strict_frac=np.array([np.mean(d <= 1e-9) for d in dev]),
                        member_max=np.array([d[d < 1e38].max() for d in dev]), dev_x87=dx.astype(np.float32))
    print("wrote %s: %d points in %.0f s; fraction of (point, bin) where every member is within 1e-9: %.4f" %
          (a.out, len(idx), time.time() - t0, np.mean(dev.max(axis=0) <= 1e-9)))


if __name__ == "__main__":
    main()
