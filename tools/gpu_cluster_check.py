#!/usr/bin/env python
"""Cluster integrator (integrate_cluster_kernel: opts.integrator = 0 for N_bins_E > 224, or 4) against the diagonal-owned
kernel (3) at several fine binnings: largest relative difference, finite masks, and the integrator's CUDA-event time for
one point and for a batch.  Usage (GPU box):  python tools/gpu_cluster_check.py [N_E ...]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nusiprop_b200 as nb  # noqa: E402

sizes = [int(a) for a in sys.argv[1:]] or [300, 500, 900, 1100, 2000]
rng = np.random.default_rng(11)
for N in sizes:
    npts = 6 if N <= 1100 else 3
    mphi, g = 10 ** rng.uniform(1, 4, npts), 10 ** rng.uniform(-5, 0, npts)
    mphi[0], g[0] = 100.0, 1e-2
    cfg = dict(N_bins_E=N, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    res, tms = {}, {}
    for integ in (3, 4):
        ctx = nb.Context(integrator=integ, **cfg)
        t0 = time.time()
        for _ in range(2):
            res[integ] = ctx.evolve(mphi, g, 0.1)
            tms[integ] = ctx.last_timing()["ms_integrate"]
        one = ctx.evolve(mphi[:1], g[:1], 0.1)
        t1 = ctx.last_timing()["ms_integrate"]
        S = ctx.N_steps_z - 1
        ctx.close()
        tms[integ] = (tms[integ], t1)
        assert np.array_equal(one[0][0], res[integ][0][0], equal_nan=True), "one point alone differs from the same point in a batch"
    a, b = res[4][0], res[3][0]
    ok = np.isfinite(b) & (b != 0)
    rel = np.max(np.abs(a[ok] - b[ok]) / np.abs(b[ok])) if ok.any() else 0.0
    print("N=%d S=%d: cluster vs diag max rel %.2e, same finite mask %s, zeros equal %s, bit-identical %.3f, fla max rel %.2e | integrate ms (%d points / 1 point): diag %.2f / %.2f, cluster %.2f / %.2f" % (
        N, S, rel, np.array_equal(np.isfinite(a), np.isfinite(b)), np.array_equal(a == 0, b == 0), np.mean(a[ok] == b[ok]),
        np.max(np.abs(res[4][1][ok] - res[3][1][ok]) / np.abs(res[3][1][ok])), npts, tms[3][0], tms[3][1], tms[4][0], tms[4][1]), flush=True)
