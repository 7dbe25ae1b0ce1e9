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

sizes = [int(a) for a in sys.argv[1:] if a.isdigit()] or [300, 500, 900, 1100, 2000]
# batches: where the cluster kernel (4 SMs per point) beats the one-CTA-per-point kernel
if "--batch" in sys.argv:
    for N in (600, 700, 800, 993):
        cfg = dict(N_bins_E=N, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
        line = "N=%d:" % N
        for npts in (8, 32, 148, 592, 4096):
            mphi, g = np.repeat(10 ** np.linspace(1.5, 3.5, (npts + 7) // 8), 8)[:npts], np.tile(10 ** np.linspace(-4, -1, 8), (npts + 7) // 8)[:npts]
            t = {}
            for integ, var in ((3, 0), (4, 0), (4, 12)):
                ctx = nb.Context(integrator=integ, fine_variant=var, **cfg)
                for _ in range(2):
                    ctx.evolve(mphi, g, 0.1)
                    t[(integ, var)] = ctx.last_timing()["ms_integrate"]
                ctx.close()
            line += "  %d pts diag %.2f cluster %.2f (clusters of 4: %.2f)" % (npts, t[(3, 0)], t[(4, 0)], t[(4, 12)])
        print(line, flush=True)
    sys.exit(0)

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

# role ablation at N_E = 2000 (timing switches, wrong results): where a block's time goes
if 2000 in sizes:
    for var in (0, 11, 21, 22, 23):
        ctx = nb.Context(integrator=4, fine_variant=var, N_bins_E=2000, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
        for _ in range(2):
            ctx.evolve([100.0], [1e-2], 0.1)
            t1 = ctx.last_timing()["ms_integrate"]
        ctx.evolve(np.full(37, 100.0), 10 ** np.linspace(-4, -1, 37), 0.1)
        t37 = ctx.last_timing()["ms_integrate"]
        ctx.evolve(np.full(148, 100.0), 10 ** np.linspace(-4, -1, 148), 0.1)
        t148 = ctx.last_timing()["ms_integrate"]
        ctx.close()
        print("variant %d (20 + bits: 1 no bulk, 2 no chain, 4 no helper, 8 no table copies): 1 point %.2f ms, 37 points %.2f ms, 148 points %.2f ms" % (var, t1, t37, t148), flush=True)
