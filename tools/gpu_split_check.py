#!/usr/bin/env python
"""Role-split integrator (opts.integrator = 0) against the diagonal-owned (3) and the serial reference-shaped (2) kernels:
largest relative difference per configuration, and the integrator's CUDA-event time on an 8192-point shard."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nusiprop_b200 as nb
import bench

rng = np.random.default_rng(5)
mphi, g = 10 ** rng.uniform(1, 4, 64), 10 ** rng.uniform(-5, 0, 64)
cfgs = [dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2), dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2, majorana=0),
        dict(N_bins_E=97, lEmin=5.0, lEmax=8.0, zmax=2.0, flav=1), dict(N_bins_E=120, lEmin=5.0, lEmax=8.5, zmax=3.0, flav=1),
        dict(N_bins_E=60, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=0), dict(N_bins_E=224, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2), dict(N_bins_E=33, lEmin=4.0, lEmax=6.0, zmax=5.0, flav=2),
        dict(N_bins_E=100, lEmin=9.0, lEmax=14.0, zmax=5.0, flav=2), dict(N_bins_E=210, lEmin=4.0, lEmax=7.0, zmax=6.0, flav=2)]
for cfg in cfgs:
    res = {}
    for integ in (0, 3, 2):
        ctx = nb.Context(integrator=integ, **cfg)
        res[integ] = ctx.evolve(mphi, g, 0.1)
        S = ctx.N_steps_z - 1
        ctx.close()
    a, b, c = res[0][0], res[3][0], res[2][0]
    ok = np.isfinite(c) & (c != 0)
    same_nan = np.array_equal(np.isfinite(a), np.isfinite(c))
    r_ab = np.max(np.abs(a[ok] - b[ok]) / np.abs(b[ok])) if ok.any() else 0
    r_ac = np.max(np.abs(a[ok] - c[ok]) / np.abs(c[ok])) if ok.any() else 0
    r_bc = np.max(np.abs(b[ok] - c[ok]) / np.abs(c[ok])) if ok.any() else 0
    print("N=%d S=%d maj=%s: split vs diag %.2e, split vs serial %.2e, diag vs serial %.2e, same finite mask %s, bitwise-equal frac %.3f, status equal %s" % (
        cfg["N_bins_E"], S, cfg.get("majorana", 1), r_ab, r_ac, r_bc, same_nan, np.mean(a[ok] == c[ok]), np.array_equal(res[0][2], res[2][2])), flush=True)
# timing on the benchmark shard
m, gg = bench.shard(0)
ref = None
for integ, var in ((3, 0), (0, 0), (0, 21), (0, 22), (0, 23), (0, 11), (0, 13)):
    ctx = nb.Context(integrator=integ, fine_variant=var, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    for _ in range(3):
        out = ctx.evolve(m, gg, 0.1)
        tm = ctx.last_timing()
    if ref is None:
        ref = out[0]
    print("integrator %d variant %d: integrate %.3f ms, max rel diff vs diag %.2e" % (integ, var, tm["ms_integrate"], np.max(np.abs(out[0] - ref) / np.abs(ref))), flush=True)
    ctx.close()
