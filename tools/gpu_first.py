"""First GPU session: smoke, parity of fluxes / tables / source at N_E=200, FP64 peak, timing."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
ge.smoke()
import nusiprop_b200 as nb
import oracle_ref as R
out = {}
print("fp64 peak TFLOP/s:", nb.measure_fp64_peak()); sys.stdout.flush()
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
# source term
ctx = nb.Context(**cfg)
s_dev = ctx.source(); s_ref = R.source(200, 4.0, 9.0, 5.0)
r = R.relerr(s_dev, s_ref)
print("source: max rel %.3e  per-bin max over z (first 50 bins):" % r.max(), np.array2string(r.max(axis=0)[:50], precision=1))
pts = [(100, 1e-2), (30, 1e-1), (300, 1.0), (1e3, 1e-2)]
for maj, nr, no in [(1, 1, 1), (0, 1, 1), (1, 0, 1), (1, 1, 0)]:
    c = nb.Context(majorana=maj, non_resonant=nr, normal_ordering=no, **cfg)
    t0 = time.time()
    fm, ff, st = c.evolve([p[0] for p in pts], [p[1] for p in pts], 0.1)
    dt = time.time() - t0
    for i, (mphi, g) in enumerate(pts):
        rm, rf, _ = R.evolve(mphi, g, 0.1, majorana=maj, non_resonant=nr, normal_ordering=no, **cfg)
        e = np.maximum(R.relerr(fm[i], rm), R.relerr(ff[i], rf)).max(axis=0)
        print("maj=%d nr=%d no=%d mphi=%g g=%g: max rel %.3e  bins>1e-9: %d/200  status=%d  worst bin %d" % (maj, nr, no, mphi, g, e.max(), (e > 1e-9).sum(), st[i], e.argmax()))
    print("   host evolve of 4 points took %.3f s, timing %s" % (dt, c.last_timing())); sys.stdout.flush()
    if (maj, nr, no) == (1, 1, 1):
        G, A, T, st = c.tables(100.0, 1e-2, 0.1)
        G0, A0, T0 = R.tables(100.0, 1e-2, 0.1, **cfg)
        iu = np.triu_indices(c.NT, 1)
        rt = R.relerr(T[0], T0)[iu]
        print("   tables(100,1e-2): Gamma %.2e aT %.2e alpha max %.2e med %.2e >1e-9 %d/%d" % (R.relerr(G[0], G0).max(), R.relerr(A[0], A0).max(), rt.max(), np.median(rt), (rt > 1e-9).sum(), rt.size))
    c.close()
# timing on a batch
import torch
c = nb.Context(**cfg)
for n in (148, 1184, 4096):
    mphi = 10 ** np.random.default_rng(0).uniform(1, 4, n); g = 10 ** np.random.default_rng(1).uniform(-5, 0, n)
    P = c.upload_points(mphi, g, 0.1)
    fm = torch.empty((n, 3, 200), dtype=torch.float64, device="cuda"); ff = torch.empty_like(fm); st = torch.zeros(n, dtype=torch.int32, device="cuda")
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.time()
        c.evolve_points(P, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), 0)
        torch.cuda.synchronize(); dt = time.time() - t0
        print("n=%d rep=%d wall %.4f s  -> %.1f evolutions/s ; %s" % (n, rep, dt, n / dt, c.last_timing())); sys.stdout.flush()
    P.close()
