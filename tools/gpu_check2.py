"""wave integrator vs serial integrator (bitwise), tiled vs per-entry alpha (bitwise), timing."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import nusiprop_b200 as nb
import oracle_ref as R
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
rng = np.random.default_rng(5)
n = 64
mphi = 10 ** rng.uniform(1, 4, n); g = 10 ** rng.uniform(-5, 0, n)
for maj in (1, 0):
    c = nb.Context(majorana=maj, **cfg)
    os.environ.pop("NUSI_SERIAL_INTEGRATOR", None)
    fm1, ff1, st1 = c.evolve(mphi, g, 0.1)
    os.environ["NUSI_SERIAL_INTEGRATOR"] = "1"
    fm2, ff2, st2 = c.evolve(mphi, g, 0.1)
    os.environ.pop("NUSI_SERIAL_INTEGRATOR", None)
    print("maj=%d wave vs serial integrator: bitwise equal mass=%s fla=%s  max rel %.3e" % (maj, np.array_equal(fm1, fm2, equal_nan=True), np.array_equal(ff1, ff2, equal_nan=True), np.nanmax(R.relerr(fm1, fm2))))
    c.close()
os.environ["NUSI_NO_TILES"] = "1"
c0 = nb.Context(**cfg)
os.environ.pop("NUSI_NO_TILES")
c1 = nb.Context(**cfg)
G0, A0, T0, s0 = c0.tables(mphi[:4], g[:4], 0.1); G1, A1, T1, s1 = c1.tables(mphi[:4], g[:4], 0.1)
print("tiled vs per-entry alpha tables bitwise:", np.array_equal(T0, T1, equal_nan=True), "rows:", np.array_equal(G0, G1), np.array_equal(A0, A1))
# oracle parity on a few points
for i in range(3):
    rm, rf, _ = R.evolve(mphi[i], g[i], 0.1, **cfg)
    fm, ff, st = c1.evolve(mphi[i], g[i], 0.1)
    print("point mphi=%.3g g=%.3g: max rel vs oracle %.3e" % (mphi[i], g[i], max(R.relerr(fm[0], rm).max(), R.relerr(ff[0], rf).max())))
# N_E=2000 path smoke (one point) + N=300 default
for N, lo, hi in ((300, 12.0, 17.0), (2000, 4.0, 9.0)):
    cc = nb.Context(N_bins_E=N, lEmin=lo, lEmax=hi, zmax=5.0)
    t0 = time.time(); fm, ff, st = cc.evolve(100.0 if N == 2000 else 1e7, 1e-2 if N == 2000 else 0.1, 0.1); dt = time.time() - t0
    print("N_E=%d: NT=%d Nz=%d finite=%s status=%s wall %.3fs timing %s" % (N, cc.NT, cc.N_steps_z, np.isfinite(fm).all(), st, dt, cc.last_timing()))
    cc.close()
for n in (1184, 4096):
    mp = 10 ** np.random.default_rng(0).uniform(1, 4, n); gg = 10 ** np.random.default_rng(1).uniform(-5, 0, n)
    P = c1.upload_points(mp, gg, 0.1)
    fm = torch.empty((n, 3, 200), dtype=torch.float64, device="cuda"); ff = torch.empty_like(fm); st = torch.zeros(n, dtype=torch.int32, device="cuda")
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.time()
        c1.evolve_points(P, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), 0)
        torch.cuda.synchronize(); dt = time.time() - t0
        print("n=%d rep=%d wall %.4f s  -> %.1f evolutions/s ; %s" % (n, rep, dt, n / dt, c1.last_timing())); sys.stdout.flush()
    P.close()
