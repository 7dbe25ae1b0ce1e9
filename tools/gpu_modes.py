"""Flux parity of the alpha-builder modes against the oracle, next to the reference's own noise floor
(strict oracle vs the same reference built with -O2 -mfma -ffp-contract=fast), on a seeded sample of
the S-scan; and timing per mode."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import nusiprop_b200 as nb
import oracle_ref as R
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, ROOT)
import bench
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
npts = int(sys.argv[1]) if len(sys.argv) > 1 else 48
pts = bench.cpu_sample_points(npts)
mphi = np.array([p[0] for p in pts]); g = np.array([p[1] for p in pts])
t0 = time.time()
ref = [R.evolve(m, gg, 0.1, **cfg)[1] for m, gg in pts]
reff = [R.evolve(m, gg, 0.1, variant="_fast", **cfg)[1] for m, gg in pts]
print("oracle time %.1fs" % (time.time() - t0))
ref = np.array(ref); reff = np.array(reff)
floor = R.relerr(reff, ref).max(axis=1)  # [pts, N]
res = {}
for mode in (0, 1, 2):
    c = nb.Context(elementary_mode=mode, **cfg)
    fm, ff, st = c.evolve(mphi, g, 0.1)
    res[mode] = R.relerr(ff, ref).max(axis=1)
    c.close()
print("%-22s %10s %10s %10s %10s" % ("point", "ref-floor", "mode0", "mode1", "mode2"))
for i in np.argsort(mphi):
    print("mphi=%8.1f g=%8.2e %10.1e %10.1e %10.1e %10.1e   bins>1e-9: %3d %3d %3d %3d" % (mphi[i], g[i], floor[i].max(), res[0][i].max(), res[1][i].max(), res[2][i].max(),
          (floor[i] > 1e-9).sum(), (res[0][i] > 1e-9).sum(), (res[1][i] > 1e-9).sum(), (res[2][i] > 1e-9).sum()))
tot = floor.size
print("fraction of (point,bin) within 1e-9: ref-vs-ref %.4f  mode0 %.4f mode1 %.4f mode2 %.4f" % tuple((x <= 1e-9).sum() / tot for x in (floor, res[0], res[1], res[2])))
for k in (1, 4, 10):
    tolr = np.maximum(1e-9, k * floor)
    print("within max(1e-9, %d x floor): mode0 %.4f mode1 %.4f mode2 %.4f" % ((k,) + tuple((res[m] <= tolr).sum() / tot for m in (0, 1, 2))))
n = 4096
mp = 10 ** np.random.default_rng(0).uniform(1, 4, n); gg = 10 ** np.random.default_rng(1).uniform(-5, 0, n)
for mode in (0, 1, 2):
    c = nb.Context(elementary_mode=mode, **cfg)
    P = c.upload_points(mp, gg, 0.1)
    fm = torch.empty((n, 3, 200), dtype=torch.float64, device="cuda"); ff = torch.empty_like(fm); st = torch.zeros(n, dtype=torch.int32, device="cuda")
    for rep in range(2):
        torch.cuda.synchronize(); t0 = time.time()
        c.evolve_points(P, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), 0)
        torch.cuda.synchronize(); dt = time.time() - t0
    print("mode %d n=%d wall %.4f s -> %.1f ev/s ; %s" % (mode, n, dt, n / dt, c.last_timing()))
    P.close(); c.close()
