import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import nusiprop_b200 as nb
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
n = 4096
mp = 10 ** np.random.default_rng(0).uniform(1, 4, n); gg = 10 ** np.random.default_rng(1).uniform(-5, 0, n)
for mode in ([int(a) for a in sys.argv[1:]] or [1]):
    c = nb.Context(elementary_mode=mode, **cfg)
    P = c.upload_points(mp, gg, 0.1)
    fm = torch.empty((n, 3, 200), dtype=torch.float64, device="cuda"); ff = torch.empty_like(fm); st = torch.zeros(n, dtype=torch.int32, device="cuda")
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.time()
        c.evolve_points(P, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), 0)
        torch.cuda.synchronize(); dt = time.time() - t0
    print("mode %d n=%d wall %.4f s -> %.1f ev/s ; %s  checksum %.17g" % (mode, n, dt, n / dt, c.last_timing(), float(torch.nansum(ff).item())))
    P.close(); c.close()
