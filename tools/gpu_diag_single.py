import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import nusiprop_b200 as nb, oracle_ref as R
G = os.path.join(ROOT, "tests", "golden")
z = np.load(G + "/oracle_flux.npz"); fl = np.load(G + "/oracle_floor.npz"); meta = json.load(open(G + "/oracle_flux_cases.json"))
for name, c in meta.items():
    c = dict(c); mphi, g, mntot = c.pop("mphi"), c.pop("g"), c.pop("mntot")
    out = []
    for mode in (0, 1):
        ctx = nb.Context(phiphi=False, elementary_mode=mode, **c)
        fm, ff, st = ctx.evolve(mphi, g, mntot); ctx.close()
        r = np.maximum(R.relerr(fm[0], z[name + "/mass"]), R.relerr(ff[0], z[name + "/fla"])).max(axis=0)
        f = np.maximum(R.relerr(fl[name + "/mass"], z[name + "/mass"]), R.relerr(fl[name + "/fla"], z[name + "/fla"])).max(axis=0)
        fs = np.maximum(R.noise_floor(z[name + "/mass"], [fl[name + "/mass"], fl[name + "/mass_ulp"]]), R.noise_floor(z[name + "/fla"], [fl[name + "/fla"], fl[name + "/fla_ulp"]]))
        out.append("m%d: max %.1e n>1e-9 %3d beyond10xfloor %3d beyond4xsmooth %3d worst r/fs %.1f" % (mode, r.max(), (r > 1e-9).sum(), (r > np.maximum(1e-9, 10 * f)).sum(), (r > np.maximum(1e-9, 4 * fs)).sum(), (r / np.maximum(fs, 1e-10)).max()))
    print("%-36s floor max %.1e n>1e-9 %3d | %s | %s" % (name, f.max(), (f > 1e-9).sum(), out[0], out[1]))
