#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/ from the ORACLE (unmodified reference
headers + refshim, oracle/_ref) and from mpmath.  Run in the container that has /root/reference:

    make -C oracle && python tools/gen_golden.py

  specfun_mpmath.json   50-digit mpmath values of Re Li2(x), complex Li2(z), Li2/Li3 on [-1,0]
  oracle_flux.npz       evolve() outputs of the oracle for the BASELINE configs at parity-test sizes
  oracle_tables.npz     Gamma / alphaTilde / alpha tables of the oracle for a few points (N_E=40)
  oracle_floor.npz      the same flux cases from two benign variants of the SAME reference sources: rebuilt with
                        -O2 -mfma -ffp-contract=fast ("<case>/mass", "/fla") and with +-1 ulp applied to a quarter of
                        its libm results ("<case>/mass_ulp", "/fla_ulp"): the reference's own noise floor
"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_ref as R
import mpmath as mp

OUT = os.path.join(ROOT, "tests", "golden")
mp.mp.dps = 50


def specfun():
    rng = np.random.default_rng(20261104)
    xs = np.concatenate([rng.uniform(-3, 3, 120), -10 ** rng.uniform(-8, 6, 60), 10 ** rng.uniform(-8, 6, 60),
                         [0.5, 1.0, 2.0, -1.0, 0.25, 1.01, 1.0000001, 0.9999999, 0.75, 0.8, 100.0, -100.0]])
    real = [[float(x), float(mp.polylog(2, mp.mpf(float(x))).real)] for x in xs]
    pts = []
    for _ in range(250):
        r = 10 ** rng.uniform(-6, 4); th = rng.uniform(-np.pi, np.pi)
        pts.append((r * np.cos(th), r * np.sin(th)))
    for _ in range(100):
        pts.append((rng.uniform(-3, 3), 10 ** rng.uniform(-12, 0) * rng.choice([-1, 1])))
    for _ in range(60):
        th = rng.uniform(-np.pi, np.pi); r = 1 + rng.uniform(-0.03, 0.03)
        pts.append((r * np.cos(th), r * np.sin(th)))
    for x in [-2.0, -0.5, 0.3, 0.9, 1.0, 1.5, 7.0]:
        pts.append((x, 0.0))
    cplx = []
    for x, y in pts:
        x, y = float(x), float(y)
        if y == 0.0:  # GSL convention on the real axis
            v = mp.mpc(mp.polylog(2, mp.mpf(x)).real, -mp.pi * mp.log(x) if x >= 1 else 0)
        else:
            v = mp.polylog(2, mp.mpc(x, y))
        cplx.append([x, y, float(v.real), float(v.imag)])
    neg = -10 ** rng.uniform(-300, 0, 150)
    neg = np.concatenate([neg, [-1.0, -0.5, -0.999, -1e-300]])
    li2n = [[float(x), float(mp.polylog(2, mp.mpf(float(x))))] for x in neg]
    li3n = [[float(x), float(mp.polylog(3, mp.mpf(float(x))))] for x in neg]
    json.dump({"li2_real": real, "cli2": cplx, "li2_neg": li2n, "li3_neg": li3n}, open(os.path.join(OUT, "specfun_mpmath.json"), "w"))


FLUX_CASES = {
    # name: (mphi, g, mntot, cfg)
    "testpy_massless": (5e6, 1e-6, float(0.0 + np.sqrt(7.42e-5) + np.sqrt(2.514e-3)),
                        dict(majorana=True, non_resonant=False, normal_ordering=True, N_bins_E=100, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)),
    "testcpp_default": (6e5, 0.01, 0.1, dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=100, lEmin=9.0, lEmax=14.0, zmax=5.0, flav=2)),
}
for (mphi, g) in [(100.0, 1e-2), (30.0, 1e-1), (300.0, 1.0), (1e3, 1e-2)]:
    for maj in (True, False):
        for nr in (True, False):
            for no in (True, False):
                FLUX_CASES["single_m%g_g%g_maj%d_nr%d_no%d" % (mphi, g, maj, nr, no)] = (
                    mphi, g, 0.1, dict(majorana=maj, non_resonant=nr, normal_ordering=no, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2))
FLUX_CASES["flav0_N60"] = (100.0, 1e-2, 0.1, dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=60, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=0))
FLUX_CASES["flav1_mn0.3_N60"] = (50.0, 3e-2, 0.3, dict(majorana=True, non_resonant=True, normal_ordering=False, N_bins_E=60, lEmin=4.0, lEmax=9.0, zmax=3.0, flav=1))
FLUX_CASES["default_binning_N300"] = (1e7, 0.1, 0.1, dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=300, lEmin=12.0, lEmax=17.0, zmax=5.0, flav=2))


def fluxes():
    out, floor, meta = {}, {}, {}
    for name, (mphi, g, mntot, cfg) in FLUX_CASES.items():
        fm, ff, en = R.evolve(mphi, g, mntot, **cfg)
        fm2, ff2, _ = R.evolve(mphi, g, mntot, variant="_fast", **cfg)
        fm3, ff3, _ = R.evolve(mphi, g, mntot, variant="_ulp", **cfg)
        out[name + "/mass"], out[name + "/fla"], out[name + "/E"] = fm, ff, en
        floor[name + "/mass"], floor[name + "/fla"] = fm2, ff2
        floor[name + "/mass_ulp"], floor[name + "/fla_ulp"] = fm3, ff3
        meta[name] = dict(mphi=mphi, g=g, mntot=mntot, **cfg)
        print(name, "floor max rel: fma %.2e  ulp %.2e" % (max(R.relerr(fm2, fm).max(), R.relerr(ff2, ff).max()),
                                                           max(R.relerr(fm3, fm).max(), R.relerr(ff3, ff).max())))
    np.savez_compressed(os.path.join(OUT, "oracle_flux.npz"), **out)
    np.savez_compressed(os.path.join(OUT, "oracle_floor.npz"), **floor)
    json.dump(meta, open(os.path.join(OUT, "oracle_flux_cases.json"), "w"), indent=1)


def tables():
    out = {}
    cfg = dict(N_bins_E=40, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    for (mphi, g) in [(100.0, 1e-2), (30.0, 1e-1), (1e3, 1e-2)]:
        for maj, nr, no in [(1, 1, 1), (0, 1, 1), (1, 0, 1), (1, 1, 0)]:
            G, A, T = R.tables(mphi, g, 0.1, majorana=maj, non_resonant=nr, normal_ordering=no, **cfg)
            key = "m%g_g%g_maj%d_nr%d_no%d" % (mphi, g, maj, nr, no)
            out[key + "/Gamma"], out[key + "/alphaTilde"], out[key + "/alpha"] = G, A, T
    np.savez_compressed(os.path.join(OUT, "oracle_tables.npz"), **out)
    src = R.source(200, 4.0, 9.0, 5.0)
    np.savez_compressed(os.path.join(OUT, "oracle_source_N200.npz"), source=src)


if __name__ == "__main__":
    R.build_oracle()
    specfun()
    fluxes()
    tables()
    print("golden fixtures written to", OUT)
