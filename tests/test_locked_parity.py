"""Strict parity made provable: the libm-locked pair of builds.

    O2 = oracle/_ref/libnusi_ref_locked.so   the UNMODIFIED reference; every log / log1p / atan / atanh / carg / cabs / clog
                                             of its own code routed to nusiprop_b200/csrc/nusi_locked_math.h (IEEE basic
                                             operations only: the same bits under g++ and nvcc), its two GSL dilogarithms
                                             served by the product's Li2 (one shared __host__ __device__ source)
    P2 = the product's formulas built with -DNUSI_LOCKED_MATH: on the host (tests/hostcheck, g++) and on the GPU
         (nusiprop_b200/lib/libnusiprop_b200_locked.so, nvcc)

With the transcendental functions shared, whatever differs between O2 and P2 is the PRODUCT's arithmetic.  The strict
builder (elementary_mode 0) must then reproduce the reference's tables bit for bit -- on the host and on the GPU -- and the
residual of the fast builders (modes 2 / 3) against it is their measured cost.  Test infrastructure only: the product
library proper contains none of this."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
D, I, DP = C.c_double, C.c_int, C.POINTER(C.c_double)
V = "_locked"
GRID = dict(N_bins_E=40, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
POINTS = [(100.0, 1e-2), (30.0, 1e-1), (1e3, 1e-2), (5e3, 5e-2)]
CONFIGS = [(1, 1, 1), (0, 1, 1), (1, 1, 0)]  # (majorana, non_resonant, normal_ordering)
CASES = [(m, g) + c for (m, g) in POINTS for c in CONFIGS]


def _p(a):
    return a.ctypes.data_as(DP)


# ---------------------------------------------------------------------------------------------------------------------
# the shared elementary functions themselves
# ---------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def lk():
    out = os.path.join(HERE, "hostcheck", "_build")
    os.makedirs(out, exist_ok=True)
    src, so = os.path.join(out, "locked_math_check.cpp"), os.path.join(out, "liblocked_math_check.so")
    open(src, "w").write('#include "nusi_locked_math.h"\nextern "C" {\n' + "".join(
        "double lk_%s(double x) { return nusi_locked::%s(x); }\n" % (f, f) for f in ("log", "log1p", "atan", "atanh")) +
        "double lk_atan2(double y, double x) { return nusi_locked::atan2(y, x); }\n"
        "double lk_hypot(double x, double y) { return nusi_locked::hypot(x, y); }\n}\n")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", os.path.join(ROOT, "nusiprop_b200", "csrc"), "-o", so, src])
    L = C.CDLL(so)
    for n in ("lk_log", "lk_log1p", "lk_atan", "lk_atanh"):
        getattr(L, n).restype, getattr(L, n).argtypes = D, [D]
    for n in ("lk_atan2", "lk_hypot"):
        getattr(L, n).restype, getattr(L, n).argtypes = D, [D, D]
    return L


def test_locked_elementary_functions_are_accurate(lk):
    """< 1.5 ulp against mpmath on the ranges the path uses: the locked oracle is as good a libm as any other."""
    mp = pytest.importorskip("mpmath")
    mp.mp.prec = 200
    rng = np.random.default_rng(1)

    def worst(f, ref, xs):
        w = 0.0
        for x in xs:
            a = f(*x) if isinstance(x, tuple) else f(x)
            r = ref(*[mp.mpf(v) for v in x]) if isinstance(x, tuple) else ref(mp.mpf(x))
            if r != 0:
                w = max(w, float(abs((mp.mpf(a) - r) / mp.mpf(float(np.spacing(abs(float(r))))))))
        return w

    n = 1500
    assert worst(lk.lk_log, mp.log, [float(v) for v in np.r_[10 ** rng.uniform(-300, 300, n), 1 + rng.uniform(-1e-3, 1e-3, n), rng.uniform(0.5, 2, n)]]) < 1.0
    assert worst(lk.lk_log1p, mp.log1p, [float(v) for v in np.r_[rng.uniform(-0.999, 5, n), 10 ** rng.uniform(-30, 10, n), -10 ** rng.uniform(-30, -0.01, n)]]) < 1.0
    assert worst(lk.lk_atan, mp.atan, [float(v) for v in np.r_[rng.uniform(-5, 5, n), 10 ** rng.uniform(-20, 20, n)]]) < 1.0
    yx = [(float(a), float(b)) for a, b in zip(rng.normal(size=n) * 10 ** rng.uniform(-5, 5, n), rng.normal(size=n) * 10 ** rng.uniform(-5, 5, n))]
    assert worst(lk.lk_atan2, lambda y, x: mp.atan2(y, x), yx) < 1.5
    assert worst(lk.lk_atanh, mp.atanh, [float(v) for v in np.r_[rng.uniform(-0.9999, 0.9999, n), 10 ** rng.uniform(-20, -0.001, n)]]) < 1.5
    assert worst(lk.lk_hypot, lambda x, y: mp.sqrt(x * x + y * y), [(float(a), float(b)) for a, b in rng.normal(size=(n, 2))]) < 1.5
    # the conventions the reference relies on
    assert lk.lk_atan2(0.0, -1.0) == np.pi and lk.lk_atan2(-0.0, -1.0) == -np.pi and lk.lk_log(1.0) == 0.0 and lk.lk_log1p(0.0) == 0.0
    assert lk.lk_log(0.0) == -np.inf and np.isnan(lk.lk_log(-1.0))


# ---------------------------------------------------------------------------------------------------------------------
# host: the product's strict formulas (g++) against O2, per table entry
# ---------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def hc_locked():
    out = os.path.join(HERE, "hostcheck", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libnusi_hostcheck_locked.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-DNUSI_HOSTCHECK_XSEC", "-DNUSI_LOCKED_MATH", "-shared", "-fPIC",
                           "-I", os.path.join(ROOT, "nusiprop_b200", "csrc"), "-o", so, os.path.join(HERE, "hostcheck", "hostcheck.cpp")])
    L = C.CDLL(so)
    L.hc_tables.argtypes = [D, D, D, DP, DP, I, I, I, I, DP, DP, DP, DP, DP, C.POINTER(I)]
    return L


def _edges(R, no):
    N = GRID["N_bins_E"]
    nz, Emin, Emax, Ec, z = R.grid(N, GRID["lEmin"], GRID["lEmax"], GRID["zmax"], variant=V)
    NT = N + nz - 2
    lo, hi = np.zeros(NT), np.zeros(NT)
    lo[:N], hi[:N] = Emin, Emax
    for i in range(N, NT):  # nuSIprop.hpp:229-233
        lo[i] = Emin[N - 1] * (1 + z[i - N + 1])
        hi[i] = Emax[N - 1] * (1 + z[i - N + 1])
    return NT, lo, hi


def _oracle_tables(R, mphi, g, maj, nr, no):
    return R.tables(mphi, g, 0.1, bool(maj), bool(nr), bool(no), variant=V, **GRID)


@pytest.mark.parametrize("mphi,g,maj,nr,no", CASES)
def test_host_strict_formulas_equal_the_locked_oracle_bit_for_bit(hc_locked, oracle, mphi, g, maj, nr, no):
    """Gamma, alphaTilde and every alpha entry, including the ill-conditioned corner (5e3 eV, 0.05): the strict transcription
    performs the reference's operations in the reference's order (reference nuSIprop.hpp:759-1520, aux.hpp)."""
    NT, lo, hi = _edges(oracle, no)
    mn, U2f = oracle.masses(0.1, bool(no)), oracle.pmns2(bool(no))[GRID["flav"]].copy()
    Ga = g * g * mphi / (16.0 * np.pi) if maj else g * g * mphi / (8.0 * np.pi)
    G, A, T, w = np.zeros(NT), np.zeros(NT), np.zeros((NT, NT)), I()
    hc_locked.hc_set_shared_logs(0)
    hc_locked.hc_tables(mphi, g, Ga, _p(mn), _p(U2f), maj, nr, 0, NT, _p(lo), _p(hi), _p(G), _p(A), _p(T), C.byref(w))
    G0, A0, T0 = _oracle_tables(oracle, mphi, g, maj, nr, no)
    assert np.array_equal(G, G0) and np.array_equal(A, A0)
    iu = np.triu_indices(NT, 1)
    assert np.array_equal(T[iu], T0[iu]), "alpha: %.4f of the entries bit-identical, max rel %.2e" % (np.mean(T[iu] == T0[iu]), oracle.relerr(T, T0)[iu].max())


def test_locked_oracle_is_an_ordinary_member_of_the_reference_ensemble(oracle):
    """O2 differs from the strict oracle O1 (glibc + the independent GSL stand-in) by libm-level noise only: no more than
    the +-1 ulp ensemble members do."""
    cfg = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=100, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    for mphi, g in ((100.0, 1e-2), (30.0, 0.1), (3000.0, 0.05)):
        a = oracle.evolve(mphi, g, 0.1, **cfg)[0]
        b = oracle.evolve(mphi, g, 0.1, variant=V, **cfg)[0]
        u = max(oracle.relerr(oracle.evolve(mphi, g, 0.1, variant=v, **cfg)[0], a).max() for v in ("_ulp", "_ulp2", "_fast"))
        assert oracle.relerr(b, a).max() <= max(1e-11, 2.0 * u), (mphi, g, oracle.relerr(b, a).max(), u)


# ---------------------------------------------------------------------------------------------------------------------
# GPU: the locked build of the CUDA library against O2
# ---------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def locked_lib(nusi):
    from nusiprop_b200 import _lib, build
    path = build.build_locked_library()
    return _lib.load_path(path)


@pytest.mark.gpu
@pytest.mark.parametrize("mphi,g,maj,nr,no", CASES)
def test_gpu_strict_builder_equals_the_locked_oracle(nusi, locked_lib, oracle, mphi, g, maj, nr, no):
    """elementary_mode 0 on the GPU (nvcc, -fmad=false) against O2 (g++, -ffp-contract=off), per table entry: <= 1e-13
    demanded (VERDICT), bit-identical expected -- both sides run the same source for every transcendental."""
    ctx = nusi.Context(majorana=maj, non_resonant=nr, normal_ordering=no, elementary_mode=0, _library=locked_lib, **GRID)
    G, A, T, st = ctx.tables([mphi], [g], 0.1)
    ctx.close()
    G0, A0, T0 = _oracle_tables(oracle, mphi, g, maj, nr, no)
    iu = np.triu_indices(T0.shape[0], 1)
    rel = max(oracle.relerr(G[0], G0).max(), oracle.relerr(A[0], A0).max(), oracle.relerr(T[0], T0)[iu].max())
    same = np.mean(np.r_[G[0] == G0, A[0] == A0, T[0][iu] == T0[iu]])
    print("GPU strict vs O2 (%g, %g, maj %d, no %d): max rel %.2e, %.5f of the entries bit-identical" % (mphi, g, maj, no, rel, same))
    assert rel <= 1e-13
    assert same == 1.0


SCAN16 = [(m, g) for m in (10.0, 100.0, 3e3, 1e4) for g in (1e-5, 1e-3, 3e-2, 1.0)]
CFG200 = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)


@pytest.mark.gpu
def test_gpu_fluxes_against_the_locked_oracle_on_16_scan_points(nusi, locked_lib, oracle):
    """BASELINE configs[4] binning, 16 points of the scan plane including m_phi >= 3e3 eV where the reference is
    ill-conditioned.  With libm noise removed:
      strict builder + reference-shaped serial integrator   the reference's evolve() on the GPU: <= 1e-13
      strict builder + default (role-split) integrator      + rounding of the split regeneration sums: <= 1e-12
      default builder (mode 3)                              the measured cost of the fast builder -- reported, and gated by
                                                            the same per-bin criterion as the headline scan."""
    mphi, g = np.array([p[0] for p in SCAN16]), np.array([p[1] for p in SCAN16])
    ref = np.array([oracle.evolve(float(a), float(b), 0.1, variant=V, **CFG200)[0] for a, b in SCAN16])
    out = {}
    for name, kw in (("strict_serial", dict(elementary_mode=0, integrator=2)), ("strict", dict(elementary_mode=0)), ("default", dict())):
        ctx = nusi.Context(_library=locked_lib, **CFG200, **kw)
        fm, ff, st = ctx.evolve(mphi, g, 0.1)
        ctx.close()
        assert np.isfinite(fm).all() and np.all(st & ~7 == 0)
        out[name] = oracle.relerr(fm, ref).max(axis=(1, 2))
        print("%-14s vs O2: worst point %.2e, per point %s" % (name, out[name].max(), " ".join("%.0e" % v for v in out[name])))
    assert out["strict_serial"].max() <= 1e-13
    assert out["strict"].max() <= 1e-12
    # the fast builder: within 1e-9 wherever the reference itself is (its +-1 ulp / FMA members), within 4x its own spread elsewhere
    for i, (a, b) in enumerate(SCAN16):
        if out["default"][i] > 1e-9:
            base = oracle.evolve(a, b, 0.1, **CFG200)[0]
            spread = max(oracle.relerr(oracle.evolve(a, b, 0.1, variant=v, **CFG200)[0], base).max() for v in ("_fast", "_ulp", "_ulp2", "_fastulp"))
            assert out["default"][i] <= min(0.1, 4.0 * spread), (a, b, out["default"][i], spread)
