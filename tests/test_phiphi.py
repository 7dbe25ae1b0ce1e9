"""Double-scalar (phi phi) channel, BASELINE configs[2]: the CUDA table generator
against the reference's own integrand + scipy dblquad (xsec/funcs.c, xsec/tables_phiphi.py), and
evolve() with phiphi = true against the oracle reading THE SAME table files."""
import ctypes as C
import os
import subprocess
import warnings

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
D = C.c_double


@pytest.fixture(scope="module")
def hcpp():
    """Host compile of the generator's __host__ __device__ quadrature (test aid, never loaded by the package)."""
    out = os.path.join(HERE, "hostcheck", "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libnusi_hostcheck_pp.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", os.path.join(ROOT, "nusiprop_b200", "csrc"),
                           "-o", so, os.path.join(HERE, "hostcheck", "hostcheck.cpp")])
    L = C.CDLL(so)
    L.hc_pp_integrand.restype = L.hc_pp_alpha_node.restype = L.hc_pp_alphatilde_node.restype = D
    L.hc_pp_integrand.argtypes = [D, D]
    L.hc_pp_alpha_node.argtypes = [D, D, D]
    L.hc_pp_alphatilde_node.argtypes = [D, D]
    return L


@pytest.fixture(scope="module")
def ref_integrand(oracle):
    p = os.path.join(ROOT, "oracle", "_ref", "funcs.so")
    if not os.path.exists(p):
        pytest.skip("oracle/_ref/funcs.so not built")
    lib = C.CDLL(p)
    f = lib["dsigma_phiphi_over_tauphi"]  # the signature scipy recognises as double f(int, double*), tables_phiphi.py:15-17
    f.restype = D
    f.argtypes = (C.c_int, D)
    direct = lib["dsigma_phiphi_over_tauphi"]
    direct.restype = D
    direct.argtypes = (C.c_int, C.POINTER(D))
    f.direct = direct
    return f


def _ref_alpha(f, s, n, d):  # xsec/tables_phiphi.py:44-57, tighter tolerances than its defaults
    import scipy.integrate as integ
    delta = 10 ** d
    smin = s / delta
    tmin = -smin / delta ** n
    tplus = tmin * delta
    v = integ.dblquad(f, tplus, tmin, lambda t: max(smin, 4), lambda t: s, epsabs=1e-14, epsrel=1e-11)[0]
    return 0.0 if v < 1e-37 else v


def _ref_alphatilde(f, t, d):  # xsec/tables_phiphi.py:25-34
    import scipy.integrate as integ
    delta = 10 ** d
    return integ.dblquad(f, t, t / delta, lambda x: max(-x, 4, -x ** 2 / (1 + x)), lambda x: -t, epsabs=1e-14, epsrel=1e-11)[0]


def test_generator_integrand_is_the_reference_integrand(hcpp, ref_integrand):
    rng = np.random.default_rng(3)
    for s, t in zip(10 ** rng.uniform(np.log10(4.0001), 4, 400), -10 ** rng.uniform(-2, 4, 400)):
        a = hcpp.hc_pp_integrand(s, t)
        b = ref_integrand.direct(2, (D * 2)(s, t))
        assert a == pytest.approx(b, rel=1e-12, abs=1e-300)


def test_generator_quadrature_vs_scipy_dblquad(hcpp, ref_integrand):
    """Fixed Gauss-Legendre panels vs adaptive QUADPACK on the reference's integrand, including nodes
    where the kinematic boundary s*(t) cuts through the integration box (n = 1) and the threshold s -> 4."""
    warnings.simplefilter("ignore")
    for s in (5.0, 12.0, 40.0, 300.0, 1e4):
        for n in (1, 2, 5, 40):
            for d in (0.005, 0.05):
                b = _ref_alpha(ref_integrand, s, n, d)
                a = hcpp.hc_pp_alpha_node(s, float(n), d)
                assert a == pytest.approx(b, rel=2e-6, abs=1e-13), (s, n, d)
    for t in (-4.0, -12.0, -30.0, -100.0, -1e4):
        for d in (0.005, 0.05):
            assert hcpp.hc_pp_alphatilde_node(-t, d) == pytest.approx(_ref_alphatilde(ref_integrand, t, d), rel=1e-7, abs=1e-15), (t, d)


def _read_table(path, D_):
    return np.fromfile(path, dtype=np.float32).reshape(-1, D_ + 1)


@pytest.mark.gpu
def test_generator_small_tables_format_and_values(nusi, hcpp, tmp_path):
    at, al = str(tmp_path / "at.bin"), str(tmp_path / "al.bin")
    nusi.generate_phiphi_tables(at, al, n_t=40, n_s=24, n_n=12, n_d=6)
    A, B = _read_table(at, 2), _read_table(al, 3)
    assert A.shape == (40 * 6, 3) and B.shape == (24 * 12 * 6, 4)
    # record order: last coordinate fastest (interp.hpp:259-290)
    assert np.all(A[:6, 0] == A[0, 0]) and np.all(np.diff(A[:6, 1]) > 0) and A[0, 0] == 4.0 and A[-1, 0] == 1e4
    assert np.array_equal(B[:6 * 12, 1].reshape(12, 6)[:, 0], np.arange(1, 13)) and np.all(B[:6, 2] == A[:6, 1])
    assert B[0, 0] == 4.0 and B[-1, 0] == 1e4 and np.float32(0.005) == A[0, 1] and np.float32(0.05) == A[5, 1]
    # device values == the same quadrature compiled for the host (float32 storage)
    s_nodes = 10 ** np.linspace(np.log10(4.0), 4.0, 24)
    d_nodes = np.linspace(0.005, 0.05, 6)
    t_nodes = 10 ** np.linspace(np.log10(4.0), 4.0, 40)
    for i0 in (1, 5, 11, 23):
        for i1 in (0, 1, 4, 11):
            for i2 in (0, 3, 5):
                ref = hcpp.hc_pp_alpha_node(s_nodes[i0], float(i1 + 1), d_nodes[i2])
                assert B[(i0 * 12 + i1) * 6 + i2, 3] == pytest.approx(ref, rel=3e-7, abs=1e-44)
    for i0 in (0, 7, 20, 39):
        for i1 in (0, 5):
            assert A[i0 * 6 + i1, 2] == pytest.approx(hcpp.hc_pp_alphatilde_node(t_nodes[i0], d_nodes[i1]), rel=3e-7, abs=1e-44)
    # the product loads tables of any size in this format
    ctx = nusi.Context(N_bins_E=60, lEmin=4.0, lEmax=9.0, phiphi=True, alphatilde_phiphi_path=at, alpha_phiphi_path=al)
    ctx.close()


@pytest.mark.gpu
def test_evolve_with_phiphi_vs_oracle_on_the_same_tables(nusi, oracle, tmp_path):
    """Full-size tables (the only size the reference constructor loads, nuSIprop.hpp:168-169) generated on
    the GPU, then evolve() with phiphi = true on both sides."""
    xs = tmp_path / "xsec"
    xs.mkdir()
    at, al = str(xs / "alphatilde_phiphi.bin"), str(xs / "alpha_phiphi.bin")
    sec = nusi.generate_phiphi_tables(at, al)
    assert os.path.getsize(at) == 5000 * 100 * 3 * 4 and os.path.getsize(al) == 1000 * 1000 * 100 * 4 * 4
    print("generator kernels: %.2f s for 1.005e8 nodes" % sec)
    cfg = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    pts = [(100.0, 1e-2, 0.1), (30.0, 1e-1, 0.1), (300.0, 1.0, 0.1), (1e3, 1e-2, 0.1)]
    cwd = os.getcwd()
    os.chdir(str(tmp_path))  # the reference opens "xsec/*.bin" relative to the CWD
    try:
        ref = [oracle.evolve(m, g, mn, phiphi=True, **cfg) for (m, g, mn) in pts[:2]]
        ref_off = oracle.evolve(*pts[0], phiphi=False, **cfg)
        ref_d = oracle.evolve(*pts[0], phiphi=True, **dict(cfg, majorana=False))
    finally:
        os.chdir(cwd)
    ctx = nusi.Context(phiphi=True, alphatilde_phiphi_path=at, alpha_phiphi_path=al, **cfg)
    fm, ff, st = ctx.evolve([p[0] for p in pts], [p[1] for p in pts], [p[2] for p in pts])
    ctx.close()
    assert np.all(st & 8 == 0) and np.isfinite(fm).all()
    for i in range(2):
        assert oracle.relerr(fm[i], ref[i][0]).max() <= 1e-9, (pts[i], oracle.relerr(fm[i], ref[i][0]).max())
        assert oracle.relerr(ff[i], ref[i][1]).max() <= 1e-9
    # the channel is really on: it changes the flux well beyond the tolerance
    assert oracle.relerr(ref[0][0], ref_off[0]).max() > 1e-6
    ctx = nusi.Context(phiphi=True, alphatilde_phiphi_path=at, alpha_phiphi_path=al, **dict(cfg, majorana=False))
    fd = ctx.evolve(*pts[0])[0]
    ctx.close()
    assert oracle.relerr(fd[0], ref_d[0]).max() <= 1e-9
