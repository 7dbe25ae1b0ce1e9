"""The reference-facing surfaces above the C ABI: the drop-in C++ class nuSIprop::calculate_flux
(nusiprop_b200/cxx/nuSIprop.hpp, driven through tests/dropin/consumer.cpp, which uses only the
reference's interface -- cf. reference test.cpp:4-31) and the Cython class nuSIprop.pyprop
(nusiprop_b200/cython, cf. reference nuSIprop.pyx)."""
import json
import os
import subprocess
import sys
import warnings

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(HERE, "golden")
CASE = "single_m100_g0.01_maj1_nr1_no1"


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="module")
def consumer(nusi):
    from nusiprop_b200.build import build_dropin
    return build_dropin()


@pytest.fixture(scope="module")
def cymod(consumer):
    sys.path.insert(0, os.path.join(ROOT, "nusiprop_b200", "cython"))
    import nuSIprop
    return nuSIprop


def _args(c, mode=None, si=2.0, norm=1.0, phiphi=0):
    a = [repr(float(c["mphi"])), repr(float(c["g"])), repr(float(c["mntot"])), repr(si), repr(norm), str(int(c["majorana"])),
         str(int(c["non_resonant"])), str(int(c["normal_ordering"])), str(c["N_bins_E"]), repr(float(c["lEmin"])),
         repr(float(c["lEmax"])), repr(float(c["zmax"])), str(c["flav"]), str(phiphi)]
    return a + ([mode] if mode else [])


def _case():
    return json.load(open(os.path.join(GOLD, "oracle_flux_cases.json")))[CASE]


# ---- no GPU needed ---------------------------------------------------------------------------
def test_dropin_grid_and_getter_messages(consumer, oracle):
    c = _case()
    r = subprocess.run([consumer] + _args(c, "grid"), capture_output=True, text=True)
    assert r.returncode == 0
    lines = r.stdout.split("\n")
    en = np.array([float(x) for x in lines[: c["N_bins_E"]]])
    _, _, _, Ec, _ = oracle.grid(c["N_bins_E"], c["lEmin"], c["lEmax"], c["zmax"])
    assert np.array_equal(en, Ec)  # get_energy() before any evolve(), no GPU involved
    assert lines[c["N_bins_E"]] == "oob 0 0"
    # the reference's stderr texts (nuSIprop.hpp:364-378, :414-426)
    assert "You asked for the energy at the bin -1<0! Zero will be returned." in r.stderr
    assert "You asked for the flux of the mass eigenstate 3, not in [0,1,2]. Zero will be returned." in r.stderr


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_dropin_has_no_cpu_path(consumer, cymod):
    r = subprocess.run([consumer] + _args(_case()), capture_output=True, text=True)
    assert r.returncode == 1 and "no usable CUDA device" in r.stderr and r.stdout == ""
    p = cymod.pyprop(100.0, 1e-2, 0.1, 2.0, phiphi=False, N_bins_E=50, lEmin=4.0, lEmax=9.0)
    with pytest.raises(RuntimeError, match="no usable CUDA device"):
        p.evolve()


def test_cython_wrapper_surface(cymod, oracle):
    """Constructor keywords / defaults and pre-evolve behaviour of reference nuSIprop.pyx:47-58, :92-118."""
    p = cymod.pyprop(mphi=100.0, g=1e-2, mntot=0.1, si=2.0, norm=1, majorana=True, non_resonant=True, normal_ordering=True,
                     N_bins_E=60, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2, phiphi=False)
    _, _, _, Ec, _ = oracle.grid(60, 4.0, 9.0, 5.0)
    assert np.array_equal(p.get_energies(), Ec)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        assert p.get_flux().shape == (3, 60) and np.all(p.get_flux() == 0) and np.all(p.get_flux_fla() == 0)
        assert len(w) == 3 and "You have not evolved the neutrino flux! Zero flux will be returned." in str(w[0].message)
    p.set_parameters(g=1e-3, mntot=0.2)
    for name in ("evolve", "get_flux", "get_flux_fla", "interp_flux_el", "interp_flux_mu", "interp_flux_ta", "get_energies",
                 "check_energy_conservation", "set_parameters"):
        assert callable(getattr(p, name))
    # phiphi defaults to True in the Python wrapper (pyx:52): the tables are looked up in the constructor
    with pytest.raises(RuntimeError):
        cymod.pyprop(100.0, 1e-2, 0.1, 2.0, N_bins_E=60, lEmin=4.0, lEmax=9.0)


def test_freestream_energy_host_function(nusi, oracle):
    import ctypes as C
    from nusiprop_b200 import _lib
    lib = _lib.load()
    nz, Emin, Emax, Ec, z = oracle.grid(60, 4.0, 9.0, 5.0)
    from nusiprop_b200 import freestream

    class _Ctx(object):
        zmax = z[-1]
        opts = dict(lEmin=4.0, lEmax=9.0)
    for si, norm in ((2.0, 6.0), (2.5, 1.0)):
        e = C.c_double()
        assert lib.nusi_freestream_energy(z[-1], 4.0, 9.0, si, norm, C.byref(e)) == 0
        assert e.value == pytest.approx(freestream.energy_FS(_Ctx, si, norm), rel=1e-13)


# ---- on the B200 -----------------------------------------------------------------------------
@pytest.mark.gpu
def test_dropin_class_evolve_vs_oracle(consumer):
    c = _case()
    z = np.load(os.path.join(GOLD, "oracle_flux.npz"))
    r = subprocess.run([consumer] + _args(c), capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = np.array([[float(x) for x in ln.split()] for ln in r.stdout.strip().split("\n")])
    assert out.shape == (c["N_bins_E"], 7)
    import oracle_ref
    assert oracle_ref.relerr(out[:, 1:4].T, z[CASE + "/mass"]).max() <= 1e-9
    assert oracle_ref.relerr(out[:, 4:7].T, z[CASE + "/fla"]).max() <= 1e-9


@pytest.mark.gpu
def test_dropin_class_copy_semantics_and_member_updates(consumer):
    r = subprocess.run([consumer] + _args(_case(), "copy"), capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "copy_equal 1 original_changed 1", (r.stdout, r.stderr)


@pytest.mark.gpu
def test_dropin_class_fatal_paths(consumer):
    c = dict(_case(), mntot=0.01)
    r = subprocess.run([consumer] + _args(c), capture_output=True, text=True)
    assert r.returncode == 1 and "No neutrino mass spectrum was found" in r.stderr
    r = subprocess.run([consumer] + _args(_case(), phiphi=1), capture_output=True, text=True, cwd="/tmp")
    assert r.returncode == 1 and "table" in r.stderr  # missing xsec/*.bin relative to CWD, as the reference


@pytest.mark.gpu
def test_dropin_check_energy_conservation(consumer, oracle):
    c = dict(_case(), N_bins_E=60)
    r = subprocess.run([consumer] + _args(c, "conservation", si=2.0, norm=6.0), capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ref = oracle.lib().nusi_ref_check_energy_conservation(100.0, 1e-2, 0.1, 2.0, 6.0, 1, 1, 1, 60, 4.0, 9.0, 5.0, 2)
    assert float(r.stdout) == pytest.approx(ref, rel=1e-8)


@pytest.mark.gpu
def test_cython_pyprop_and_scan(cymod, nusi, oracle):
    cfg = dict(N_bins_E=60, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
    p = cymod.pyprop(100.0, 1e-2, 0.1, 2.0, norm=6, phiphi=False, **cfg)
    p.evolve()
    rm, rf, en = oracle.evolve(100.0, 1e-2, 0.1, **cfg)
    assert oracle.relerr(p.get_flux(), rm).max() < 1e-9 and oracle.relerr(p.get_flux_fla(), rf).max() < 1e-9
    e = np.array([2e4, 3e6])
    assert np.all(p.interp_flux_mu(e) > 0)
    p.set_parameters(g=0.1)
    p.evolve()
    rm2, _, _ = oracle.evolve(100.0, 0.1, 0.1, **cfg)
    assert oracle.relerr(p.get_flux(), rm2).max() < 1e-9
    # batch entry point == the ctypes path, bit for bit
    mphi, g = 10 ** np.linspace(1.5, 3, 7), 10 ** np.linspace(-4, -1, 7)
    fm, ff, st = cymod.evolve_scan(mphi, g, 0.1, **cfg)
    ctx = nusi.Context(**cfg)
    fm2, ff2, st2 = ctx.evolve(mphi, g, 0.1)
    ctx.close()
    assert np.array_equal(fm, fm2) and np.array_equal(ff, ff2) and np.array_equal(st, st2)
    with pytest.raises(RuntimeError, match="No neutrino mass spectrum"):
        q = cymod.pyprop(100.0, 1e-2, 0.01, 2.0, phiphi=False, **cfg)
        q.evolve()


@pytest.mark.gpu
def test_dropin_power_law_source_closes_energy_conservation(consumer, cymod):
    """source_model = 1 (extra trailing constructor argument): the upstream power-law source (reference hpp:648-657);
    check_energy_conservation() of the C++ class and of the Cython wrapper then measures the discretisation error."""
    c = dict(_case(), N_bins_E=200, lEmin=12.0, lEmax=17.0)
    args = _args(c, "conservation", si=2.5, norm=3.0)
    args[0:2] = ["5e6", "1e-9"]  # mphi, g: free streaming
    r = subprocess.run([consumer] + args + ["1"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    cons = float(r.stdout)
    assert -0.03 < cons < 0
    p = cymod.pyprop(5e6, 1e-9, 0.1, 2.5, 3.0, N_bins_E=200, lEmin=12.0, lEmax=17.0, zmax=5.0, phiphi=False, source_model=1)
    assert p.check_energy_conservation() == pytest.approx(cons, rel=1e-12)
    en = p.get_energies()
    tot = p.get_flux().sum(axis=0)
    assert np.allclose(tot / (3.0 * (en / 1e14) ** -2.5), tot[0] / (3.0 * (en[0] / 1e14) ** -2.5), rtol=1e-9)
