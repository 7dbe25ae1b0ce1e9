"""C-ABI surface and host logic that need no GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol(nusi):
    from nusiprop_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "nusiprop_b200.h")).read()
    declared = set(re.findall(r"\b(nusi_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = C.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), "symbol %s declared in include/nusiprop_b200.h but not exported" % name
    assert declared == set(_lib.SIGNATURES), "ctypes table and header disagree: %s" % (declared ^ set(_lib.SIGNATURES))
    assert _lib.load().nusi_abi_version() == 4


def test_opts_defaults_match_reference_constructor(nusi):
    from nusiprop_b200 import _lib
    o = _lib.NusiOpts()
    assert _lib.load().nusi_opts_default(C.byref(o)) == 0
    # reference nuSIprop.hpp:61-65
    assert (o.majorana, o.non_resonant, o.normal_ordering, o.N_bins_E, o.lEmin, o.lEmax, o.zmax, o.flav, o.phiphi) == (1, 1, 1, 300, 12.0, 17.0, 5.0, 2, 0)


@pytest.mark.parametrize("mntot,no", [(0.1, True), (0.1, False), (0.06, True), (0.3, False), (1.0, True)])
def test_masses_match_oracle(nusi, oracle, mntot, no):
    mn = nusi.masses(mntot, no)
    ref = oracle.masses(mntot, no)
    assert np.array_equal(mn, ref), (mn, ref)  # same quartic solve, same acceptance test (aux.hpp:38-46)
    assert abs(mn.sum() - mntot) < 1e-12


def test_masses_reject_unphysical_sum(nusi):
    with pytest.raises(ValueError):
        nusi.masses(0.01, True)  # below sqrt(dm21)+sqrt(dm31): the reference prints and exit(1)s (aux.hpp:48-49)


def test_testpy_mass_sum_is_on_the_knife_edge(nusi, oracle):
    mntot = float(0.0 + np.sqrt(7.42e-5) + np.sqrt(2.514e-3))
    assert np.array_equal(nusi.masses(mntot, True), oracle.masses(mntot, True))


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(nusi):
    """Without a CUDA device the product refuses to run; it never routes through a CPU path."""
    with pytest.raises(nusi.NusiError, match="no usable CUDA device"):
        nusi.Context(N_bins_E=50)
    with pytest.raises(nusi.NusiError):
        nusi.pyprop(100.0, 1e-2, 0.1, 2.0, phiphi=False)
    with pytest.raises(nusi.NusiError):
        nusi.specfun(0, np.array([0.5]))


def test_strerror(nusi):
    from nusiprop_b200 import _lib
    L = _lib.load()
    assert L.nusi_strerror(0) == b"ok"
    assert b"CPU" in L.nusi_strerror(-2)


def test_freestream_diagnostics_match_reference_integrals(oracle):
    """flux_FS_E0 / energy_FS (nuSIprop.hpp:666-744) are kept on the host; compare through
    check_energy_conservation of a free-streaming configuration (g -> 0, resonant-only)."""
    from nusiprop_b200 import freestream

    class Ctx(object):
        zmax = None
        opts = dict(lEmin=4.0, lEmax=9.0)
    nz, Emin, Emax, Ec, z = oracle.grid(50, 4.0, 9.0, 5.0)
    Ctx.zmax = float(z[-1])
    E_FS = freestream.energy_FS(Ctx, 2.0, 6.0)
    fm, ff, en = oracle.evolve(1e2, 1e-8, 0.1, si=2.0, norm=6.0, non_resonant=False, N_bins_E=50, lEmin=4.0, lEmax=9.0)
    E_int = sum((np.log(Emax[i]) - np.log(Emin[i])) * Ec[i] ** 2 * fm[k][i] for i in range(50) for k in range(3))
    ref = oracle.lib().nusi_ref_check_energy_conservation(1e2, 1e-8, 0.1, 2.0, 6.0, 1, 0, 1, 50, 4.0, 9.0, 5.0, 2)
    assert (E_int - E_FS) / E_FS == pytest.approx(ref, rel=1e-10)
