"""nusi_energy_integral_static -- the entry point that allocates nothing on the device (include/nusiprop_b200.h; no
counterpart in the reference) -- against nusi_energy_integral, the existing
entry point, through the C ABI.  Same reduction, so the tolerance is the bitwise one of the other device-vs-device tests.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

CFG = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


@pytest.mark.gpu
def test_energy_integral_static_equals_energy_integral_on_two_calls_in_a_row(nusi):
    import torch
    from nusiprop_b200 import _lib
    lib = _lib.load()
    ctx = nusi.Context(phiphi=False, **CFG)
    N = ctx.N_bins_E
    n1, n2 = 96, 37
    rng = np.random.default_rng(7)
    mphi, g = 10 ** rng.uniform(1, 3, n1), 10 ** rng.uniform(-5, -1, n1)
    pts = ctx.upload_points(mphi, g, 0.1)
    fm = torch.empty((n1, 3, N), dtype=torch.float64, device="cuda:0")
    ctx.evolve_points(pts, fm.data_ptr(), 0, 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(fm).all()) and float(fm.max()) > 0

    # existing entry point (grows a cudaMalloc'ed staging buffer inside the context)
    want1, want2 = np.zeros(n1), np.zeros(n2)
    assert lib.nusi_energy_integral(ctx._h, n1, C.c_void_p(fm.data_ptr()), _dp(want1), None) == 0
    second = fm[n1 - n2:].contiguous()  # the second call: fewer, different points -- stale slots of the first must not show
    assert lib.nusi_energy_integral(ctx._h, n2, C.c_void_p(second.data_ptr()), _dp(want2), None) == 0
    assert np.all(want1 > 0) and not np.array_equal(want1[:n2], want2)

    # the new entry point, two calls in a row into caller-owned host arrays; device memory in use must not move
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    got1, got2 = np.full(n1, -1.0), np.full(n2, -1.0)
    assert lib.nusi_energy_integral_static(ctx._h, n1, C.c_void_p(fm.data_ptr()), _dp(got1), None) == 0
    assert lib.nusi_energy_integral_static(ctx._h, n2, C.c_void_p(second.data_ptr()), _dp(got2), None) == 0
    free1 = torch.cuda.mem_get_info(0)[0]
    assert np.array_equal(got1, want1), np.abs(got1 / want1 - 1).max()
    assert np.array_equal(got2, want2), np.abs(got2 / want2 - 1).max()
    assert free1 == free0, "device memory in use changed across the no-allocation calls: %d -> %d" % (free0, free1)
    # and against the reference's serial loop (nuSIprop.hpp:351-355) to summation-order rounding, like tests/test_powerlaw_source.py
    Emin, Ec, Emax = ctx.energies()
    host = ((np.log(Emax) - np.log(Emin)) * Ec ** 2 * fm.cpu().numpy().sum(axis=1)).sum(axis=1)
    assert np.allclose(got1, host, rtol=1e-13)
    # the Python wrapper is the same call
    assert np.array_equal(ctx.energy_integral_static(fm.data_ptr(), n1), want1)

    # sizes beyond the stated maximum are rejected with a clear error, and nothing is written
    limit = 65536
    big = np.full(4, -1.0)
    rc = lib.nusi_energy_integral_static(ctx._h, limit + 1, C.c_void_p(fm.data_ptr()), _dp(big), None)
    assert rc == -1 and np.all(big == -1.0)
    msg = lib.nusi_last_error().decode()
    assert "NUSI_STATIC_MAX_POINTS" in msg and str(limit) in msg, msg
    assert lib.nusi_energy_integral_static(ctx._h, 0, C.c_void_p(fm.data_ptr()), _dp(big), None) == -1
    pts.close()
    ctx.close()
