"""BASELINE configs[4] under a parity gate: the 256 x 256 (mphi, g) scan at N_E = 200, evolved shard by shard in
scan order with the DEFAULT builder (elementary_mode 3: alpha_group_kernel / rows_group_kernel on groups of 32), against
the oracle on the seeded 1024-point subsample (reference path nuSIprop.hpp:176-337).  Criterion and tiers:
tests/scan_parity.py."""
import numpy as np
import pytest

import scan_parity as SP

CFG200 = dict(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)


def test_fixture_is_the_oracle(oracle):
    """The committed fixture is what the oracle (and its ensemble) produce here, now: a few points live (not gpu)."""
    fx = SP.load_fixture()
    assert fx["mass"].shape == (1024, 3, 200) and fx["floor"].shape == (1024, 200) and len(set(fx["index"].tolist())) == 1024
    assert np.array_equal(fx["index"], np.random.default_rng(20261104).choice(65536, 1024, replace=False))
    for p in (0, 511, 1023):
        fm = oracle.evolve(float(fx["mphi"][p]), float(fx["g"][p]), 0.1, **CFG200)[0]
        assert np.array_equal(fm, fx["mass"][p])
        dev = np.zeros(200)
        for v in fx["members"]:
            dev = np.maximum(dev, oracle.relerr(oracle.evolve(float(fx["mphi"][p]), float(fx["g"][p]), 0.1, variant=str(v), **CFG200)[0], fm).max(axis=0))
        assert np.array_equal(dev.astype(np.float32), fx["floor"][p])
    # the corners of the scan the advisor asked about are inside the sample's range
    assert fx["mphi"].min() < 12 and fx["mphi"].max() > 9e3 and fx["g"].max() > 0.9 and fx["g"].min() < 1.1e-5


def test_criterion_rejects_a_wrong_result():
    """The gate is not vacuous: a relative perturbation of 1e-8 in well-conditioned bins fails it."""
    fx = SP.load_fixture()
    ok = SP.evaluate(fx["mass"], fx)
    assert ok["n_failed"] == 0 and ok["strict_frac"] == 1.0
    bad = fx["mass"].copy()
    p, b = np.unravel_index(np.argmin(fx["floor"] + (fx["mass"][:, 0, :] == 0)), fx["floor"].shape)
    bad[p, :, b] *= 1 + 1e-8
    rep = SP.evaluate(bad, fx)
    assert rep["n_failed"] == 1 and rep["worst_over_tol"] > 5


@pytest.mark.gpu
def test_headline_scan_matches_the_oracle_in_scan_order_with_the_default_builder(nusi):
    fx = SP.load_fixture()
    ctx = nusi.Context(**CFG200)
    got, st, info = SP.evolve_fixture_points_in_their_shards(ctx, fx)
    ctx.close()
    # the grouped scan kernels really produced these numbers: every shard is 256 groups of 32 points, in the caller's order
    assert info == dict(n_groups=256, n_grouped_points=8192, reordered=False), info
    assert np.all(st & ~7 == 0)
    rep = SP.evaluate(got, fx)
    print("scan parity:", rep)
    # every bin within max(1e-9, 4 x the reference's own spread in THAT bin) -- no widening over neighbouring bins
    assert rep["n_failed"] == 0, rep
    # and within 1e-9 at least as often as a rebuild of the reference agrees with the reference
    assert rep["strict_frac"] >= rep["ref_strict_frac"], rep
    assert rep["strict_frac"] > 0.92


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [0, 2])
def test_headline_sample_other_builders(nusi, mode):
    """The strict builder (mode 0) and the compact builder (mode 2, = mode 3 bit for bit) on the same sample."""
    fx = SP.load_fixture()
    ctx = nusi.Context(elementary_mode=mode, **CFG200)
    fm, ff, st = ctx.evolve(fx["mphi"], fx["g"], 0.1)
    ctx.close()
    rep = SP.evaluate(fm, fx)
    print("mode", mode, rep)
    assert rep["n_failed"] == 0 and rep["strict_frac"] >= rep["ref_strict_frac"], rep


@pytest.mark.gpu
def test_scan_runner_is_the_product_path_and_equals_context_evolve_bitwise(nusi):
    """nusiprop_b200.scan.ScanRunner -- what bench.py --gpus N times -- at world size 1: device-resident evolve in
    2048-point pieces, same bits as one Context.evolve() of the whole shard; evolve_scan_distributed() is built on it."""
    import bench
    from nusiprop_b200.scan import ScanRunner, evolve_scan_distributed
    mphi, g = bench.shard(3)
    mphi, g = mphi[:5000], g[:5000]  # ragged last piece
    ctx = nusi.Context(**CFG200)
    ref = ctx.evolve(mphi, g, 0.1)
    run = ScanRunner(ctx, mphi, g, 0.1)
    for _ in range(2):  # a second step reuses the uploaded points and the output tensors
        fm, ff, st = run.step().result()
        assert np.array_equal(fm, ref[0]) and np.array_equal(ff, ref[1]) and np.array_equal(st, ref[2])
    run.close()
    ctx.close()
    fm, ff, st = evolve_scan_distributed(mphi[:300], g[:300], 0.1, ctx_kwargs=CFG200)
    assert np.array_equal(fm, ref[0][:300]) and np.array_equal(ff, ref[1][:300])


@pytest.mark.gpu
def test_fast_builder_against_the_strict_builder_on_a_full_shard(nusi):
    """Breadth beyond the 1024-point fixture: elementary_mode 0 is the reference's arithmetic (bit-identical tables and
    fluxes within 7e-15 of the libm-locked oracle, tests/test_locked_parity.py), so the default builder can be measured
    against it on EVERY point of a shard of the headline scan.  The tiers must look like the fixture's: well over 90 % of the
    bins within 1e-9, nothing off by more than a few per cent, and the large deviations only where the reference is
    ill-conditioned (m_phi >~ 1e3 eV with g >~ 1e-2: there two builds of the reference differ by up to 2.5e-2)."""
    import bench
    mphi, g = bench.shard(5)
    ctx = nusi.Context(**CFG200)
    fast = ctx.evolve(mphi, g, 0.1)
    ctx.close()
    ctx = nusi.Context(elementary_mode=0, **CFG200)
    strict = ctx.evolve(mphi, g, 0.1)
    ctx.close()
    assert np.array_equal(fast[2] & ~7, strict[2] & ~7) and np.isfinite(fast[0]).all() and np.isfinite(strict[0]).all()
    r = SP.relerr(fast[0], strict[0]).max(axis=(1, 2))  # per point
    rb = SP.relerr(fast[0], strict[0]).max(axis=1)      # per (point, bin)
    for lo, hi in ((10, 1e2), (1e2, 1e3), (1e3, 1e4 + 1)):
        m = (mphi >= lo) & (mphi < hi)
        print("mode 3 vs mode 0, %g <= m_phi < %g eV: %d points, bins within 1e-9 %.4f, worst %.2e" % (lo, hi, m.sum(), np.mean(rb[m] <= 1e-9), r[m].max()))
    print("all 8192 points: bins within 1e-9 %.4f, worst %.2e at (%.1f, %.2e)" % (np.mean(rb <= 1e-9), r.max(), mphi[np.argmax(r)], g[np.argmax(r)]))
    assert np.mean(rb <= 1e-9) > 0.92
    assert r.max() < 0.05
    assert r[mphi < 1e2].max() < 1e-9  # the well-conditioned third of the scan: north_star's tolerance with no floor term
