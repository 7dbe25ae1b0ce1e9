"""Integrator configurations for fine binning (NUSI_FINE_VARIANT=0: 256 threads per point throughout; default: 512
threads for N_E > 256): time and bitwise comparison with the bin-owned wave kernel.  usage: gpu_fine_variants.py n N_E"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nusiprop_b200 as nb
NE = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
ctx = nb.Context(N_bins_E=NE, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
mphi = 10 ** np.linspace(1.5, 3.0, n); g = 10 ** np.linspace(-3, -1, n)
os.environ["NUSI_WAVE_INTEGRATOR"] = "1"
ref = ctx.evolve(mphi, g, 0.1)[1]
del os.environ["NUSI_WAVE_INTEGRATOR"]
for v in (0, 1):
    os.environ["NUSI_FINE_VARIANT"] = str(v)
    for rep in range(2):
        fm, ff, st = ctx.evolve(mphi, g, 0.1)
        tm = ctx.last_timing()
    same = np.array_equal(ff, ref, equal_nan=True)
    rel = np.nanmax(np.abs(ff - ref) / np.maximum(np.abs(ref), 1e-300))
    print("N_E=%d n=%d variant %d: integrate %.1f ms, us/z-step/point %.2f, bitwise == wave kernel: %s (max rel %.1e)" % (
        NE, n, v, tm["ms_integrate"], 1e3 * tm["ms_integrate"] / n / (ctx.N_steps_z - 1), same, rel))
