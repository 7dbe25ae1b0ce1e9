#!/usr/bin/env python
"""Time the kernels of every build under nusiprop_b200/lib/variants/ (tools/build_variant.sh) on the benchmark shard and
compare its fluxes with the product library's bit for bit."""
import glob
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nusiprop_b200 as nb  # noqa: E402
from nusiprop_b200 import _lib  # noqa: E402
import bench  # noqa: E402

m, g = bench.shard(0)
cfg = dict(N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2)
ref = None
for path in [None] + sorted(glob.glob(os.path.join(ROOT, "nusiprop_b200", "lib", "variants", "libnusiprop_b200_*.so"))):
    ctx = nb.Context(_library=_lib.load_path(path) if path else None, **cfg)
    for _ in range(3):
        out = ctx.evolve(m, g, 0.1)
        tm = ctx.last_timing()
    ctx.close()
    if ref is None:
        ref = out
    print("%-40s rows %.3f alpha %.3f integrate %.3f ms; bitwise equal to the product: %s" % (
        os.path.basename(path) if path else "product", tm["ms_rows"], tm["ms_alpha"], tm["ms_integrate"], np.array_equal(out[0], ref[0])), flush=True)
