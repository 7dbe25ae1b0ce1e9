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
import torch  # noqa: E402


def step_ms(lib, **kw):
    """CUDA-event time of one device-resident step of the shard (kernels of different streams may overlap)."""
    ctx = nb.Context(_library=lib, **cfg, **kw)
    pts = ctx.upload_points(m, g, np.full(m.size, 0.1))
    fm = torch.empty((m.size, 3, 200), dtype=torch.float64, device="cuda")
    ff, st = torch.empty_like(fm), torch.zeros(m.size, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        ctx.evolve_points(pts, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(3):
        ctx.evolve_points(pts, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), s)
    e1.record()
    torch.cuda.synchronize()
    out = fm.cpu().numpy()
    pts.close()
    ctx.close()
    return e0.elapsed_time(e1) / 3, out


ref = None
for path in [None] + sorted(glob.glob(os.path.join(ROOT, "nusiprop_b200", "lib", "variants", "libnusiprop_b200_*.so"))):
    ctx = nb.Context(_library=_lib.load_path(path) if path else None, **cfg)
    for _ in range(3):
        out = ctx.evolve(m, g, 0.1)
        tm = ctx.last_timing()
    ctx.close()
    if ref is None:
        ref = out
    lib_ = _lib.load_path(path) if path else None
    t1, o1 = step_ms(lib_)
    t2, o2 = step_ms(lib_, two_stream=1)
    print("%-40s step %.3f ms; builder and rows+integrator on two streams %.3f ms (bitwise equal %s)" % ("", t1, t2, np.array_equal(o1, o2) and np.array_equal(o1, ref[0])))
    print("%-40s rows %.3f alpha %.3f integrate %.3f ms; bitwise equal to the product: %s" % (
        os.path.basename(path) if path else "product", tm["ms_rows"], tm["ms_alpha"], tm["ms_integrate"], np.array_equal(out[0], ref[0])), flush=True)
