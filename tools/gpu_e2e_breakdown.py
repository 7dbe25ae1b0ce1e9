#!/usr/bin/env python
"""Where the end-to-end path (nusi_evolve_batch, host buffers) spends the time it takes beyond the device-resident path:
host preparation of the points, the pipeline itself, the exposed device->host tail.  One rank per GPU under torchrun
(all ranks run concurrently, as in bench.py's e2e leg), or a single process.
    python tools/gpu_e2e_breakdown.py            |   torchrun --nproc-per-node 8 tools/gpu_e2e_breakdown.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import nusiprop_b200 as nb  # noqa: E402
import bench  # noqa: E402
from nusiprop_b200.api import _dp, _ip  # noqa: E402

world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.cuda.set_device(local)
ctx = nb.Context(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2, device=local)
mphi, g = bench.shard(rank % 8)
n, N = mphi.size, 200
mntot = np.full(n, 0.1)
lib = ctx._lib
h_fm = torch.empty((n, 3, N), dtype=torch.float64, pin_memory=True).numpy()
h_ff = torch.empty((n, 3, N), dtype=torch.float64, pin_memory=True).numpy()
h_st = np.zeros(n, dtype=np.int32)
d_fm = torch.empty((n, 3, N), dtype=torch.float64, device="cuda")
d_ff = torch.empty_like(d_fm)
d_st = torch.zeros(n, dtype=torch.int32, device="cuda")


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def timed(fn, reps=5):
    fn()
    sync()
    t = []
    for _ in range(reps):
        sync()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        t.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(t))


def create_destroy():
    p = ctx.upload_points(mphi, g, mntot)
    p.close()


pts = ctx.upload_points(mphi, g, mntot)
stream = torch.cuda.current_stream().cuda_stream
res = {
    "points_create+destroy (host prep + H2D)": timed(create_destroy),
    "evolve_points (device-resident step)": timed(lambda: ctx.evolve_points(pts, d_fm.data_ptr(), d_ff.data_ptr(), d_st.data_ptr(), stream)),
    "evolve_points + one D2H of everything": timed(lambda: (ctx.evolve_points(pts, d_fm.data_ptr(), d_ff.data_ptr(), d_st.data_ptr(), stream),
                                                             torch.from_numpy(h_fm).copy_(d_fm, non_blocking=True), torch.from_numpy(h_ff).copy_(d_ff, non_blocking=True))),
    "D2H of everything alone": timed(lambda: (torch.from_numpy(h_fm).copy_(d_fm, non_blocking=True), torch.from_numpy(h_ff).copy_(d_ff, non_blocking=True))),
    "nusi_evolve_batch, both fluxes (e2e)": timed(lambda: lib.nusi_evolve_batch(ctx._h, n, _dp(mphi), _dp(g), _dp(mntot), _dp(h_fm), _dp(h_ff), _ip(h_st))),
    "nusi_evolve_batch, flavour fluxes only": timed(lambda: lib.nusi_evolve_batch(ctx._h, n, _dp(mphi), _dp(g), _dp(mntot), None, _dp(h_ff), _ip(h_st))),
}
c1 = nb.Context(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=200, lEmin=4.0, lEmax=9.0, zmax=5.0, flav=2, device=local, host_copy=1)
res["nusi_evolve_batch, both fluxes, host_copy=1 (copies after the last kernel)"] = timed(
    lambda: lib.nusi_evolve_batch(c1._h, n, _dp(mphi), _dp(g), _dp(mntot), _dp(h_fm), _dp(h_ff), _ip(h_st)))
c1.close()
for r in range(world):
    if r == rank:
        print("rank %d/%d: " % (rank, world) + "; ".join("%s %.2f ms" % kv for kv in res.items()), flush=True)
    if world > 1:
        dist.barrier()
pts.close()
ctx.close()
if world > 1:
    dist.destroy_process_group()
