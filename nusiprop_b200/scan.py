"""Parameter scans sharded statically over the GPUs of one box.

Each (mphi, g, mntot) point is an independent evolve() (reference: one calculate_flux object per
point), so the only exchange is the final gather of the results.  Point p of the scan belongs to
rank p mod world_size (strided: the cost of a point varies smoothly over the grid).  One process
per GPU, `torch.distributed` for the gather (NCCL on GPUs; the host-side logic is backend agnostic
and is tested with gloo on CPU).
"""
import numpy as np


def shard_indices(n, rank, world):
    """Indices of the scan points owned by `rank`."""
    return np.arange(rank, n, world)


def unshard(parts, n, world):
    """Inverse of the strided sharding: parts[r] holds the rows of rank r, returns [n, ...]."""
    first = np.asarray(parts[0])
    out = np.empty((n,) + first.shape[1:], dtype=first.dtype)
    for r in range(world):
        out[shard_indices(n, r, world)] = np.asarray(parts[r])
    return out


def evolve_scan_distributed(mphi, g, mntot, ctx_kwargs=None, evolve_fn=None, device=None):
    """Run a scan across the ranks of the default process group (or alone when it is not
    initialised).  Every rank passes the FULL point arrays; rank 0 returns
    (flux_mass[n,3,N], flux_fla[n,3,N], status[n]), other ranks return None.

    evolve_fn(mphi, g, mntot) -> (flux_mass, flux_fla, status) replaces the GPU context in the CPU
    tests of the sharding logic; by default a nusiprop_b200.Context on `device` (LOCAL_RANK) is used.
    """
    import torch
    import torch.distributed as dist
    mphi, g, mntot = (np.ascontiguousarray(np.broadcast_to(np.asarray(a, dtype=np.float64), np.broadcast(mphi, g, mntot).shape)).ravel()
                      for a in (mphi, g, mntot))
    n = mphi.size
    on = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank() if on else 0
    world = dist.get_world_size() if on else 1
    idx = shard_indices(n, rank, world)
    ctx = None
    if evolve_fn is None:
        import os
        from .api import Context
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        ctx = Context(device=device, **(ctx_kwargs or {}))
        evolve_fn = ctx.evolve
    fm, ff, st = evolve_fn(mphi[idx], g[idx], mntot[idx])
    if ctx is not None:
        ctx.close()
    if world == 1:
        return fm, ff, st
    use_cuda = dist.get_backend() == "nccl"
    dev = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    # shards differ in length by at most one row: pad to the longest, gather, trim
    nmax = (n + world - 1) // world
    N = fm.shape[-1]

    def pad(a, shape, dtype):
        t = torch.zeros(shape, dtype=dtype, device=dev)
        t[: a.shape[0]] = torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(dev)
        return t
    tf = pad(np.concatenate([fm, ff], axis=1), (nmax, 6, N), torch.float64)
    ts = pad(st, (nmax,), torch.int32)
    gl_f = [torch.empty_like(tf) for _ in range(world)] if rank == 0 else None
    gl_s = [torch.empty_like(ts) for _ in range(world)] if rank == 0 else None
    dist.gather(tf, gl_f, dst=0)
    dist.gather(ts, gl_s, dst=0)
    if rank != 0:
        return None
    pf = [gl_f[r][: len(shard_indices(n, r, world))].cpu().numpy() for r in range(world)]
    ps = [gl_s[r][: len(shard_indices(n, r, world))].cpu().numpy() for r in range(world)]
    allf = unshard(pf, n, world)
    return allf[:, :3], allf[:, 3:], unshard(ps, n, world)
