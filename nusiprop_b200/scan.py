"""Parameter scans sharded statically over the GPUs of one box.

Each (mphi, g, mntot) point is an independent evolve() (reference: one calculate_flux object per
point), so the only exchange is the final gather of the results.  Point p of the scan belongs to
rank p mod world_size (strided: the cost of a point varies smoothly over the grid).  One process
per GPU, `torch.distributed` for the gather (NCCL over NVLink on GPUs; the host-side logic is backend
agnostic and is tested with gloo on CPU).

`ScanRunner` is the device-resident form: the shard's points are prepared and uploaded once
(`Context.upload_points`), every `step()` evolves them with the outputs left in HBM and gathers the
fluxes to rank 0 piece by piece -- the gather of piece k (an asynchronous NCCL gather, issued as soon as
the piece's integrator is queued) runs under the kernels of piece k+1, so only the last piece's gather is
exposed.  `evolve_scan_distributed` is the one-call form built on it.  bench.py --gpus N times
`ScanRunner.step()`.
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


def _dist_state():
    import torch.distributed as dist
    on = dist.is_available() and dist.is_initialized()
    return dist, (dist.get_rank() if on else 0), (dist.get_world_size() if on else 1)


class ScanRunner(object):
    """One rank's share of a scan, resident on its GPU.

    Every rank passes the FULL point arrays (they are cheap); the runner keeps rank::world of them.  Outputs per
    step: this rank's `flux_mass`, `flux_fla` [n_local, 3, N] and `status` [n_local] as torch device tensors; on
    rank 0 additionally `gathered` = per-rank flavour fluxes (and mass fluxes when gather_mass), each padded to
    the longest shard.  pieces: how many pieces the shard is cut into for the overlapped gather (default: the
    library's pipeline slices of 2048 points)."""

    PIECE = 2048

    def __init__(self, ctx, mphi, g, mntot, gather_mass=False, piece=None):
        import torch
        self.torch = torch
        self.dist, self.rank, self.world = _dist_state()
        mphi, g, mntot = (np.ascontiguousarray(np.broadcast_to(np.asarray(a, dtype=np.float64), np.broadcast(mphi, g, mntot).shape)).ravel()
                          for a in (mphi, g, mntot))
        self.n = mphi.size
        self.idx = shard_indices(self.n, self.rank, self.world)
        self.ctx = ctx
        self.N = ctx.N_bins_E
        self.dev = torch.device("cuda", ctx.device)
        self.n_local = len(self.idx)
        self.n_max = (self.n + self.world - 1) // self.world  # shards differ in length by at most one row
        piece = piece or self.PIECE
        self.cuts = [(a, min(a + piece, self.n_local)) for a in range(0, self.n_local, piece)]
        # a scan row must not be cut in the middle of a (mphi, m_nu) group more often than needed: pieces are the
        # library's own slice length, so the groups are the ones a single batch would get
        self.points = [ctx.upload_points(mphi[self.idx[a:b]], g[self.idx[a:b]], mntot[self.idx[a:b]]) for a, b in self.cuts]
        self.gather_mass = bool(gather_mass)
        C = 6 if gather_mass else 3
        # [n_max][C][N]: flavour fluxes first (what a scan consumer reads), mass fluxes behind them when asked for
        self.flux_mass = torch.empty((self.n_max, 3, self.N), dtype=torch.float64, device=self.dev)
        self.flux_fla = torch.empty((self.n_max, 3, self.N), dtype=torch.float64, device=self.dev)
        self.status = torch.zeros(self.n_max, dtype=torch.int32, device=self.dev)
        if self.n_local < self.n_max:  # padding rows of a short shard: defined values
            self.flux_mass[self.n_local:].zero_()
            self.flux_fla[self.n_local:].zero_()
        self.gathered = None
        self.gathered_mass = None
        self.gathered_status = None
        if self.world > 1 and self.rank == 0:
            self.gathered = [torch.empty_like(self.flux_fla) for _ in range(self.world)]
            self.gathered_status = [torch.empty_like(self.status) for _ in range(self.world)]
            if gather_mass:
                self.gathered_mass = [torch.empty_like(self.flux_mass) for _ in range(self.world)]
        # every rank must issue the same sequence of collectives: cut the padded length, not the local one
        self.gcuts = [(a, min(a + piece, self.n_max)) for a in range(0, self.n_max, piece)]

    def step(self, gather=True):
        """Evolve the shard (enqueue only) and, with more than one rank, gather to rank 0.  Returns when everything is
        queued and the asynchronous gathers have been waited for on the current stream (no host synchronisation)."""
        torch, N = self.torch, self.N
        stream = torch.cuda.current_stream(self.dev).cuda_stream
        works = []
        for k, (a, b) in enumerate(self.cuts):
            self.ctx.evolve_points(self.points[k], self.flux_mass[a:b].data_ptr(), self.flux_fla[a:b].data_ptr(), self.status[a:b].data_ptr(), stream)
            if self.world > 1 and gather and k < len(self.gcuts):
                works += self._gather_piece(k)
        if self.world > 1 and gather:
            for k in range(len(self.cuts), len(self.gcuts)):
                works += self._gather_piece(k)
            for w in works:
                w.wait()  # stream-level wait for NCCL work; does not block the host
        return self

    def _gather_piece(self, k):
        a, b = self.gcuts[k]
        dist, r0 = self.dist, self.rank == 0
        ws = [dist.gather(self.flux_fla[a:b], [t[a:b] for t in self.gathered] if r0 else None, dst=0, async_op=True)]
        if self.gather_mass:
            ws.append(dist.gather(self.flux_mass[a:b], [t[a:b] for t in self.gathered_mass] if r0 else None, dst=0, async_op=True))
        if k == len(self.gcuts) - 1:  # the status words once, at the end
            ws.append(dist.gather(self.status, self.gathered_status if r0 else None, dst=0, async_op=True))
        return ws

    def result(self):
        """Rank 0: (flux_mass or None, flux_fla [n,3,N], status [n]) as host arrays in the scan's order (flux_mass only
        with gather_mass, or world 1); other ranks: None.  Synchronises."""
        self.torch.cuda.synchronize(self.dev)
        if self.world == 1:
            return self.flux_mass[: self.n].cpu().numpy(), self.flux_fla[: self.n].cpu().numpy(), self.status[: self.n].cpu().numpy()
        if self.rank != 0:
            return None
        cnt = [len(shard_indices(self.n, r, self.world)) for r in range(self.world)]
        ff = unshard([self.gathered[r][: cnt[r]].cpu().numpy() for r in range(self.world)], self.n, self.world)
        st = unshard([self.gathered_status[r][: cnt[r]].cpu().numpy() for r in range(self.world)], self.n, self.world)
        fm = None
        if self.gather_mass:
            fm = unshard([self.gathered_mass[r][: cnt[r]].cpu().numpy() for r in range(self.world)], self.n, self.world)
        return fm, ff, st

    def close(self):
        for p in self.points:
            p.close()
        self.points = []


def evolve_scan_distributed(mphi, g, mntot, ctx_kwargs=None, evolve_fn=None, device=None):
    """Run a scan across the ranks of the default process group (or alone when it is not
    initialised).  Every rank passes the FULL point arrays; rank 0 returns
    (flux_mass[n,3,N], flux_fla[n,3,N], status[n]), other ranks return None.

    Default: a nusiprop_b200.Context on `device` (LOCAL_RANK) and a ScanRunner (device-resident evolve, one
    overlapped gather over NCCL).  evolve_fn(mphi, g, mntot) -> (flux_mass, flux_fla, status) replaces the GPU
    context in the CPU tests of the sharding logic (gloo): host arrays, one gather.
    """
    import torch
    dist, rank, world = _dist_state()
    mphi, g, mntot = (np.ascontiguousarray(np.broadcast_to(np.asarray(a, dtype=np.float64), np.broadcast(mphi, g, mntot).shape)).ravel()
                      for a in (mphi, g, mntot))
    n = mphi.size
    if evolve_fn is None:
        import os
        from .api import Context
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        ctx = Context(device=device, **(ctx_kwargs or {}))
        run = ScanRunner(ctx, mphi, g, mntot, gather_mass=True)
        try:
            return run.step().result()
        finally:
            run.close()
            ctx.close()
    idx = shard_indices(n, rank, world)
    fm, ff, st = evolve_fn(mphi[idx], g[idx], mntot[idx])
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
