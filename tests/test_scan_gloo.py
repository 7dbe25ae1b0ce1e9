"""world_size-2 gloo test of the scan sharding / gather logic (the only multi-GPU step of the path)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp


def _fake_evolve(mphi, g, mntot):
    n, N = len(mphi), 7
    fm = np.stack([np.outer(mphi, np.arange(1, N + 1)) * (k + 1) for k in range(3)], axis=1)
    ff = fm + g[:, None, None]
    st = (mphi > 50).astype(np.int32)
    return fm, ff, st


def _worker(rank, world, port, n, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nusiprop_b200.scan import evolve_scan_distributed
    rng = np.random.default_rng(3)
    mphi, g = rng.uniform(1, 100, n), rng.uniform(0, 1, n)
    out = evolve_scan_distributed(mphi, g, 0.1, evolve_fn=_fake_evolve)
    if rank == 0:
        fm0, ff0, st0 = _fake_evolve(mphi, g, np.full(n, 0.1))
        q.put(bool(np.array_equal(out[0], fm0) and np.array_equal(out[1], ff0) and np.array_equal(out[2], st0)))
    else:
        q.put(out is None)
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [8, 11])
def test_sharded_scan_gathers_in_point_order(n):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(res)


def test_shard_roundtrip():
    from nusiprop_b200.scan import shard_indices, unshard
    n, world = 13, 4
    data = np.arange(n * 2).reshape(n, 2)
    parts = [data[shard_indices(n, r, world)] for r in range(world)]
    assert np.array_equal(unshard(parts, n, world), data)
    assert sorted(np.concatenate([shard_indices(n, r, world) for r in range(world)]).tolist()) == list(range(n))
