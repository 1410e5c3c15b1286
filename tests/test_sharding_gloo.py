"""Multi-rank host logic on CPU: world_size 2 and 3 over gloo (127.0.0.1).  Each rank takes the slabs / level chunks the
plan gives it, "regrids" them (the oracle stands in for the GPU here: this test is about partitioning and gathering, the
kernels have their own GPU tests) and rank 0 reassembles; the result must equal the single-process result."""
import os
import socket

import numpy as np
import pytest

from fimex_b200 import sharding


def test_plan_covers_everything_once():
    for n_slabs, nz, world in [(67, 65, 8), (67, 65, 1), (3, 65, 8), (1, 65, 8), (2, 5, 8), (8, 1, 3), (5, 65, 2)]:
        p = sharding.plan(n_slabs, nz, world)
        assert len(p) == world
        seen = np.zeros((n_slabs, nz), int)
        for items in p:
            for (t, z0, z1) in items:
                seen[t, z0:z1] += 1
        assert (seen == 1).all(), (n_slabs, nz, world)
    # 67 slabs on 8 ranks: 9/8 imbalance
    sizes = sorted(len(x) for x in sharding.plan(67, 65, 8))
    assert sizes == [8, 8, 8, 8, 8, 9, 9, 9]
    # fewer slabs than ranks: the level axis is split so that nobody idles
    assert all(len(x) >= 1 for x in sharding.plan(1, 65, 8))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_slabs, nz, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys
        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        import oracle_lib as ol
        rng = np.random.default_rng(0)  # same tables and data on every rank (the cache is replicated)
        in_x, in_y, out_x, out_y = 23, 19, 31, 17
        px, py = rng.uniform(-1, in_x, out_x * out_y), rng.uniform(-1, in_y, out_x * out_y)
        data = rng.normal(size=(n_slabs, nz, in_y, in_x)).astype(np.float32)
        local = {}
        for (t, z0, z1) in sharding.plan(n_slabs, nz, world)[rank]:
            local[(t, z0, z1)], _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, data[t, z0:z1])
        merged = sharding.gather_pieces(local, dst=0)
        if rank == 0:
            slabs = sharding.assemble(merged, n_slabs, nz)
            want = [ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, data[t])[0] for t in range(n_slabs)]
            ok = all(np.array_equal(np.nan_to_num(a, nan=-1), np.nan_to_num(b, nan=-1)) for a, b in zip(slabs, want))
            q.put(("ok" if ok else "mismatch", len(merged)))
        else:
            assert merged is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_slabs,nz", [(2, 5, 4), (3, 1, 7), (2, 2, 3)])
def test_two_and_three_ranks_over_gloo(world, n_slabs, nz):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_slabs, nz, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    status, n = q.get(timeout=10)
    assert status == "ok" and n >= n_slabs
