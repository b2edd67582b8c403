"""Host-side logic of the multi-GPU path on the CPU: partition arithmetic and the
rank-order merge of per-rank results, exercised with world_size 2 over gloo."""
import os
import socket

import numpy as np
import pytest


def test_shard_bounds_cover_the_range_and_match_the_library_rule():
    from particle_simulation_b200 import sharding as S
    for n in (1, 127, 128, 2000, 39440, 65536, 262144, 1048576):
        for world in (1, 2, 3, 4, 8):
            b = [S.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[r][1] == b[r + 1][0] for r in range(world - 1))
            assert S.padded_size(n, world) % (S.TILE * world) == 0
            sizes = {hi - lo for lo, hi in b if hi - lo > 0 and hi != n}
            assert len(sizes) <= 1                         # equal messages for the all-gather


def test_reference_slices():
    from particle_simulation_b200 import sharding as S
    # get_particle_index (program:305-323): 2000 particles on 3 threads -> 666, 666, 668
    assert [S.reference_slice(2000, t, 3) for t in range(3)] == [(0, 666), (666, 1332), (1332, 2000)]


def _worker(rank, world, port, n, q):
    import torch
    import torch.distributed as dist
    from particle_simulation_b200 import sharding as S
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # the 128-byte communicator id travels by broadcast, as bench.py does with the NCCL id
        uid = torch.arange(128, dtype=torch.uint8) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
        dist.broadcast(uid, 0)
        lo, hi = S.shard_bounds(n, rank, world)
        vel = np.zeros((n, 3))
        vel[lo:hi] = np.arange(lo, hi)[:, None] + np.array([0.0, 0.25, 0.5])   # what a rank's get_frame fills
        t = torch.from_numpy(vel)
        parts = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        bounds = [S.shard_bounds(n, r, world) for r in range(world)]
        merged = S.merge_owned([p.numpy() for p in parts], bounds, n)
        want = np.arange(n)[:, None] + np.array([0.0, 0.25, 0.5])
        # max over ranks of a timing, as bench.py reduces it
        tmax = torch.tensor([float(rank + 1)]); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        q.put((rank, bool(np.array_equal(merged, want)), bool(uid[127] == 127), float(tmax)))
    finally:
        dist.destroy_process_group()


def test_two_rank_merge_over_gloo():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 2000, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True, True, 2.0), (1, True, True, 2.0)]
