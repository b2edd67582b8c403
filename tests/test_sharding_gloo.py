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


def _trigger_worker(rank, world, port, n, slices, seed, q):
    import torch
    import torch.distributed as dist
    from particle_simulation_b200 import sharding as S
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(seed)                   # every rank holds the same displacements, as on the GPUs
        verdicts = []
        for trial in range(40):
            disp = rng.uniform(0.0, 0.2, n)
            movers = rng.choice(n, 2, replace=False)
            disp[movers] = rng.uniform(0.2, 0.4, 2)         # two large ones: same slice or not decides the verdict
            active = (rng.uniform(size=n) > 0.1).astype(np.uint8)
            mine = torch.from_numpy(S.slice_pairs_of_rank(disp, active, n, rank, world, slices))
            parts = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(parts, mine)                    # the 2 x slices values per rank of the step's NCCL group
            got = S.merged_trigger(np.stack([p.numpy() for p in parts]), 0.5)
            # the reference: each thread slice on its own (single process, all particles)
            want = False
            for k in range(slices):
                a, b = S.reference_slice(n, k, slices)
                v = np.sort(disp[a:b][active[a:b].astype(bool)])[::-1]
                want = want or (len(v) >= 2 and v[0] + v[1] > 0.5)
            verdicts.append((got, want))
        q.put((rank, all(g == w for g, w in verdicts), sum(w for _, w in verdicts)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("slices", [1, 3])
def test_per_slice_rebuild_trigger_merged_over_two_ranks(slices):
    """The reference evaluates its rebuild trigger per NB_OF_THREADS slice; a slice may span ranks.  Per-rank
    per-slice top-2 pairs, all-gathered and merged slice by slice, must give the reference's verdict."""
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_trigger_worker, args=(r, 2, port, 2000, slices, 17, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    fired = res[0][2]
    assert 0 < fired < 40, "test lost its point: the verdict should vary between trials (%d of 40 fired)" % fired
