"""Run under torchrun with N >= 2 ranks (one GPU each): the two ways of exchanging the position shards — peer-memory
stores from inside the fused force kernel, and the NCCL all-gather (PSIM_GATHER=nccl) — move the same bytes, so a
run must end in the SAME state bit for bit in both, and twice in a row in the first.  Long runs with rebuilds and
particles leaving the box: a store that arrives late or lands in the wrong buffer shows up as a difference.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/exchange_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    import particle_simulation_b200 as P
    from particle_simulation_b200 import native
    from particle_simulation_b200 import generate as G
    from conftest import two_cubes

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def fresh_uid():
        buf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            buf.copy_(torch.frombuffer(bytearray(native.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(buf, 0)
        return bytes(buf.cpu().numpy().tobytes())

    def run(mode, algo, ids, pos, vel, kw, chunks, precision):
        if mode == "nccl":
            os.environ["PSIM_GATHER"] = "nccl"
        else:
            os.environ.pop("PSIM_GATHER", None)
        sim = P.Simulation(P.make_params(**kw), ids, pos, vel, precision, device=local, rank=rank, world=world,
                           nccl_unique_id=fresh_uid())
        if algo == "cells":
            sim.set_option(P.OPT_FORCE_ALGO, P.FORCE_CELL_BINNED)
            sim.set_positions(pos)
        for c in chunks:
            sim.advance(c)
        _, x, v, a = sim.frame()
        d, fo, fn = sim.aux()
        cnt = sim.counters()
        out = (x.copy(), v[sim.i_lo:sim.i_hi].copy(), a.copy(), fn[sim.i_lo:sim.i_hi].copy(), d[sim.i_lo:sim.i_hi].copy(),
               int(cnt.rebuilds), int(cnt.peer_push_steps))
        sim.close()
        return out

    cases = []
    ids, pos, vel, area = G.synthetic(65536)
    cases.append(("lattice_65536 drifting", ids, pos, vel + np.array([8.0, 0.0, 0.0]), dict(length_of_area=area), (1, 199, 150, 50)))
    ids, pos, vel = two_cubes(side=6, speed=40.0, origin=3.0)
    cases.append(("two cubes colliding, particles leaving", ids, pos, vel, dict(length_of_area=(30.0, 30.0, 30.0), delta_t=0.002), (1, 299, 100)))
    ok = True
    for name, ids, pos, vel, kw, chunks in cases:
        for algo in ("allpairs", "cells"):
            for precision in (8, 4):
                a = run("peer", algo, ids, pos, vel, kw, chunks, precision)
                b = run("nccl", algo, ids, pos, vel, kw, chunks, precision)
                c = run("peer", algo, ids, pos, vel, kw, chunks, precision)
                same_ab = all(np.array_equal(p, q) for p, q in zip(a[:5], b[:5])) and a[5] == b[5]
                same_ac = all(np.array_equal(p, q) for p, q in zip(a[:5], c[:5])) and a[5] == c[5]
                good = same_ab and same_ac and b[6] == 0
                t = torch.tensor([1.0 if good else 0.0], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MIN)
                if rank == 0:
                    print("%s %s, %s, f%d, %d ranks, %d steps: peer stores (%d steps) vs NCCL all-gather %s, peer stores twice %s; "
                          "rebuilds %d, inactive %d" % ("ok      " if t.item() == 1.0 else "MISMATCH", name, algo, 8 * precision, world,
                                                        sum(chunks), a[6], "identical" if same_ab else "DIFFER",
                                                        "identical" if same_ac else "DIFFER", a[5], int((a[2] == 0).sum())), flush=True)
                elif not good:
                    print("rank %d: MISMATCH %s %s f%d ab=%s ac=%s" % (rank, name, algo, 8 * precision, same_ab, same_ac), flush=True)
                ok = ok and t.item() == 1.0
    if rank == 0 and ok:
        print("exchange check ok (%d ranks)" % world)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
