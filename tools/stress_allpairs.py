"""Stress one all-pairs kernel: many force-only launches and stepping chunks with an error check and
an audit after each, to catch intermittent launch failures and dropped pairs.

    python tools/stress_allpairs.py [--variant 0] [--n 262144] [--precision 8] [--iters 300] [--steps 5]

Every iteration: psim_set_positions (force-only launch, MODE 0) -> forces back -> filter_pairs against the exact
O(N^2) count every `--audit-every` iterations -> psim_advance(steps) (fused launches, MODE 2 + final MODE 1).
The first failing call is reported with its iteration; a CUDA launch failure poisons the context, so the process
exits there.  Written for round 1's intermittent failure of the pipelined tensor-core kernel (an odd TMA ring let
an MMA warp pass an mbarrier phase early, csrc/psim_allpairs_tc.cuh "NPAIR must be EVEN"), whose one observed
failure was in exactly this call sequence (bench.py's e2e loop).  --variant takes a psim_allpairs_kernel value.
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import particle_simulation_b200 as P                      # noqa: E402
from particle_simulation_b200 import generate as G       # noqa: E402
from particle_simulation_b200 import native              # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--n", type=int, default=262144)
    ap.add_argument("--precision", type=int, default=8, choices=(4, 8))
    ap.add_argument("--iters", type=int, default=300)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--audit-every", type=int, default=25)
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args()

    ids, pos, vel, area = G.synthetic(args.n)
    rt = native.rt(args.precision)
    sim = P.Simulation(P.make_params(length_of_area=area), ids, pos, vel, args.precision, device=args.device)
    sim.set_option(P.OPT_ALLPAIRS_KERNEL, args.variant)
    x = pos.astype(rt)
    f_first = None
    t0 = time.perf_counter()
    for it in range(args.iters):
        where = "set_positions"
        try:
            sim.set_positions(x)
            where = "forces"
            f = sim.forces()
            if f_first is None:
                f_first = f.copy()
            elif not np.array_equal(f, f_first):            # same positions, same kernel: bit-identical forces
                print("iteration %d: forces differ from the first evaluation (max |d| = %g)" % (it, np.abs(f - f_first).max()))
                return 1
            if it % args.audit_every == 0:
                where = "audit"
                c = sim.counters()
                exact = sim.pairs_in_range()
                if c.filter_pairs >= 0 and c.filter_pairs != exact:
                    print("iteration %d: filter evaluated %d in-range pairs, exact count %d" % (it, c.filter_pairs, exact))
                    return 1
            where = "advance"
            sim.advance(args.steps)
        except native.PsimError as e:
            print("iteration %d: %s failed: %s" % (it, where, e))
            return 2
        if it % 50 == 49:
            print("iteration %d ok (%.1f s)" % (it + 1, time.perf_counter() - t0), flush=True)
    c = sim.counters()
    print("ok: %d iterations of kernel %d at N=%d f%d: %d force launches, 0 failures, forces bit-identical, %.1f s" % (
        args.iters, args.variant, args.n, 8 * args.precision, c.force_launches, time.perf_counter() - t0))
    sim.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
