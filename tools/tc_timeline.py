"""Print the clock64 timeline of block 0 of the tensor-core all-pairs kernel (a library built with
-DPSIM_TC_PROF, selected through PSIM_LIBRARY): when each MMA warp saw its barriers and committed, when each
consumer warp saw the accumulator, had it in registers and released it.

    PSIM_LIBRARY=particle_simulation_b200/_build/libpsim_b200_prof.so python tools/tc_timeline.py [--variant 0] [--n 262144]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import particle_simulation_b200 as P                      # noqa: E402
from particle_simulation_b200 import generate as G       # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--n", type=int, default=262144)
    ap.add_argument("--precision", type=int, default=8)
    args = ap.parse_args()
    ids, pos, vel, area = G.synthetic(args.n)
    sim = P.Simulation(P.make_params(length_of_area=area), ids, pos, vel, args.precision)
    sim.set_option(P.OPT_ALLPAIRS_KERNEL, args.variant)
    for _ in range(2):
        sim.set_positions(pos)          # one force-only launch each; the kernel prints its timeline
        sim.synchronize()
    sim.close()


if __name__ == "__main__":
    main()
