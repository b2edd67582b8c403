#!/usr/bin/env python
"""Seeded counterpart of the reference's python_src/write_particle_file.py: same
arguments and output format, plus --seed (the reference draws its thermal noise
from an unseeded global RNG, write_particle_file.py:48).

  python tools/write_particle_file.py -o block.dat -n 64 32 32 -s 49.5 49.5 49.5 -vt 0.1 --seed 12345
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from particle_simulation_b200 import formats, generate  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("-o", "--output", required=True, help="path to output file")
ap.add_argument("-n", "--nb_of_particles", type=int, required=True, nargs=3, help="number of particles in x, y, z direction")
ap.add_argument("-s", "--start_point", type=float, default=[0.0, 0.0, 0.0], nargs=3, help="center of particle mass in x, y, z direction")
ap.add_argument("-r", "--radius", type=float, default=1.0, help="radius of particle")
ap.add_argument("-v", "--velocity", type=float, default=[0.0, 0.0, 0.0], nargs=3, help="General velocity of all particles")
ap.add_argument("-vt", "--thermal_noise", type=float, default=0.0, help="Amount of thermal noise of particles, i.e. temperature")
ap.add_argument("-a", "--append", action="store_true", help="Append to existing file")
ap.add_argument("--seed", type=int, default=generate.DEFAULT_SEED)
# accepted for command-line compatibility; the reference parses but never uses them
ap.add_argument("-e", "--epsilon", type=float, default=1.0)
ap.add_argument("-m", "--mass", type=float, default=1.0)
args = ap.parse_args()
if not args.append and os.path.exists(args.output):
    sys.exit("refusing to overwrite %s (the reference opens it with mode 'x')" % args.output)
ids, pos, vel = generate.lattice_block(args.nb_of_particles, args.start_point, args.velocity, args.thermal_noise,
                                       radius=args.radius, seed=args.seed)
formats.write_particles(args.output, ids, pos, vel, append=args.append)
print("%d particles -> %s" % (len(ids), args.output))
