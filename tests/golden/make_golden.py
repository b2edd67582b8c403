"""Generates the golden fixtures of tests/golden/ (run from the repo root):

    python tests/golden/make_golden.py

The reference is Fortran and cannot be run here, and it ships no
golden outputs of its own (F5).  The vectors below are therefore produced by a
THIRD implementation that shares no code with the oracle or with the CUDA
path: plain numpy, brute force over all pairs, float64, written directly from
the formulas of fortran_src/module_stoermer_verlet_method.f08:68-81 (constants),
:117-129 (pair force), :155-162 (drift) and :199-202 (kick).  The scenario keeps
every particle inside the box, so the neighbour-list machinery of the reference
has no observable effect and brute force is its exact physical equivalent.

Files written:
  two_cubes.nml / two_cubes.dat     inputs in the reference's own formats
  two_cubes_golden.npz              forces at t=0; positions, velocities, forces after K steps;
                                    kinetic/potential energy and momentum at both times
  pair_force_kat.json               closed-form pair force at the radii of SURVEY.md 8(c)
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

K_STEPS = 40
PARAMS = dict(sigma=1.0, epsilon=1.0, mass=1.0, r_cut=2.5, delta_t=0.001,
              length_of_area=(60.0, 60.0, 60.0), length_of_cell=(3.0, 3.0, 3.0))


def constants(p):
    r_min = 2.0 ** (1.0 / 6.0) * p["sigma"]
    r_cut = p["r_cut"] * p["sigma"]
    s = (r_min / r_cut) ** 6
    F_rcut = (s / (r_cut * r_cut)) * (1.0 - s)
    tmf = 0.5 * p["delta_t"] / p["mass"]
    return r_min, r_cut, F_rcut, tmf


def forces(x, p):
    r_min, r_cut, F_rcut, _ = constants(p)
    d = x[:, None, :] - x[None, :, :]                  # r_ij = x_i - x_j
    r = np.sqrt((d * d).sum(-1))
    np.fill_diagonal(r, np.inf)
    inside = r < r_cut
    s = np.where(inside, (r_min / r) ** 6, 0.0)
    F = np.where(inside, (s / (r * r)) * (1.0 - s) - F_rcut, 0.0)
    # force on i = - sum_j r_ij * 12 eps (F_rij - F_rcut)
    return -(d * (12.0 * p["epsilon"] * F)[:, :, None]).sum(1), int(inside.sum())


def energy(x, v, p):
    r_min, r_cut, F_rcut, _ = constants(p)
    eps = p["epsilon"]
    d = x[:, None, :] - x[None, :, :]
    r2 = (d * d).sum(-1)
    iu = np.triu_indices(len(x), 1)
    r2 = r2[iu]
    r2 = r2[r2 < r_cut * r_cut]
    sc = (r_min / r_cut) ** 6
    C = -(eps * (sc * sc - 2 * sc) - 6 * eps * F_rcut * r_cut * r_cut)
    s = (r_min * r_min / r2) ** 3
    U = (eps * (s * s - 2 * s) - 6 * eps * F_rcut * r2 + C).sum()
    K = 0.5 * p["mass"] * (v * v).sum()
    return K, U, p["mass"] * v.sum(0)


def main():
    from particle_simulation_b200 import formats, generate

    rng = np.random.Generator(np.random.PCG64(2024))
    a = generate.lattice_block((6, 6, 6), (20.0, 20.0, 20.0), (5.0, 5.0, 5.0), 0.1, rng=rng)
    b = generate.lattice_block((6, 6, 6), (31.0, 31.0, 31.0), (-5.0, -5.0, -5.0), 0.1, rng=rng)
    ids = np.concatenate([a[0], b[0]]); x = np.concatenate([a[1], b[1]]); v = np.concatenate([a[2], b[2]])
    formats.write_particles(os.path.join(HERE, "two_cubes.dat"), ids, x, v)
    with open(os.path.join(HERE, "two_cubes.nml"), "w") as f:
        f.write("&ENVIRONMENTPARAMS\n NB_OF_THREADS   = 1,\n SIGMA           = 1.0,\n EPSILON         = 1.0,\n"
                " MASS            = 1.0,\n R_CUT           = 2.5,\n LENGTH_OF_AREA  = 60.0, 60.0, 60.0,\n"
                " LENGTH_OF_CELL  = 3.0, 3.0, 3.0,\n DELTA_T         = 0.001,\n WRITE_TIME_STEP = 20,\n"
                " MAX_TIME_STEP   = %d\n /\n" % K_STEPS)
    # read back what the file holds (4 decimals), exactly what every implementation will see
    ids, x, v = formats.read_particles(os.path.join(HERE, "two_cubes.dat"))
    p = PARAMS
    _, _, _, tmf = constants(p)
    f0, pairs0 = forces(x, p)
    e0 = energy(x, v, p)
    f_new = f0.copy()
    for _ in range(K_STEPS):
        f_old = f_new
        x = x + p["delta_t"] * (v + tmf * f_old)
        f_new, pairs = forces(x, p)
        v = v + tmf * (f_new + f_old)
    e1 = energy(x, v, p)
    np.savez_compressed(os.path.join(HERE, "two_cubes_golden.npz"), ids=ids, f0=f0, pairs0=pairs0, k_steps=K_STEPS,
                        x=x, v=v, f=f_new, f_old=f_old, pairs=pairs,
                        e0=np.array([e0[0], e0[1], *e0[2]]), e1=np.array([e1[0], e1[1], *e1[2]]))
    # closed form pair force on i along r_ij, shipped parameters
    r_min, r_cut, F_rcut, _ = constants(p)
    kat = {}
    for r in (0.95, 1.0, r_min, 1.5, 2.0, 2.49, 2.5, 3.0):
        s = (r_min / r) ** 6
        kat[repr(r)] = 0.0 if r >= r_cut else -(r * 12.0 * p["epsilon"] * ((s / (r * r)) * (1.0 - s) - F_rcut))
    with open(os.path.join(HERE, "pair_force_kat.json"), "w") as f:
        json.dump({"r_min": r_min, "r_cut": r_cut, "F_rcut": F_rcut, "time_mass_factor": tmf, "force_on_i_along_rij": kat}, f, indent=1)
    print("golden written: N=%d, pairs at t=0 %d, after %d steps %d" % (len(ids), pairs0, K_STEPS, pairs))
    print("energy drift", (e1[0] + e1[1]) - (e0[0] + e0[1]))


if __name__ == "__main__":
    main()
