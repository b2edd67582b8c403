"""Golden vectors on the reference's OWN shipped case (BASELINE.json configs[0]):
particle_files/test_particles.dat (N = 2,000: two 10^3 lattice cubes approaching each other) with
environment_params.nml.  Run from the repo root, in the container that has /root/reference:

    python tests/golden/make_golden_shipped.py

The reference's files exist only there (the GPU box has no /root/reference), so the inputs travel inside the
fixture: ids, positions and velocities exactly as the file holds them (4 decimals), the namelist values, and the
outputs of the independent numpy brute force of make_golden.py (float64, all pairs) — forces at t = 0 and
positions, velocities and forces after WRITE_TIME_STEP = 100 steps, i.e. the state of the reference's first
frame (program_particle_simulation_openmp.f08:109-115).  Every particle stays inside the box over that
horizon (0.5 length units of travel), so the reference's rebuilds (displacement resets) have no physical effect
and brute force is its exact physical equivalent.

A second fixture covers particle_files/more_test_particles.dat (N = 39,440: a 16^3 body at speed 10 towards a
35,344-particle slab), forces at t = 0 only.  A dense pair matrix would need 12 GB, so the pairs come from a
KD-tree (scipy.spatial.cKDTree.query_pairs, r <= r_cut) and the strict r < r_cut test and the force formula are
then applied in numpy float64, accumulated with np.add.at — again sharing no code with the oracle or CUDA path.

Writes tests/golden/shipped_test_particles_golden.npz and shipped_more_test_particles_golden.npz
(binary; no reference text is copied into the repo).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

REF = "/root/reference"


def main():
    import make_golden as mg
    from particle_simulation_b200 import formats

    nml = formats.read_namelist(os.path.join(REF, "environment_params.nml"))
    ids, x, v = formats.read_particles(os.path.join(REF, "particle_files", "test_particles.dat"))
    p = dict(sigma=nml["sigma"], epsilon=nml["epsilon"], mass=nml["mass"], r_cut=nml["r_cut"], delta_t=nml["delta_t"],
             length_of_area=tuple(nml["length_of_area"]), length_of_cell=tuple(nml["length_of_cell"]))
    k_steps = int(nml["write_time_step"])
    x0, v0 = x.copy(), v.copy()
    _, _, _, tmf = mg.constants(p)
    f0, pairs0 = mg.forces(x, p)
    e0 = mg.energy(x, v, p)
    f_new = f0.copy()
    for _ in range(k_steps):
        f_old = f_new
        x = x + p["delta_t"] * (v + tmf * f_old)
        f_new, pairs = mg.forces(x, p)
        v = v + tmf * (f_new + f_old)
    e1 = mg.energy(x, v, p)
    area = np.array(p["length_of_area"])
    assert (x0 > 0).all() and (x0 < area).all() and (x > 0).all() and (x < area).all(), "a particle left the box"
    np.savez_compressed(
        os.path.join(HERE, "shipped_test_particles_golden.npz"),
        ids=ids, x0=x0, v0=v0, f0=f0, pairs0=pairs0, k_steps=k_steps, x=x, v=v, f=f_new, pairs=pairs,
        e0=np.array([e0[0], e0[1], *e0[2]]), e1=np.array([e1[0], e1[1], *e1[2]]),
        nb_of_threads=int(nml["nb_of_threads"]), sigma=p["sigma"], epsilon=p["epsilon"], mass=p["mass"], r_cut=p["r_cut"],
        delta_t=p["delta_t"], length_of_area=area, length_of_cell=np.array(p["length_of_cell"]))
    print("golden written: N=%d, ordered pairs in range at t=0 %d, after %d steps %d" % (len(ids), pairs0, k_steps, pairs))
    print("energy drift", (e1[0] + e1[1]) - (e0[0] + e0[1]), "of", e0[0] + e0[1])


def forces_kdtree(x, p, mg):
    from scipy.spatial import cKDTree
    r_min, r_cut, F_rcut, _ = mg.constants(p)
    pairs = cKDTree(x).query_pairs(r_cut * (1.0 + 1e-12), output_type="ndarray")     # i < j, a hair generous
    d = x[pairs[:, 0]] - x[pairs[:, 1]]                                            # r_ij = x_i - x_j
    r = np.sqrt((d * d).sum(-1))
    keep = r < r_cut                                                               # stoermer:121, strict
    pairs, d, r = pairs[keep], d[keep], r[keep]
    s = (r_min / r) ** 6
    fij = -(d * (12.0 * p["epsilon"] * ((s / (r * r)) * (1.0 - s) - F_rcut))[:, None])   # force on i
    f = np.zeros_like(x)
    np.add.at(f, pairs[:, 0], fij)
    np.add.at(f, pairs[:, 1], -fij)
    return f, 2 * len(pairs)


def main_more():
    import make_golden as mg
    from particle_simulation_b200 import formats

    nml = formats.read_namelist(os.path.join(REF, "environment_params.nml"))
    ids, x, v = formats.read_particles(os.path.join(REF, "particle_files", "more_test_particles.dat"))
    p = dict(sigma=nml["sigma"], epsilon=nml["epsilon"], mass=nml["mass"], r_cut=nml["r_cut"], delta_t=nml["delta_t"],
             length_of_area=tuple(nml["length_of_area"]), length_of_cell=tuple(nml["length_of_cell"]))
    area = np.array(p["length_of_area"])
    assert (x > 0).all() and (x < area).all()
    f0, pairs0 = forces_kdtree(x, p, mg)
    # cross-check the KD-tree route against the dense route of make_golden.py on the first body
    sub = slice(0, 4096)
    near = np.zeros(len(x), bool); near[sub] = True
    fd, _ = mg.forces(x[sub], p)
    from scipy.spatial import cKDTree
    touching = cKDTree(x[~near]).query_ball_point(x[sub], p["r_cut"] * p["sigma"] * 1.001, return_length=True)
    iso = touching == 0                                                            # rows of body 1 with no neighbour in body 2
    assert iso.any() and np.abs(fd[iso] - f0[sub][iso]).max() <= 1e-12 * np.abs(fd).max()
    np.savez_compressed(
        os.path.join(HERE, "shipped_more_test_particles_golden.npz"),
        ids=ids, x0=x.astype(np.float32), v0=v.astype(np.float32), f0=f0, pairs0=pairs0,
        nb_of_threads=int(nml["nb_of_threads"]), sigma=p["sigma"], epsilon=p["epsilon"], mass=p["mass"], r_cut=p["r_cut"],
        delta_t=p["delta_t"], length_of_area=area, length_of_cell=np.array(p["length_of_cell"]))
    # x0/v0 are stored as float32 to keep the fixture small: the file's 4-decimal values below 100 survive the round
    # trip through float32 exactly AFTER re-rounding to 4 decimals, which the loader does
    back = np.round(x.astype(np.float32).astype(np.float64), 4)
    assert np.array_equal(back, x) and np.array_equal(np.round(v.astype(np.float32).astype(np.float64), 4), v)
    print("golden written: N=%d, ordered pairs in range at t=0 %d, max |f| %.6g" % (len(ids), pairs0, np.abs(f0).max()))


if __name__ == "__main__":
    main()
    main_more()
