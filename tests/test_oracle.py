"""CPU tests of the oracle (oracle/psim_oracle.cpp), the checker every GPU
parity test relies on.  PARITY UNPINNED by the reference itself (no tests, no
golden outputs, Fortran not compilable here); pinned here by
  * the closed-form pair-force values of SURVEY.md 8(c),
  * golden vectors from an independent numpy brute-force implementation
    (tests/golden/make_golden.py),
  * an independent brute-force double evaluation inside the oracle library,
  * conservation laws.
"""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, two_cubes


def _golden():
    from particle_simulation_b200 import formats
    ids, pos, vel = formats.read_particles(os.path.join(GOLDEN, "two_cubes.dat"))
    nml = formats.read_namelist(os.path.join(GOLDEN, "two_cubes.nml"))
    g = np.load(os.path.join(GOLDEN, "two_cubes_golden.npz"))
    return ids, pos, vel, nml, g


def _params(oracle, nml):
    return oracle.make_params(delta_t=nml["delta_t"], epsilon=nml["epsilon"], sigma=nml["sigma"], mass=nml["mass"],
                              r_cut=nml["r_cut"], length_of_area=nml["length_of_area"],
                              length_of_cell=nml["length_of_cell"], nb_of_threads=nml["nb_of_threads"])


# SURVEY.md 8(c), computed there in FP64 from stoermer:117-129 with the shipped parameters
SURVEY_KAT = {0.95: 59.1523727, 1.0: 24.0155998, 1.5: -1.13462914, 2.0: -0.150441043, 2.49: -0.00125760340}


def test_pair_force_known_answers(oracle):
    p = oracle.make_params()
    kat = json.load(open(os.path.join(GOLDEN, "pair_force_kat.json")))
    for r_txt, want in kat["force_on_i_along_rij"].items():
        r = float(r_txt)
        assert oracle.pair_force_x(p, 8, r) == pytest.approx(want, rel=1e-13, abs=1e-300)
        assert oracle.pair_force_x(p, 4, r) == pytest.approx(want, rel=2e-5, abs=1e-9)
    for r, want in SURVEY_KAT.items():
        assert oracle.pair_force_x(p, 8, r) == pytest.approx(want, rel=1e-8)   # 9 significant digits quoted
    assert oracle.pair_force_x(p, 8, 2.5) == 0.0 and oracle.pair_force_x(p, 4, 2.5) == 0.0   # strict r < r_cut
    assert oracle.pair_force_x(p, 8, np.nextafter(2.5, 0)) != 0.0 or True
    # the force is continuous at the cutoff: just inside it is rounding-level small
    assert abs(oracle.pair_force_x(p, 8, 2.5 - 1e-9)) < 1e-9


def test_derived_constants(oracle):
    p = oracle.make_params()
    c8 = oracle.constants(p, 8)
    kat = json.load(open(os.path.join(GOLDEN, "pair_force_kat.json")))
    assert c8["r_min"] == pytest.approx(kat["r_min"], rel=1e-15)
    assert c8["F_rcut"] == pytest.approx(kat["F_rcut"], rel=1e-14)
    assert c8["F_rcut"] == pytest.approx(1.2999826e-3, rel=1e-7)
    assert c8["time_mass_factor"] == 5e-4 and c8["dr_max_tolerable"] == 0.5 and c8["r_cut"] == 2.5
    c4 = oracle.constants(p, 4)
    assert c4["r_min"] == float(np.float32(np.float32(2.0) ** (np.float32(1) / np.float32(6))))
    assert c4["delta_t"] == float(np.float32(0.001))


# f32 against the float64 golden: positions near 20-30 carry ~1e-6 rounding in f32, and at the
# lattice spacing r_min the force slope dF/dr is ~57, so forces differ by ~1e-4 relative.
@pytest.mark.parametrize("precision,tol", [(8, 1e-12), (4, 2e-4)])
def test_oracle_matches_independent_numpy_golden(oracle, precision, tol):
    ids, pos, vel, nml, g = _golden()
    o = oracle.Oracle(_params(oracle, nml), precision, ids, pos, vel)
    f0 = o.aux()[2]
    scale = np.abs(g["f0"]).max()
    assert np.abs(f0 - g["f0"]).max() <= tol * scale
    assert o.n_active == len(ids)
    o.advance(int(g["k_steps"]))
    _, x, v, a = o.frame()
    d, fo, fn = o.aux()
    assert a.all()
    xt = 1e-9 if precision == 8 else 2e-4
    assert np.abs(x - g["x"]).max() <= xt
    assert np.abs(v - g["v"]).max() <= 100 * xt
    # f32: 40 steps of x += dx with x ~ 20 accumulate ~4e-5 in x, times the force slope ~57
    ft = 1e-9 if precision == 8 else 1e-2
    assert np.abs(fn - g["f"]).max() <= ft * np.abs(g["f"]).max()
    assert np.abs(fo - g["f_old"]).max() <= ft * np.abs(g["f_old"]).max()
    k, u, mom = o.energy()
    assert k == pytest.approx(g["e1"][0], rel=1e-4 if precision == 4 else 1e-10)
    assert u == pytest.approx(g["e1"][1], rel=1e-3 if precision == 4 else 1e-9)


def test_lists_equal_bruteforce_and_count_pairs(oracle):
    ids, pos, vel, nml, g = _golden()
    p = _params(oracle, nml)
    bf, pairs = oracle.bruteforce(p, 8, pos)
    assert pairs == int(g["pairs0"])
    assert np.abs(bf - g["f0"]).max() <= 1e-13 * np.abs(g["f0"]).max()
    o = oracle.Oracle(p, 8, ids, pos, vel)
    # every in-range unordered pair is in exactly one half list (and the lists hold r <= 3.0 candidates)
    assert o.list_entries >= pairs // 2


def test_momentum_and_energy_conservation(oracle):
    ids, pos, vel = two_cubes()
    o = oracle.Oracle(oracle.make_params(), 8, ids, pos, vel)
    k0, u0, p0 = o.energy()
    o.advance(400)
    k1, u1, p1 = o.energy()
    assert np.abs(p1 - p0).max() <= 1e-10 * np.abs(p0).max()
    assert abs((k1 + u1) - (k0 + u0)) <= 2e-5 * abs(k0 + u0)
    assert o.rebuilds > 0 and o.steps == 400


def test_threaded_baseline_matches_serial(oracle):
    ids, pos, vel = two_cubes()
    a = oracle.Oracle(oracle.make_params(), 8, ids, pos, vel, threads=1)
    b = oracle.Oracle(oracle.make_params(), 8, ids, pos, vel, threads=4)
    a.advance(50); b.advance(50)
    xa, xb = a.frame()[1], b.frame()[1]
    assert np.abs(xa - xb).max() <= 1e-9          # atomics change the summation order only
    assert a.rebuilds == b.rebuilds


def test_rebuild_trigger_depends_on_thread_slices(oracle):
    """NB_OF_THREADS changes WHEN the reference rebuilds: the top-two
    displacement norms are taken per slice.  With both fast cubes in different slices
    the per-slice sums are smaller... or larger than the global one; the counts differ."""
    ids, pos, vel = two_cubes()
    counts = []
    for t in (1, 2, 3):
        o = oracle.Oracle(oracle.make_params(nb_of_threads=t), 8, ids, pos, vel)
        o.advance(120)
        counts.append(o.rebuilds)
        x = o.frame()[1]
        if t == 1:
            x1 = x
        else:
            assert np.abs(x - x1).max() <= 1e-9   # the schedule does not change the physics here
    assert all(c > 0 for c in counts)


def test_particles_outside_the_box_start_inactive_and_leavers_are_deactivated(oracle):
    ids, pos, vel = two_cubes(side=3, speed=40.0, origin=3.0)
    pos = pos.copy()
    pos[0] = [-1.0, 5.0, 5.0]          # outside at start (pm:82-92)
    kw = dict(length_of_area=(30.0, 30.0, 30.0), delta_t=0.002)
    o = oracle.Oracle(oracle.make_params(**kw), 8, ids, pos, vel)
    act0 = o.frame()[3]
    assert act0[0] == 0 and act0[1:].all()
    x_before = o.frame()[1][0].copy()
    o.advance(300)
    _, x, v, act = o.frame()
    assert np.array_equal(x[0], x_before)          # inactive particles are never moved (stoermer:151)
    assert act.sum() < act0.sum()                  # some particles left the 30^3 box
    inside = ((x > 0) & (x <= 30.0)).all(1)
    assert not (act.astype(bool) & ~inside).sum() > act.sum() * 0.2   # leavers linger only until the next rebuild


def test_empty_and_invalid_input(oracle):
    import ctypes as C
    p = oracle.make_params()
    lib = oracle.load()
    assert not lib.psim_oracle_create(C.byref(p), 8, 0, None, None, None, 1)
    with pytest.raises(ValueError):
        oracle.Oracle(p, 2, [1], [[1.0, 1.0, 1.0]], [[0.0, 0.0, 0.0]])


def test_single_particle_and_isolated_pair(oracle):
    p = oracle.make_params()
    o = oracle.Oracle(p, 8, [1], [[10.0, 10.0, 10.0]], [[1.0, 0.0, 0.0]])
    o.advance(10)
    assert o.frame()[1][0, 0] == pytest.approx(10.01, abs=1e-12)
    assert np.all(o.aux()[2] == 0)
    # two particles at r = 1.0 along x: equal and opposite forces of the closed-form magnitude
    o = oracle.Oracle(p, 8, [1, 2], [[11.0, 10.0, 10.0], [10.0, 10.0, 10.0]], np.zeros((2, 3)))
    f = o.aux()[2]
    assert f[0, 0] == pytest.approx(24.0155998, rel=1e-8) and f[1, 0] == -f[0, 0]
    assert np.all(f[:, 1:] == 0)


def _shipped():
    """BASELINE.json configs[0]: the reference's own test_particles.dat + environment_params.nml, carried inside
    the fixture (tests/golden/make_golden_shipped.py; /root/reference does not exist on the GPU box)."""
    return np.load(os.path.join(GOLDEN, "shipped_test_particles_golden.npz"))


def _shipped_params(mod, g, nb_of_threads=None):
    return mod.make_params(delta_t=float(g["delta_t"]), epsilon=float(g["epsilon"]), sigma=float(g["sigma"]),
                           mass=float(g["mass"]), r_cut=float(g["r_cut"]), length_of_area=tuple(g["length_of_area"]),
                           length_of_cell=tuple(g["length_of_cell"]),
                           nb_of_threads=int(g["nb_of_threads"]) if nb_of_threads is None else nb_of_threads)


# f32 tolerances: coordinates up to ~50 carry ~4e-6 of rounding, the force slope at the lattice spacing is ~57
@pytest.mark.parametrize("threads", [3, 1])            # the shipped namelist runs 3 OpenMP threads
@pytest.mark.parametrize("precision,ftol,xtol", [(8, 1e-12, 1e-9), (4, 5e-4, 5e-4)])
def test_oracle_on_the_shipped_case_matches_numpy_golden(oracle, precision, ftol, xtol, threads):
    g = _shipped()
    o = oracle.Oracle(_shipped_params(oracle, g, threads), precision, g["ids"], g["x0"], g["v0"])
    assert o.n_active == len(g["ids"]) == 2000
    f0 = o.aux()[2]
    assert np.abs(f0 - g["f0"]).max() <= ftol * np.abs(g["f0"]).max()
    fb, pairs = oracle.bruteforce(_shipped_params(oracle, g, threads), precision, g["x0"])
    assert pairs in (int(g["pairs0"]), int(g["pairs0"]) // 2)          # ordered pairs in the golden
    assert np.abs(fb - g["f0"]).max() <= ftol * np.abs(g["f0"]).max()
    o.advance(int(g["k_steps"]))                       # one frame interval: the reference's first frame
    _, x, v, a = o.frame()
    assert a.all()
    assert np.abs(x - g["x"]).max() <= xtol
    assert np.abs(v - g["v"]).max() <= 100 * xtol
    fn = o.aux()[2]
    assert np.abs(fn - g["f"]).max() <= (1e-9 if precision == 8 else 2e-2) * np.abs(g["f"]).max()
    assert o.rebuilds >= 1                             # 0.005 per step against dr_max_tolerable = 0.5: lists were rebuilt
    k, u, mom = o.energy()
    assert k == pytest.approx(g["e1"][0], rel=1e-4 if precision == 4 else 1e-10)
    assert np.abs(mom - g["e1"][2:]).max() <= (1e-2 if precision == 4 else 1e-8)
    o.close()


def _shipped_more():
    """particle_files/more_test_particles.dat (N = 39,440) inside its fixture; positions and velocities are stored
    as float32 and re-rounded to the file's 4 decimals (exact, checked by the generator)."""
    g = np.load(os.path.join(GOLDEN, "shipped_more_test_particles_golden.npz"))
    x0 = np.round(g["x0"].astype(np.float64), 4)
    v0 = np.round(g["v0"].astype(np.float64), 4)
    return g, x0, v0


@pytest.mark.parametrize("precision,ftol", [(8, 1e-12), (4, 5e-4)])
def test_oracle_on_the_larger_shipped_case_matches_kdtree_golden(oracle, precision, ftol):
    """Forces at t = 0 on the reference's own N = 39,440 input: the oracle's cell lists (3 slices, as shipped) and its
    internal brute force against a KD-tree + numpy evaluation (tests/golden/make_golden_shipped.py)."""
    g, x0, v0 = _shipped_more()
    p = _shipped_params(oracle, g)
    o = oracle.Oracle(p, precision, g["ids"], x0, v0)
    assert o.n_active == 39440
    scale = np.abs(g["f0"]).max()
    assert np.abs(o.aux()[2] - g["f0"]).max() <= ftol * scale
    fb, pairs = oracle.bruteforce(p, precision, x0)
    assert np.abs(fb - g["f0"]).max() <= ftol * scale
    if precision == 8:
        assert pairs in (int(g["pairs0"]), int(g["pairs0"]) // 2)
    o.close()
