"""GPU parity tests proper: the CUDA path, called through the C ABI
(particle_simulation_b200.native -> libpsim_b200.so), against the CPU oracle on
the same seeded inputs.

Tolerances.  A pair's contribution is computed with the same
correctly rounded operations on both sides, so one force evaluation differs by
summation order only:
  f64: |f_gpu - f_oracle| <= 1e-12 * max|f|   f32: <= 1e-5 * max|f|
Trajectories are chaotic; K-step comparisons use short horizons and
  f64: 1e-9 absolute on positions,  f32: 2e-4.
"""
import os

import numpy as np
import pytest

from conftest import two_cubes

pytestmark = pytest.mark.gpu

FTOL = {4: 1e-5, 8: 1e-12}
XTOL = {4: 2e-4, 8: 1e-9}


def _ref_files():
    import os
    from conftest import GOLDEN
    return os.path.join(GOLDEN, "two_cubes_432.dat")


ALGOS = ["allpairs", "allpairs_cuda_core", "cells"]

KERNEL = {"allpairs": 0, "allpairs_difference_form": 1, "allpairs_cuda_core": 5,
          "allpairs_round1_alias_3": 3, "allpairs_round1_alias_6": 6}


def _select(psim, sim, algo):
    sim.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED if algo == "cells" else psim.FORCE_ALL_PAIRS)
    if algo in KERNEL:
        sim.set_option(psim.OPT_ALLPAIRS_KERNEL, KERNEL[algo])


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("variant", ["allpairs", "allpairs_difference_form", "allpairs_cuda_core",
                                     "allpairs_round1_alias_3", "allpairs_round1_alias_6", "cells"])
def test_initial_forces_match_oracle(psim, oracle, precision, variant):
    ids, pos, vel = two_cubes()
    p = psim.make_params()
    op = oracle.make_params()
    o = oracle.Oracle(op, precision, ids, pos, vel)
    _, _, f_ref = o.aux()
    sim = psim.Simulation(p, ids, pos, vel, precision)
    _select(psim, sim, variant)
    sim.set_positions(pos)
    f = sim.forces()
    scale = np.abs(f_ref).max()
    assert scale > 0
    assert np.abs(f - f_ref).max() <= FTOL[precision] * scale
    # and against the independent brute-force double evaluation
    bf, pairs = oracle.bruteforce(op, precision, pos)
    assert np.abs(f - bf).max() <= 4 * FTOL[precision] * scale
    assert sim.pairs_in_range() == pairs
    if variant in ("allpairs", "allpairs_round1_alias_3", "allpairs_round1_alias_6"):
        assert sim.counters().filter_pairs == pairs          # the tensor-core filter dropped no pair
    sim.close()


def test_removed_kernel_values_are_refused(psim):
    ids, pos, vel = two_cubes(side=3)
    sim = psim.Simulation(psim.make_params(), ids, pos, vel, 8)
    for value in (2, 4, 7, -1):
        with pytest.raises(psim.PsimError) as e:
            sim.set_option(psim.OPT_ALLPAIRS_KERNEL, value)
        assert e.value.status == 1
    sim.close()


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("threads", [1, 3])
def test_k_steps_match_oracle(psim, oracle, precision, fused, algo, threads):
    """threads = NB_OF_THREADS: the shipped namelist says 3, which changes the
    rebuild trigger (evaluated per thread slice, program:71,86,92)."""
    ids, pos, vel = two_cubes()
    p = psim.make_params(nb_of_threads=threads)
    o = oracle.Oracle(oracle.make_params(nb_of_threads=threads), precision, ids, pos, vel)
    sim = psim.Simulation(p, ids, pos, vel, precision)
    sim.set_option(psim.OPT_FUSED_STEP, fused)
    _select(psim, sim, algo)
    K = 60
    for chunk in (1, 2, K - 3):
        o.advance(chunk)
        sim.advance(chunk)
    _, x_ref, v_ref, a_ref = o.frame()
    _, x, v, a = sim.frame()
    d_ref, fo_ref, fn_ref = o.aux()
    d, fo, fn = sim.aux()
    assert np.array_equal(a, a_ref)
    assert np.abs(x - x_ref).max() <= XTOL[precision]
    assert np.abs(v - v_ref).max() <= XTOL[precision] * 100
    fs = np.abs(fn_ref).max()
    assert np.abs(fn - fn_ref).max() <= 1e3 * FTOL[precision] * fs
    assert np.abs(fo - fo_ref).max() <= 1e3 * FTOL[precision] * fs
    assert np.abs(d - d_ref).max() <= XTOL[precision]
    c = sim.counters()
    assert c.steps == K and c.rebuilds == o.rebuilds
    sim.close()


@pytest.mark.parametrize("algo", ["allpairs", "cells"])
def test_rebuild_schedule_depends_on_nb_of_threads_like_the_reference(psim, oracle, algo):
    """Two fast particles at the two ends of the array: with one trigger slice their displacements add up and
    the rebuild fires at step 9; with NB_OF_THREADS = 3 they sit in different slices (program:305-323) and it
    fires at step 13.  The device must follow the oracle step by step in both cases."""
    ids, pos, vel = two_cubes()
    vel = vel.copy()
    vel[0] = (-30.0, 0.0, 0.0)
    vel[-1] = (30.0, 0.0, 0.0)
    first = {}
    for threads in (1, 3):
        o = oracle.Oracle(oracle.make_params(nb_of_threads=threads), 8, ids, pos, vel)
        sim = psim.Simulation(psim.make_params(nb_of_threads=threads), ids, pos, vel, 8)
        _select(psim, sim, algo)
        for step in range(1, 31):
            o.advance(1)
            sim.advance(1)
            assert sim.counters().rebuilds == o.rebuilds, "step %d, NB_OF_THREADS %d" % (step, threads)
            if o.rebuilds and threads not in first:
                first[threads] = step
        _, x_ref, _, a_ref = o.frame()
        _, x, _, a = sim.frame()
        assert np.array_equal(a, a_ref) and np.abs(x - x_ref).max() <= XTOL[8]
        sim.close()
        o.close()
    assert first[1] < first[3], "test lost its point: the schedules should differ (%r)" % (first,)


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("algo", ALGOS)
def test_particles_leaving_the_box_are_deactivated_like_the_reference(psim, oracle, precision, algo):
    # a small box: the fast cube leaves it during the run
    ids, pos, vel = two_cubes(side=4, speed=40.0, origin=3.0)
    kw = dict(length_of_area=(30.0, 30.0, 30.0), delta_t=0.002)
    o = oracle.Oracle(oracle.make_params(**kw), precision, ids, pos, vel)
    sim = psim.Simulation(psim.make_params(**kw), ids, pos, vel, precision)
    _select(psim, sim, algo)
    total = 0
    for _ in range(12):
        o.advance(25)
        sim.advance(25)
        total += 25
        _, x_ref, v_ref, a_ref = o.frame()
        _, x, v, a = sim.frame()
        # the collision is violent and chaotic: in f32, summation-order noise reaches the size of a
        # step after ~50 steps, after which WHEN a particle crosses the box edge may differ too
        if precision == 8 or total <= 50:
            assert np.array_equal(a, a_ref), "active flags differ after %d steps" % total
            assert sim.counters().rebuilds == o.rebuilds
        keep = (a_ref & a).astype(bool)
        # the collision is violent: summation-order noise grows exponentially, so positions are
        # compared over the first 100 steps only (f32) / throughout (f64); flags and rebuild
        # counts are compared throughout
        if keep.any() and (precision == 8 or total <= 50):
            assert np.abs(x[keep] - x_ref[keep]).max() <= 50 * XTOL[precision]
        # size-independent check that does not suffer from chaos: the force the stepping loop
        # holds (computed with possibly stale bins) is the brute-force force at ITS OWN positions
        f_now = sim.aux()[2]
        bf, _ = oracle.bruteforce(oracle.make_params(**kw), precision, x, a)
        scale = max(np.abs(bf).max(), 1e-30)
        assert np.abs(f_now[a.astype(bool)] - bf[a.astype(bool)]).max() <= 4 * FTOL[precision] * scale
    assert (a_ref == 0).any(), "test is vacuous: nothing left the box"
    sim.close()


@pytest.mark.parametrize("algo", ALGOS)
def test_conservation_and_energy(psim, oracle, algo):
    ids, pos, vel = two_cubes()
    sim = psim.Simulation(psim.make_params(), ids, pos, vel, 8)
    _select(psim, sim, algo)
    k0, u0, p0 = sim.energy()
    o = oracle.Oracle(oracle.make_params(), 8, ids, pos, vel)
    ko, uo, po = o.energy()
    assert abs(k0 - ko) <= 1e-9 * abs(ko) and abs(u0 - uo) <= 1e-9 * abs(uo)
    sim.advance(400)
    o.advance(400)
    k1, u1, p1 = sim.energy()
    ko1, uo1, _ = o.energy()
    assert np.abs(p1 - p0).max() <= 1e-9 * np.abs(p0).max()          # momentum: pairwise antisymmetric forces
    assert abs((k1 + u1) - (k0 + u0)) <= 2e-5 * abs(k0 + u0)          # velocity Verlet, dt = 1e-3
    # "total-energy drift matching" (north_star): the drift over the same 400 steps equals the oracle's
    drift, drift_o = (k1 + u1) - (k0 + u0), (ko1 + uo1) - (ko + uo)
    assert abs(drift - drift_o) <= 1e-9 * abs(ko + uo)
    sim.close()
    o.close()


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("case", ["lattice_65536", "hot_gas_32768", "two_cubes", "corner_of_big_box", "dense_row_chunks"])
def test_tensor_core_filter_drops_no_pair(psim, oracle, precision, case):
    """The tensor-core distance filter (split-FP16 Gram matrix) must be conservative: the ordered in-range
    pairs its exact path evaluated (psim_counters.filter_pairs) equal the exact O(N^2) count, and stay equal
    while the configuration moves; the forces equal the CPU oracle's brute-force evaluation at the same
    positions.  corner_of_big_box puts the block in the far corner of the 198 x 198 x 99 box of the
    N = 1,048,576 configuration, where the filter's margins (which grow with the distance from the box
    centre) are widest."""
    from particle_simulation_b200 import generate as G
    kw = {}
    if case == "two_cubes":
        ids, pos, vel = two_cubes()
    elif case == "lattice_65536":
        ids, pos, vel, area = G.synthetic(65536)
        kw = dict(length_of_area=area)
    elif case == "corner_of_big_box":
        ids, pos, vel = G.lattice_block((32, 32, 16), (177.0, 177.0, 88.0), thermal=0.1, seed=5)
        kw = dict(length_of_area=(198.0, 198.0, 99.0))
    elif case == "dense_row_chunks":
        # a compressed block (radius 0.85: ~75 neighbours in range) whose lattice order puts 7 to 9 of a row's
        # neighbours into 64 consecutive indices: the kernel's queue turns such a chunk into ONE "dense" entry and
        # its exact path walks all 64 columns (a branch the other cases never take); a shuffled copy of the same
        # block spreads the neighbours over all chunks instead (sparse entries, long queues, mid-block flushes)
        ids, pos, vel = G.lattice_block((16, 16, 16), (49.5, 49.5, 49.5), radius=0.85, seed=3)
        rng = np.random.default_rng(17)
        pos = pos + rng.uniform(-0.02, 0.02, pos.shape)
        perm = rng.permutation(len(ids))
        ids = np.concatenate([ids, ids + len(ids)]).astype(np.int32)
        pos = np.concatenate([pos, pos[perm] + np.array([30.0, 0.0, 0.0])])
        vel = np.concatenate([vel, vel[perm]]) * 0.0
        kw = dict(delta_t=1e-7)
    else:   # strongly perturbed lattice: many pairs close to the cut-off on both sides
        ids, pos, vel, area = G.synthetic(32768, thermal=0.1)
        rng = np.random.default_rng(11)
        pos = pos + rng.uniform(-0.45, 0.45, pos.shape)
        vel = vel * 0.0
        kw = dict(length_of_area=area, delta_t=1e-5)
    sim = psim.Simulation(psim.make_params(**kw), ids, pos, vel, precision)   # default: the tensor-core kernel
    for round_ in range(3):
        exact = sim.pairs_in_range()
        assert exact > 0
        assert sim.counters().filter_pairs == exact
        _, x, _, a = sim.frame()
        bf, pairs = oracle.bruteforce(oracle.make_params(**kw), precision, x, a)
        assert pairs == exact
        f = sim.forces()
        assert np.abs(f - bf).max() <= 4 * FTOL[precision] * np.abs(bf).max()
        sim.advance(7)
    sim.close()
