"""GPU test on the reference's own shipped case (BASELINE.json configs[0]); fixture and generator:
tests/golden/shipped_test_particles_golden.npz, tests/golden/make_golden_shipped.py.  In a file of its own that
sorts after the core parity files: it was written after the round's GPU budget was spent and has not run on a GPU yet."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("algo", ["allpairs", "cells"])
def test_shipped_case_first_frame(psim, oracle, precision, algo):
    """BASELINE.json configs[0]: the reference's own particle_files/test_particles.dat with environment_params.nml
    (N = 2,000, NB_OF_THREADS = 3), carried in tests/golden/shipped_test_particles_golden.npz.  State after
    WRITE_TIME_STEP = 100 steps — the reference's first frame — against the oracle in the same precision
    (tolerances of test_k_steps_match_oracle) and, in f64, against the independent numpy golden."""
    g = np.load(os.path.join(GOLDEN, "shipped_test_particles_golden.npz"))
    kw = dict(delta_t=float(g["delta_t"]), epsilon=float(g["epsilon"]), sigma=float(g["sigma"]), mass=float(g["mass"]),
              r_cut=float(g["r_cut"]), length_of_area=tuple(g["length_of_area"]), length_of_cell=tuple(g["length_of_cell"]),
              nb_of_threads=int(g["nb_of_threads"]))
    ids, x0, v0, K = g["ids"], g["x0"], g["v0"], int(g["k_steps"])
    o = oracle.Oracle(oracle.make_params(**kw), precision, ids, x0, v0)
    sim = psim.Simulation(psim.make_params(**kw), ids, x0, v0, precision)
    f0, f0_ref = sim.forces(), o.aux()[2]               # psim_create evaluated them (program:77)
    sim.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED if algo == "cells" else psim.FORCE_ALL_PAIRS)
    ftol = 1e-5 if precision == 4 else 1e-12
    xtol = 2e-4 if precision == 4 else 1e-9
    assert np.abs(f0 - f0_ref).max() <= ftol * np.abs(f0_ref).max()
    assert sim.pairs_in_range() == int(g["pairs0"]) or precision == 4
    o.advance(K)
    sim.advance(K)
    _, x_ref, v_ref, a_ref = o.frame()
    _, x, v, a = sim.frame()
    fn, fn_ref = sim.aux()[2], o.aux()[2]
    assert a.all() and np.array_equal(a, a_ref)
    assert np.abs(x - x_ref).max() <= xtol and np.abs(v - v_ref).max() <= 100 * xtol
    assert np.abs(fn - fn_ref).max() <= 1e3 * ftol * np.abs(fn_ref).max()
    c = sim.counters()
    assert c.steps == K and c.rebuilds == o.rebuilds == 3          # steps 28, 56, 84
    if precision == 8:
        assert np.abs(f0 - g["f0"]).max() <= 1e-12 * np.abs(g["f0"]).max()
        assert np.abs(x - g["x"]).max() <= 1e-9 and np.abs(v - g["v"]).max() <= 1e-7
        assert np.abs(fn - g["f"]).max() <= 1e-9 * np.abs(g["f"]).max()
    sim.close()
    o.close()


@pytest.mark.parametrize("precision", [4, 8])
@pytest.mark.parametrize("algo", ["allpairs", "cells"])
def test_larger_shipped_case_forces_and_short_run(psim, oracle, precision, algo):
    """particle_files/more_test_particles.dat (N = 39,440, two bodies; BASELINE.json configs[1]) from its fixture:
    forces at t = 0 against the KD-tree + numpy golden (f64) and the oracle (same precision), the exact pair
    count, the tensor-core filter's audit, then 30 steps (one rebuild) against the oracle."""
    g = np.load(os.path.join(GOLDEN, "shipped_more_test_particles_golden.npz"))
    x0 = np.round(g["x0"].astype(np.float64), 4)
    v0 = np.round(g["v0"].astype(np.float64), 4)
    kw = dict(delta_t=float(g["delta_t"]), epsilon=float(g["epsilon"]), sigma=float(g["sigma"]), mass=float(g["mass"]),
              r_cut=float(g["r_cut"]), length_of_area=tuple(g["length_of_area"]), length_of_cell=tuple(g["length_of_cell"]),
              nb_of_threads=int(g["nb_of_threads"]))
    ids = g["ids"]
    o = oracle.Oracle(oracle.make_params(**kw), precision, ids, x0, v0)
    sim = psim.Simulation(psim.make_params(**kw), ids, x0, v0, precision)
    ftol = 1e-5 if precision == 4 else 1e-12
    xtol = 2e-4 if precision == 4 else 1e-9
    f0, f0_ref = sim.forces(), o.aux()[2]                  # all-pairs (tensor-core) evaluation of psim_create
    scale = np.abs(f0_ref).max()
    assert np.abs(f0 - f0_ref).max() <= ftol * scale
    exact = sim.pairs_in_range()
    assert sim.counters().filter_pairs == exact            # the filter dropped no pair
    if precision == 8:
        assert exact == int(g["pairs0"])
        assert np.abs(f0 - g["f0"]).max() <= 1e-12 * np.abs(g["f0"]).max()
    sim.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED if algo == "cells" else psim.FORCE_ALL_PAIRS)
    K = 30                                                 # one list rebuild on the way (step 25)
    o.advance(K)
    sim.advance(K)
    _, x_ref, v_ref, a_ref = o.frame()
    _, x, v, a = sim.frame()
    fn, fn_ref = sim.aux()[2], o.aux()[2]
    assert np.array_equal(a, a_ref) and a.all()
    assert np.abs(x - x_ref).max() <= xtol and np.abs(v - v_ref).max() <= 100 * xtol
    assert np.abs(fn - fn_ref).max() <= 1e3 * ftol * np.abs(fn_ref).max()
    c = sim.counters()
    assert c.steps == K and c.rebuilds == o.rebuilds
    sim.close()
    o.close()
