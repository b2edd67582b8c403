"""Parity at the sizes that are benchmarked (BASELINE.json configs[2] and [3]: the synthetic lattice
blocks N = 65,536 and N = 262,144 of SURVEY.md 8(d)), CUDA through the C ABI against the CPU oracle.

The oracle here is the restatement of the reference's own cell-list algorithm (oracle/psim_oracle.cpp,
single thread: the reference's summation order), which is O(N) and therefore finishes in well under a
second per force evaluation at these sizes; the O(N k) brute-force sweep is added at N = 65,536 where it
takes one second.  Every force kernel that can run the job is compared: the tensor-core all-pairs kernel
(default), the CUDA-core all-pairs kernel and the cell-binned path.

Tolerances as everywhere: one force evaluation differs from the oracle by summation order
only, f64 <= 1e-12 max|f|, f32 <= 1e-5 max|f|; K-step positions f64 <= 1e-9, f32 <= 2e-4.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FTOL = {4: 1e-5, 8: 1e-12}
XTOL = {4: 2e-4, 8: 1e-9}
K_STEPS = 20

_cache = {}


def _case(oracle, n, precision):
    """Inputs and oracle answers of one (n, precision), computed once per session."""
    key = (n, precision)
    if key not in _cache:
        from particle_simulation_b200 import generate as G
        ids, pos, vel, area = G.synthetic(n)
        o = oracle.Oracle(oracle.make_params(length_of_area=area), precision, ids, pos, vel)
        f0 = o.aux()[2].copy()
        o.advance(K_STEPS)
        _, x, v, a = o.frame()
        fn = o.aux()[2].copy()
        rebuilds = o.rebuilds
        o.close()
        pairs = None
        if n <= 65536:
            bf, pairs = oracle.bruteforce(oracle.make_params(length_of_area=area), precision, pos)
            scale = np.abs(bf).max()
            assert np.abs(f0 - bf).max() <= 4 * FTOL[precision] * scale      # the two oracles agree with each other
        _cache[key] = dict(ids=ids, pos=pos, vel=vel, area=area, f0=f0, x=x.copy(), v=v.copy(), a=a.copy(), fn=fn,
                           rebuilds=rebuilds, pairs=pairs)
    return _cache[key]


def _select(psim, sim, kernel):
    if kernel == "cells":
        sim.set_option(psim.OPT_FORCE_ALGO, psim.FORCE_CELL_BINNED)
    elif kernel == "allpairs_cuda_core":
        sim.set_option(psim.OPT_ALLPAIRS_KERNEL, psim.ALLPAIRS_CUDA_CORE)


@pytest.mark.parametrize("precision", [8, 4])
@pytest.mark.parametrize("kernel", ["allpairs", "allpairs_cuda_core", "cells"])
@pytest.mark.parametrize("n", [65536, 262144])
def test_forces_and_k_steps_at_bench_sizes_match_oracle(psim, oracle, n, kernel, precision):
    c = _case(oracle, n, precision)
    sim = psim.Simulation(psim.make_params(length_of_area=c["area"]), c["ids"], c["pos"], c["vel"], precision)
    _select(psim, sim, kernel)
    sim.set_positions(c["pos"])
    f = sim.forces()
    scale = np.abs(c["f0"]).max()
    assert scale > 0
    assert np.abs(f - c["f0"]).max() <= FTOL[precision] * scale, "initial forces differ from the oracle"
    exact = sim.pairs_in_range()
    if c["pairs"] is not None:
        assert exact == c["pairs"]
    if kernel == "allpairs":      # the tensor-core filter evaluated exactly the in-range pairs: it dropped none
        assert sim.counters().filter_pairs == exact
    sim.advance(K_STEPS)
    _, x, v, a = sim.frame()
    fn = sim.aux()[2]
    assert np.array_equal(a, c["a"])
    assert sim.counters().rebuilds == c["rebuilds"]
    assert np.abs(x - c["x"]).max() <= XTOL[precision]
    assert np.abs(v - c["v"]).max() <= 100 * XTOL[precision]
    assert np.abs(fn - c["fn"]).max() <= 1e3 * FTOL[precision] * np.abs(c["fn"]).max()
    if kernel == "allpairs":
        assert sim.counters().filter_pairs == sim.pairs_in_range()
    sim.close()


def test_repeat_evaluation_is_bit_identical_at_bench_size(psim):
    """Fixed summation order (also across the j segments of the tensor-core kernel, which N = 65,536 on one
    GPU uses): the same positions give the same bits, launch after launch."""
    from particle_simulation_b200 import generate as G
    ids, pos, vel, area = G.synthetic(65536)
    sim = psim.Simulation(psim.make_params(length_of_area=area), ids, pos, vel, 8)
    f0 = sim.forces().copy()
    for _ in range(5):
        sim.set_positions(pos)
        assert np.array_equal(sim.forces(), f0)
    sim.close()
