import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


GOLDEN = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure; parity unpinned, see oracle/psim_oracle.h)."""
    from oracle import oracle_py

    oracle_py.load()
    return oracle_py


@pytest.fixture(scope="session")
def psim():
    import particle_simulation_b200 as P

    return P


def two_cubes(side=6, gap=4.0, speed=5.0, seed=7, origin=10.0):
    """A scaled-down test_particles.dat: two lattice cubes at the LJ minimum
    spacing flying at each other."""
    from particle_simulation_b200 import generate as G

    rng = np.random.Generator(np.random.PCG64(seed))
    ext = side * 2.0 ** (1.0 / 6.0)
    c0 = origin + 0.5 * ext
    ids_a, pos_a, vel_a = G.lattice_block((side,) * 3, (c0,) * 3, (speed,) * 3, 0.1, rng=rng)
    c1 = c0 + (ext + gap) / np.sqrt(3.0) * 1.0 + ext * 0.0
    c1 = origin + ext + gap + 0.5 * ext
    ids_b, pos_b, vel_b = G.lattice_block((side,) * 3, (c1,) * 3, (-speed,) * 3, 0.1, rng=rng)
    return (np.concatenate([ids_a, ids_b]), np.concatenate([pos_a, pos_b]), np.concatenate([vel_a, vel_b]))
