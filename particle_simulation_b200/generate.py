"""Seeded restatement of python_src/write_particle_file.py.

The reference script builds a simple-cubic block of particles at the
Lennard-Jones minimum spacing r_min = 2^(1/6) * radius (write_particle_file.py:32,
44, 64-69), gives every particle the same drift velocity plus `vt` times a noise
vector (:46-55, 70) and prints everything with `%.4f` (:73-74).  Its noise is
drawn from an unseeded global RNG (:48); here the generator is explicit so that
benchmark and test inputs are reproducible.
"""
from __future__ import annotations

import numpy as np

DEFAULT_SEED = 12345


def thermal_noise(rng: np.random.Generator, n: int) -> np.ndarray:
    """n noise vectors as thermal_velocity_noise() makes them
    (write_particle_file.py:46-55): three U[0,1) draws, accepted when
    0 < r = sum of squares < 1, scaled by sqrt(-2 ln r / r).  Components are
    therefore never negative, as in the reference."""
    out = np.empty((n, 3))
    filled = 0
    while filled < n:
        m = max(1024, int((n - filled) * 2.2))
        u = rng.uniform(size=(m, 3))
        r = np.sum(u * u, axis=1)
        ok = (r > 0.0) & (r < 1.0)
        u, r = u[ok], r[ok]
        take = min(len(u), n - filled)
        scale = np.sqrt(-2.0 * np.log(r[:take]) / r[:take])
        out[filled:filled + take] = u[:take] * scale[:, None]
        filled += take
    return out


def lattice_block(nb_of_particles, start_point=(0.0, 0.0, 0.0), velocity=(0.0, 0.0, 0.0),
                  thermal=0.0, radius=1.0, rng: np.random.Generator | None = None, seed: int = DEFAULT_SEED):
    """One call of the reference script: returns (ids, pos, vel) with ids
    1..N (x fastest, then y, then z; write_particle_file.py:63-69) and values
    rounded to four decimals like the `%.4f` output."""
    if rng is None:
        rng = np.random.Generator(np.random.PCG64(seed))
    nb = np.asarray(nb_of_particles, dtype=np.int64)
    r_min = radius * 2.0 ** (1.0 / 6.0)
    start = np.asarray(start_point, dtype=np.float64) - r_min * 0.5 * nb     # :44
    z, y, x = np.meshgrid(np.arange(nb[2]), np.arange(nb[1]), np.arange(nb[0]), indexing="ij")
    grid = np.stack([x.ravel(), y.ravel(), z.ravel()], axis=1).astype(np.float64)
    pos = start[None, :] + r_min * grid                                      # :69
    n = pos.shape[0]
    vel = np.asarray(velocity, dtype=np.float64)[None, :] + thermal * thermal_noise(rng, n)   # :70
    ids = np.arange(1, n + 1, dtype=np.int32)
    return ids, np.round(pos, 4), np.round(vel, 4)


# shapes, boxes and steps of the synthetic configurations.
SYNTHETIC = {
    65536: dict(nb=(64, 32, 32), area=(99.0, 99.0, 99.0), steps=1000),
    262144: dict(nb=(64, 64, 64), area=(99.0, 99.0, 99.0), steps=1000),
    1048576: dict(nb=(128, 128, 64), area=(198.0, 198.0, 99.0), steps=100),
}


def synthetic(n: int, thermal: float = 0.1, seed: int = DEFAULT_SEED):
    """(ids, pos, vel, length_of_area) of a synthetic configuration; the block
    is centred in its box.  Sizes outside SYNTHETIC get a near-cubic block in
    the smallest cubic box of 3.0-wide cells that holds it."""
    if n in SYNTHETIC:
        cfg = SYNTHETIC[n]
        nb, area = cfg["nb"], cfg["area"]
    else:
        side = int(round(n ** (1.0 / 3.0)))
        while side ** 3 < n:
            side += 1
        nb = (side, side, side)
        extent = side * 2.0 ** (1.0 / 6.0)
        cells = int(np.ceil((extent + 6.0) / 3.0))
        area = (3.0 * cells,) * 3
    centre = tuple(0.5 * a for a in area)
    ids, pos, vel = lattice_block(nb, start_point=centre, thermal=thermal, seed=seed)
    if len(ids) > n:
        ids, pos, vel = ids[:n], pos[:n], vel[:n]
    return ids, pos, vel, area
