"""Host-side arithmetic of the multi-GPU partition (DESIGN.md "Multi-GPU").

The reference splits the particle index range into contiguous slices, one per
OpenMP thread (program_particle_simulation_openmp.f08:305-323).  Here a slice
belongs to a GPU.  Slices are equal-sized in the PADDED index space (a multiple
of the j-tile per rank) so that the per-step position all-gather moves
equal-sized messages; psim_create_sharded applies exactly this rule
(csrc/psim_api.cu, SimT::init).
"""
from __future__ import annotations

import numpy as np

TILE = 128 * 8   # SimT::T x SimT::MAX_CLUSTER: a rank's j range splits into up to 8 equal tile runs


def padded_size(n: int, world: int, tile: int = TILE) -> int:
    gran = tile * world
    return (n + gran - 1) // gran * gran


def shard_bounds(n: int, rank: int, world: int, tile: int = TILE):
    """Half-open range [i_lo, i_hi) of particle indices rank `rank` owns."""
    shard = padded_size(n, world, tile) // world
    return min(n, shard * rank), min(n, shard * (rank + 1))


def reference_slice(n: int, thread_id: int, num_threads: int):
    """get_particle_index (program_...:305-323), 0-based half-open."""
    per = n // num_threads
    lo = per * thread_id
    hi = n if thread_id == num_threads - 1 else per * (thread_id + 1)
    return lo, hi


def merge_owned(parts, bounds, n: int):
    """Concatenate per-rank arrays (each full-length, valid on its own range)
    in rank order, which is array order (the order write_data emits)."""
    out = np.zeros_like(parts[0])
    for a, (lo, hi) in zip(parts, bounds):
        out[lo:hi] = a[lo:hi]
    assert bounds[0][0] == 0 and bounds[-1][1] == n
    return out


def slice_pairs_of_rank(disp_norm, active, n: int, rank: int, world: int, slices: int):
    """What slice_top2_kernel computes on one rank (csrc/psim_integrate.cuh): for every reference slice the two
    largest displacement norms among the ACTIVE particles of the slice that the rank owns (zeros when it owns
    none).  Returns an array [slices, 2]."""
    lo, hi = shard_bounds(n, rank, world)
    out = np.zeros((slices, 2))
    for k in range(slices):
        a, b = reference_slice(n, k, slices)
        a, b = max(a, lo), min(b, hi)
        if b > a:
            v = np.sort(disp_norm[a:b][active[a:b].astype(bool)])[::-1]
            out[k, :min(2, len(v))] = v[:2]
    return out


def merged_trigger(pairs_by_rank, tolerance: float) -> bool:
    """What housekeeping_kernel decides (csrc/psim_cells.cuh): per slice, the two largest values among all ranks'
    pairs; the rebuild fires when any slice's two largest sum to more than the tolerance
    (module_stoermer_verlet_method.f08:168-185 per thread slice, program_...:71,86,92)."""
    pairs = np.asarray(pairs_by_rank)                       # [world, slices, 2]
    for k in range(pairs.shape[1]):
        v = np.sort(pairs[:, k, :].ravel())[::-1]
        if v[0] + v[1] > tolerance:
            return True
    return False
