"""particle_simulation_b200 — B200-native Lennard-Jones / Stoermer-Verlet hot path.

A drop-in for one path of RobertOpitz/Particle_simulation: the pairwise force
evaluation and velocity-Verlet update of
fortran_src/module_stoermer_verlet_method.f08, driven as in
fortran_src/program_particle_simulation_openmp.f08.  The product is the CUDA
library behind include/psim.h (csrc/); this package is the thin Python host.
"""
from . import formats, generate, native, sharding  # noqa: F401
from .native import (  # noqa: F401
    ALLPAIRS_CUDA_CORE, ALLPAIRS_DIFFERENCE_FORM, ALLPAIRS_TENSOR, FORCE_ALL_PAIRS, FORCE_CELL_BINNED,
    OPT_ALLPAIRS_KERNEL, OPT_EXACT_DIFF_FILTER, OPT_FORCE_ALGO, OPT_FUSED_STEP, OPT_KERNEL_TIMING, Params, PsimError, Simulation, compute_forces, make_params,
)

__version__ = "0.1.0"
