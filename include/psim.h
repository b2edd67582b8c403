/*
 * psim.h — C ABI of the B200-native Lennard-Jones / Stoermer-Verlet hot path.
 *
 * The reference (RobertOpitz/Particle_simulation) has no FFI or plugin interface
 * of its own: its integrator module takes assumed-shape arrays of a
 * non-interoperable derived type (fortran_src/module_stoermer_verlet_method.f08:16-27,
 * :87-88, :136-137, :191-192).  This header is therefore the interface a
 * maintainer binds with ISO_C_BINDING (see INTEGRATION.md and
 * particle_simulation_b200/fortran/psim_binding.f08).  Each entry point names
 * the reference procedure(s) it stands in for.
 *
 * Conventions
 *   - every function returns a psim_status (0 = success) and never aborts;
 *   - all pointers are plain host pointers unless a name ends in _dev;
 *   - `precision` is the byte width of the real type, 4 (the reference's
 *     `rt => real32`, module_stoermer_verlet_method.f08:5) or 8; real arrays are
 *     passed as `const void*` / `void*` and hold float or double accordingly;
 *   - vectors are packed xyz, particle-major: pos[3*i + 0..2];
 *   - one host thread per handle; the library owns all device memory and streams;
 *   - there is no CPU fallback: without a CUDA device every call that would
 *     compute returns PSIM_ERR_NO_DEVICE.
 */
#ifndef PSIM_H
#define PSIM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum psim_status {
  PSIM_OK = 0,
  PSIM_ERR_INVALID_ARGUMENT = 1,
  PSIM_ERR_NO_DEVICE = 2,
  PSIM_ERR_CUDA = 3,
  PSIM_ERR_NO_PARTICLES = 4,   /* "No particles in particle file", program_particle_simulation_openmp.f08:269-270 */
  PSIM_ERR_UNSUPPORTED = 5,
  PSIM_ERR_COMM = 6,
  PSIM_ERR_IO = 7
} psim_status;

/* The ten namelist keys of ENVIRONMENTPARAMS
 * (program_particle_simulation_openmp.f08:198-200; environment_params.nml:2-11)
 * minus the two that only drive the time loop (write/max time step are
 * arguments of psim_run).  Values are held in double and rounded once to the
 * working precision, then the derived constants are formed in that precision
 * in the reference's expression order (module_stoermer_verlet_method.f08:68-81). */
typedef struct psim_params {
  double delta_t;            /* DELTA_T */
  double epsilon;            /* EPSILON  -> input_epsilon_value (program_...:219-220) */
  double sigma;              /* SIGMA    -> particle_radius */
  double mass;               /* MASS     -> particle_mass */
  double r_cut_factor;       /* R_CUT    -> r_cut_factor (r_cut = R_CUT * SIGMA) */
  double length_of_area[3];  /* LENGTH_OF_AREA */
  double length_of_cell[3];  /* LENGTH_OF_CELL */
  int32_t nb_of_threads;     /* NB_OF_THREADS: the reference evaluates its
                                rebuild trigger per thread slice
                                (program_...:71,86,92; slices :305-323); the
                                device reproduces that per-slice trigger, on
                                single-GPU and on sharded handles.  <=0 means 1. */
  int32_t reserved;
} psim_params;

/* Force evaluation strategy (all give the same forces up to summation order). */
typedef enum psim_force_algo {
  PSIM_FORCE_ALL_PAIRS = 0,  /* tiled all-pairs scan, N(N-1) ordered pair tests per step */
  PSIM_FORCE_CELL_BINNED = 1 /* particles binned by LENGTH_OF_CELL, 27-cell neighbourhood
                                (what module_particle_management.f08:52-108,179-339 achieves with lists);
                                single-GPU and sharded handles */
} psim_force_algo;

/* Which kernel runs the all-pairs force pass.  All of them evaluate the same pairs with the same arithmetic
 * (one rounding per operation in the order of stoermer:117-128); results differ by summation order only. */
typedef enum psim_allpairs_kernel {
  PSIM_ALLPAIRS_TENSOR = 0,          /* default: the distance filter of every ordered pair runs on the tensor cores
                                        (one tcgen05.mma kind::f16 per 128 x 128 pairs on split FP16 operands, FP16
                                        accumulators in TMEM, TMA-fed, packed TMEM loads and a sign test), the exact
                                        pair arithmetic on the CUDA cores for the few pairs that pass.  The round-1
                                        values 3 and 6 are accepted as aliases. */
  PSIM_ALLPAIRS_DIFFERENCE_FORM = 1, /* the distance test in difference form in the working precision (3 sub, 1 mul,
                                        2 FMA, 1 compare per pair: the literal reading of stoermer:117-121), no
                                        pre-filter; ~5x slower, kept as the plainest statement of the path */
  PSIM_ALLPAIRS_CUDA_CORE = 5        /* warp-autonomous CUDA-core kernel with the norm-expansion pre-filter in packed
                                        FP32 (3 FMAs per ordered pair); no tensor cores; ~3x slower than 0 */
} psim_allpairs_kernel;

typedef enum psim_option {
  PSIM_OPT_FORCE_ALGO = 0,      /* psim_force_algo */
  PSIM_OPT_FUSED_STEP = 1,      /* 1: force+kick+drift in one kernel (default), 0: separate kernels */
  PSIM_OPT_KERNEL_TIMING = 2,   /* 1: bracket every force-kernel launch with CUDA events */
  PSIM_OPT_ALLPAIRS_KERNEL = 3, /* psim_allpairs_kernel */
  PSIM_OPT_EXACT_DIFF_FILTER = 3 /* round-1 name of PSIM_OPT_ALLPAIRS_KERNEL (same value; kept for callers built against it) */
} psim_option;

typedef struct psim_counters {
  int64_t steps;            /* time steps advanced since create */
  int64_t rebuilds;         /* times the reference's rebuild trigger fired */
  int64_t kernel_launches;  /* kernels of this library launched since create */
  int64_t force_launches;   /* of which force kernels */
  double  force_ms;         /* summed device time of timed force launches (PSIM_OPT_KERNEL_TIMING) */
  int64_t force_timed;      /* number of force launches that were timed */
  int64_t n_active;         /* active particles after the last completed step */
  int64_t filter_pairs;     /* ordered in-range pairs (i, j) of this handle's i-slice that the last force launch
                             * sent through its exact path; tracked by the tensor-core all-pairs kernel only
                             * (-1 otherwise).  Equal to psim_count_pairs_in_range at the same positions iff
                             * the distance filter dropped no pair: the conservativeness check of the tests. */
  int64_t filter_tests;     /* ordered pair distance tests of the last force launch over this handle's i-slice:
                             * (owned i) x (n - 1) for an all-pairs kernel; for the cell-binned path the tests its
                             * 27-cell neighbourhoods needed, counted only while PSIM_OPT_KERNEL_TIMING is on
                             * (-1 otherwise).  bench.py's roofline accounting uses it. */
  int64_t peer_push_steps;  /* sharded handles: steps whose positions reached the other ranks as peer-memory stores
                             * from inside the fused force kernel (NVLink) instead of an NCCL all-gather; 0 when the
                             * ranks could not map each other's buffers (CUDA IPC), with PSIM_GATHER=nccl in the
                             * environment, and on single-GPU handles */
} psim_counters;

typedef struct psim_sim psim_sim;   /* opaque */

/* --- life cycle --------------------------------------------------------- */

/* Replaces: set_simulation_parameter (module_stoermer_verlet_method.f08:52-83),
 * the particle-array fill (program_...:226-286), initialize_neighbor_lists'
 * observable effect (particles outside the box start inactive,
 * module_particle_management.f08:82-92, 534-552) and the initial
 * compute_forces call (program_...:77).
 * id[n], pos[3n], vel[3n] are caller-owned host arrays and are copied.
 * device < 0 selects the current CUDA device. */
psim_status psim_create(const psim_params* params, int precision, int64_t n,
                        const int32_t* id, const void* pos, const void* vel,
                        int device, psim_sim** out);

/* Multi-GPU create: this process owns the contiguous i-slice `rank` of
 * `world` (the reference's thread slices, program_...:305-323, one per GPU).
 * All ranks pass the full particle arrays.  nccl_unique_id is the 128-byte
 * ncclUniqueId produced by psim_comm_unique_id on rank 0 and distributed by the
 * host (torch.distributed, MPI, a file ...).  Every sharded handle owns its own
 * communicator: use a fresh id per handle.
 * When the ranks are processes of one node whose GPUs can address each other, create also maps every rank's
 * position buffers into the others (CUDA IPC): the fused force kernel then stores x_{n+1} of its particles into
 * all of them as it finishes each block, and the per-step NCCL all-gather shrinks to one flag per rank
 * (psim_counters.peer_push_steps counts such steps).  PSIM_GATHER=nccl in the environment keeps the all-gather;
 * it is also what happens when any rank cannot map a peer.  All ranks must then call psim_advance with the
 * same step counts, as they must for the collective anyway. */
psim_status psim_create_sharded(const psim_params* params, int precision, int64_t n,
                                const int32_t* id, const void* pos, const void* vel,
                                int device, int rank, int world,
                                const void* nccl_unique_id, psim_sim** out);
psim_status psim_comm_unique_id(void* out128);

/* Replaces: clean_up (module_particle_management.f08:556-598). */
psim_status psim_destroy(psim_sim* sim);

psim_status psim_set_option(psim_sim* sim, int option, int64_t value);

/* --- the time loop ------------------------------------------------------ */

/* Replaces one or more trips of the time_integration loop
 * (program_...:80-117): compute_locations -> rebuild side effects ->
 * compute_forces -> compute_velocities, nsteps times.  Asynchronous: returns
 * once the work is enqueued.  psim_synchronize waits. */
psim_status psim_advance(psim_sim* sim, int64_t nsteps);
psim_status psim_synchronize(psim_sim* sim);

/* As psim_advance but bracketed by CUDA events on the library's own stream;
 * returns the device time of the nsteps in milliseconds (synchronous). */
psim_status psim_advance_timed(psim_sim* sim, int64_t nsteps, double* elapsed_ms);

/* --- reading state back -------------------------------------------------- */

/* Replaces the data gathered by write_data (program_...:327-365): id, location,
 * velocity and the active flag of all n particles in array order (the host
 * filters inactive ones, as write_data does at :342).  Any pointer may be NULL.
 * Synchronous. */
psim_status psim_get_frame(psim_sim* sim, int32_t* id, void* pos, void* vel,
                           uint8_t* active);

/* Test/diagnostic access to the rest of particle_type
 * (module_stoermer_verlet_method.f08:19-23): displacement, force_old, force_new. */
psim_status psim_get_aux(psim_sim* sim, void* displacement, void* force_old,
                         void* force_new);

psim_status psim_get_counters(psim_sim* sim, psim_counters* out);

/* The half-open range of particle indices this handle owns (0..n for a
 * single-GPU handle).  In a sharded run psim_get_frame / psim_get_aux /
 * snapshots fill velocity, force and displacement for this range only
 * (positions and active flags are complete on every rank); the host
 * concatenates ranks in order, which is array order. */
psim_status psim_get_shard(psim_sim* sim, int64_t* i_lo, int64_t* i_hi);

/* Replace all positions (host array pos[3n]), redo the start-up activation
 * (or take `active`, if not NULL), zero displacement and forces, and evaluate
 * the forces at the new positions, i.e. compute_forces as at
 * program_...:77.  With psim_get_aux(force_new) this is the force operator
 * on a resident handle: no allocation per call. */
psim_status psim_set_positions(psim_sim* sim, const void* pos, const uint8_t* active);

/* Frame ring for the output cadence (program_...:109-115): snapshot the
 * current (post-kick) state into pinned slot `slot` (0..PSIM_FRAME_SLOTS-1) by
 * async D2H on a side stream, without stalling the stepping stream; then wait
 * for a slot and get pointers into the pinned buffers (valid until the slot is
 * snapshotted again).  pos/vel point at `precision`-byte reals, packed xyz.
 * psim_snapshot_wait is the one entry point that may be called from a second
 * host thread while the owning thread keeps calling psim_advance /
 * psim_snapshot_async on OTHER slots (simu_b200's writer thread does): it only
 * reads the slot's event and pinned buffers.  The hand-over of the slot number
 * between the threads must itself be synchronised by the host. */
#define PSIM_FRAME_SLOTS 4
psim_status psim_snapshot_async(psim_sim* sim, int slot);
psim_status psim_snapshot_wait(psim_sim* sim, int slot, const int32_t** id,
                               const void** pos, const void** vel,
                               const uint8_t** active);

/* --- the force operator on its own --------------------------------------- */

/* Replaces: compute_forces + compute_force_between_two_particles
 * (module_stoermer_verlet_method.f08:87-132) as a stateless operator over host
 * arrays: force[3n] = sum over active j != i with r < r_cut.  active may be
 * NULL (all active).  Includes the H2D and D2H copies.  pairs_in_range (may be
 * NULL) receives the number of ordered pairs inside the cutoff. */
psim_status psim_compute_forces_host(const psim_params* params, int precision,
                                     int64_t n, const void* pos,
                                     const uint8_t* active, int algo, int device,
                                     void* force, int64_t* pairs_in_range);

/* Ordered pairs (i, j != i), both active, with r < r_cut at the current
 * positions (P_in).  Synchronous. */
psim_status psim_count_pairs_in_range(psim_sim* sim, int64_t* pairs);

/* Total kinetic and potential energy and linear momentum of the active
 * particles, accumulated in double.  The reference computes neither; the
 * potential is the one whose gradient is its force. */
psim_status psim_energy(psim_sim* sim, double* kinetic, double* potential,
                        double momentum[3]);

/* --- for callers that cannot allocate device memory ----------------------- */

/* No counterpart in the reference (there, every array is ordinary host
 * memory).  This is the drift step, compute_locations, offered to hosts that
 * will not let a library allocate on the GPU: a driver that has committed the
 * device to its own memory plan (OpenACC / CUDA Fortran arrays already
 * resident, a fixed per-process memory budget), or a stepping loop that cannot
 * afford the device-wide synchronisation of cudaMalloc / cudaFree.  It works in
 * place on packed xyz DEVICE arrays the caller owns — pos, force_old,
 * force_new, displacement are updated, vel and active (may be NULL = all
 * active) are read — and writes the rebuild verdict
 * (dr_max_first + dr_max_second > dr_max_tolerable) as 0/1 to the device int
 * *rebuild_flag_dev.  Asynchronous on `cuda_stream` (a cudaStream_t, NULL = the
 * default stream); the CUDA device that is current at the call is used.
 * The n particles of a call form ONE trigger slice, whatever
 * params->nb_of_threads says: a host that wants the reference's per-thread
 * verdicts (NB_OF_THREADS > 1, program_...:71,86,92) calls it once per slice on
 * the sub-arrays, one call after the other, and ORs the flags.
 *
 * Its reduction scratch is a fixed array inside the library image, not an
 * allocation: PSIM_STATIC_SCRATCH_BYTES per GPU that loads the library, held
 * until the process exits.  Consequences: n may not exceed
 * PSIM_STATIC_MAX_PARTICLES (PSIM_ERR_INVALID_ARGUMENT otherwise, nothing is
 * launched); all calls on a GPU share the array, so calls that may overlap on
 * the GPU (different streams or host threads) race and must be ordered by the
 * caller; the first call on each device resolves the array's device address
 * and should be made during set-up, before any stream capture begins.  Calls
 * no cudaMalloc, cudaFree or any other allocating routine. */
#define PSIM_STATIC_MAX_PARTICLES 1048576
#define PSIM_STATIC_SCRATCH_BYTES 65792
psim_status psim_compute_locations_static(const psim_params* params, int precision, int64_t n,
                                          void* pos_dev, const void* vel_dev,
                                          void* force_old_dev, void* force_new_dev,
                                          void* displacement_dev, const uint8_t* active_dev,
                                          int32_t* rebuild_flag_dev, void* cuda_stream);

/* --- measurement helpers ------------------------------------------------- */

/* Dependent-free FMA loop on every SM: the measured FMA-pipe peak of this GPU
 * in TFLOP/s for `precision` 4 or 8 (MEASURED_PEAKS.json holds no FP64/FP32
 * vector figure).  sm_mhz (may be NULL) receives the SM clock derived from
 * clock64() over the run. */
psim_status psim_measure_fma_peak(int precision, int device, double* tflops,
                                  double* sm_mhz);

/* Dense tcgen05.mma rate of this GPU in TFLOP/s: one issuing thread per SM on all
 * SMs, M128 N256 instructions back to back into two TMEM accumulators.  kind 0 =
 * kind::tf32 (K = 8 per instruction), kind 1 = kind::f16 (K = 16, what the
 * default all-pairs kernel issues).  The roofline denominator of the tensor-core
 * kernel (MEASURED_PEAKS.json has a cuBLAS bf16 figure only). */
psim_status psim_measure_tensor_peak(int kind, int device, double* tflops, double* sm_mhz);

const char* psim_status_string(int status);
/* Text of the last CUDA/NCCL error seen by this thread ("" if none). */
const char* psim_last_error(void);
const char* psim_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PSIM_H */
