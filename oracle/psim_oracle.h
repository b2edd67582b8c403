/*
 * psim_oracle.h — C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.  The
 * product (libpsim_b200.so) never links or calls it.
 *
 * PARITY UNPINNED: the reference ships no tests, golden outputs or fixtures
 *  and is Fortran 2008, for which no compiler exists on the build
 * or GPU machines, so this restatement could not be checked
 * against the reference's own executable.  It is pinned instead by closed-form
 * pair-force known answers, an independent brute-force all-pairs evaluation in
 * double, conservation laws, and byte-level format fixtures (tests/).
 *
 * The oracle restates the reference's OWN algorithm — linked cells, per-particle
 * half neighbour lists, Newton's third law, displacement-triggered rebuilds — in
 * the reference's single-thread summation order.  Citations: `stoermer` =
 * fortran_src/module_stoermer_verlet_method.f08, `pm` =
 * fortran_src/module_particle_management.f08, `program` =
 * fortran_src/program_particle_simulation_openmp.f08.
 */
#ifndef PSIM_ORACLE_H
#define PSIM_ORACLE_H

#include <stdint.h>
#include "../include/psim.h" /* psim_params only */

#ifdef __cplusplus
extern "C" {
#endif

typedef struct psim_oracle psim_oracle;

/* precision 4 or 8.  threads: OpenMP threads for the force/drift/kick loops
 * (1 = the deterministic oracle; >1 = timed baseline, f_j updated atomically to
 * repair the reference's race).  The rebuild trigger is evaluated
 * per params->nb_of_threads slices regardless of `threads`. */
psim_oracle* psim_oracle_create(const psim_params* params, int precision, int64_t n,
                                const int32_t* id, const void* pos, const void* vel,
                                int threads);
void psim_oracle_destroy(psim_oracle* o);

/* nsteps trips of the reference time loop (program:80-117). */
void psim_oracle_advance(psim_oracle* o, int64_t nsteps);

void psim_oracle_get_frame(psim_oracle* o, int32_t* id, void* pos, void* vel, uint8_t* active);
void psim_oracle_get_aux(psim_oracle* o, void* displacement, void* force_old, void* force_new);

int64_t psim_oracle_rebuilds(psim_oracle* o);
int64_t psim_oracle_steps(psim_oracle* o);
int64_t psim_oracle_n_active(psim_oracle* o);
/* number of half-list entries currently held (pairs visited per force pass) */
int64_t psim_oracle_list_entries(psim_oracle* o);
/* seconds spent inside the time loop of psim_oracle_advance since create */
double psim_oracle_loop_seconds(psim_oracle* o);

/* Derived constants in working precision, widened to double:
 * out[0..6] = r_min, r_cut, min_length_of_cell, dr_max_tolerable, F_rcut,
 * time_mass_factor, delta_t  (stoermer:68-81). */
void psim_oracle_constants(const psim_params* params, int precision, double out[7]);

/* Signed magnitude g(r) of the force on i along r_ij, so that f_i = r_ij * g / ... :
 * returns the scalar multiplying r_ij in f_i (stoermer:121-128) for a pair at
 * distance r along +x, evaluated in `precision`. */
double psim_oracle_pair_force_x(const psim_params* params, int precision, double r);

/* Independent ground truth: brute-force all-pairs force in double with the
 * same predicate (r < r_cut on active pairs), compensated summation.
 * Positions are given in `precision`; force_out is double[3n].  Returns the
 * ordered in-range pair count. */
int64_t psim_oracle_bruteforce(const psim_params* params, int precision, int64_t n,
                               const void* pos, const uint8_t* active, double* force_out);

/* Energy/momentum of the active particles in double (potential as in SURVEY.md 8(c)). */
void psim_oracle_energy(psim_oracle* o, double* kinetic, double* potential, double momentum[3]);

#ifdef __cplusplus
}
#endif
#endif
