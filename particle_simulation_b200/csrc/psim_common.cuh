// psim_common.cuh — shared types of the device code.
//
// Data layout in HBM (see DESIGN.md "Data layout"): every per-particle vector
// is a 4-wide record of the working real type R (16 B for float, 32 B for
// double) so that one particle is one aligned vector load:
//   posw[i] = (x, y, z, w)   w = pre-filter weight |x|^2 (1-k), or HUGE when the
//                            particle is inactive / padding (never in range)
//   vel[i], f_old[i], f_new[i] = (vx, vy, vz, 0) ...
//   disp[i] = (dx, dy, dz, |d|)   displacement since the last rebuild and its norm
// posw is double-buffered: a step reads x_n from one buffer and writes x_{n+1}
// to the other, so no block ever reads a half-updated position.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/psim.h"

namespace psim {

template <class R> struct Real4;
template <> struct alignas(16) Real4<float>  { float x, y, z, w; };
template <> struct alignas(32) Real4<double> { double x, y, z, w; };

template <class R> struct RealTraits;
template <> struct RealTraits<float> {
  static constexpr float unit_roundoff = 5.9604645e-8f;   // 2^-24
  static constexpr float huge = 1.0e30f;
  static constexpr int bytes = 4;
};
template <> struct RealTraits<double> {
  static constexpr double unit_roundoff = 1.1102230246251565e-16;  // 2^-53
  static constexpr double huge = 1.0e300;
  static constexpr int bytes = 8;
};

// Correctly rounded, never-contracted arithmetic: the in-range pair force and
// the integrator use exactly one rounding per operation in the reference's
// expression order, so that a pair's contribution is bit-identical to the CPU
// oracle's and only the summation order differs.
__device__ __forceinline__ float  r_add(float a, float b)   { return __fadd_rn(a, b); }
__device__ __forceinline__ float  r_sub(float a, float b)   { return __fsub_rn(a, b); }
__device__ __forceinline__ float  r_mul(float a, float b)   { return __fmul_rn(a, b); }
__device__ __forceinline__ float  r_div(float a, float b)   { return __fdiv_rn(a, b); }
__device__ __forceinline__ float  r_sqrt(float a)           { return __fsqrt_rn(a); }
__device__ __forceinline__ float  r_ceil(float a)           { return ceilf(a); }
__device__ __forceinline__ double r_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double r_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double r_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double r_div(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double r_sqrt(double a)          { return __dsqrt_rn(a); }
__device__ __forceinline__ double r_ceil(double a)          { return ceil(a); }
__device__ __forceinline__ float  r_fma(float a, float b, float c)    { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double r_fma(double a, double b, double c) { return __fma_rn(a, b, c); }

// Constants of one simulation in working precision.  The first block restates
// set_simulation_parameter (module_stoermer_verlet_method.f08:68-81); the
// second block belongs to the all-pairs pre-filter (DESIGN.md "Pre-filter").
template <class R>
struct Consts {
  R delta_t, epsilon, r_min, r_cut, min_cell, dr_max_tolerable, F_rcut, tmf;
  R area[3], cell[3];
  int nc[3];
  R one_minus_k;   // weight scale: w = |x|^2 * (1 - k)
  R cut2_pad;      // r_cut^2 (1 + delta): upper bound of what can pass the exact test
  R huge;
  R boff;          // FP64 cluster kernel: offset B that keeps est + B non-negative (0 for FP32)
  R bslack;        // absolute slack covering the roundings B introduces
  int slices;      // NB_OF_THREADS: slices of the rebuild trigger
};

template <class R>
__device__ __forceinline__ R filter_weight(R x, R y, R z, const Consts<R>& c) {
  return (x * x + y * y + z * z) * c.one_minus_k;
}

// ceiling(location / length_of_cell) against 1..nc
// (module_particle_management.f08:440-441, 503).
template <class R>
__device__ __forceinline__ bool inside_grid(R x, R y, R z, const Consts<R>& c) {
  int cx = (int)r_ceil(r_div(x, c.cell[0]));
  int cy = (int)r_ceil(r_div(y, c.cell[1]));
  int cz = (int)r_ceil(r_div(z, c.cell[2]));
  return cx >= 1 && cx <= c.nc[0] && cy >= 1 && cy <= c.nc[1] && cz >= 1 && cz <= c.nc[2];
}

// The truncated, shifted Lennard-Jones pair term of
// module_stoermer_verlet_method.f08:117-128 for particle i against j:
// subtracts `force` from (fx, fy, fz) when r < r_cut.  Returns whether the pair
// was in range.
template <class R>
__device__ __forceinline__ bool pair_force(R xi, R yi, R zi, R xj, R yj, R zj,
                                           const Consts<R>& c, R& fx, R& fy, R& fz) {
  const R dx = r_sub(xi, xj), dy = r_sub(yi, yj), dz = r_sub(zi, zj);       // :117
  const R r = r_sqrt(r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz)));   // :119
  if (!(r < c.r_cut)) return false;                                         // :121
  const R q = r_div(c.r_min, r);
  const R q2 = r_mul(q, q);
  const R s = r_mul(r_mul(q2, q2), q2);                                     // :123
  const R F = r_mul(r_div(s, r_mul(r, r)), r_sub((R)1, s));                 // :124
  const R g = r_sub(F, c.F_rcut);
  const R k12 = (R)12;
  fx = r_sub(fx, r_mul(r_mul(r_mul(dx, k12), c.epsilon), g));               // :126, :128
  fy = r_sub(fy, r_mul(r_mul(r_mul(dy, k12), c.epsilon), g));
  fz = r_sub(fz, r_mul(r_mul(r_mul(dz, k12), c.epsilon), g));
  return true;
}

// pair_force without branches, for exact paths that evaluate queued candidates (tensor-core kernel, cell kernel):
// candidate j (its record pj already loaded) of particle ig in the working precision, the reference's own arithmetic: (ax, ay, az) = 0 - the pair's
// force term (bit-identical to what pair_force subtracts) when the pair is in range, 0 otherwise.  Returns the
// verdict r < r_cut, the predicate of count_pairs_kernel (psim_allpairs.cuh): it feeds the audit counter.
template <class R, class CT>
__device__ __forceinline__ bool pair_term(const CT& c, int64_t ig, const Real4<R>& pi, int64_t jg, const Real4<R>& pj,
                                          R& ax, R& ay, R& az) {
  const R dx = r_sub(pi.x, pj.x), dy = r_sub(pi.y, pj.y), dz = r_sub(pi.z, pj.z);       // stoermer:117
  const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
  const R r = r_sqrt(r2);                                                               // :119
  const bool in = (jg != ig) & !(pj.w >= c.huge) & (r2 < c.cut2_pad) & (r < c.r_cut);   // :121
  const R rs = in ? r : (R)1;                          // keeps the unused evaluations finite
  const R q = r_div(c.r_min, rs);
  const R q2 = r_mul(q, q);
  const R s = r_mul(r_mul(q2, q2), q2);                                                 // :123
  const R F = r_mul(r_div(s, r_mul(rs, rs)), r_sub((R)1, s));                           // :124
  const R g = r_sub(F, c.F_rcut);
  const R k12 = (R)12;
  ax = in ? r_sub((R)0, r_mul(r_mul(r_mul(dx, k12), c.epsilon), g)) : (R)0;             // :126, :128
  ay = in ? r_sub((R)0, r_mul(r_mul(r_mul(dy, k12), c.epsilon), g)) : (R)0;
  az = in ? r_sub((R)0, r_mul(r_mul(r_mul(dz, k12), c.epsilon), g)) : (R)0;
  return in;
}

// (first, second) = the two largest values seen (module_stoermer_verlet_method.f08:168-173)
template <class R>
__device__ __forceinline__ void top2_push(R& first, R& second, R v) {
  if (v > first) { second = first; first = v; }
  else if (v > second) { second = v; }
}

template <class R>
__device__ __forceinline__ void top2_warp_reduce(R& first, R& second) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    R of = __shfl_xor_sync(0xffffffffu, first, off);
    R os = __shfl_xor_sync(0xffffffffu, second, off);
    // merge two (first, second) pairs
    R nf = first > of ? first : of;
    R lo = first > of ? of : first;
    R hs = second > os ? second : os;
    R ns = lo > hs ? lo : hs;
    first = nf;
    second = ns;
  }
}

}  // namespace psim
