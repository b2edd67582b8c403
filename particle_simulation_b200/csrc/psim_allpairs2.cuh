// psim_allpairs2.cuh — the production all-pairs force pass.
//
// Same contract as allpairs_step_kernel (psim_allpairs.cuh): force on every
// owned i from ALL j with the reference's pair term
// (module_stoermer_verlet_method.f08:117-129), then kick / drift in the tail.
// What differs is how the work is laid on a B200.
//
// Measured fact that shapes it (profiles/, round 1): a warp-wide shared-memory
// load delivers 16 B x 32 lanes to the register file in 4 cycles even when all
// lanes read the same address, and all four schedulers of an SM share that
// path.  With two i-particles per lane the first-generation kernel was bound by
// exactly that path, not by the FMA pipes.  Hence:
//
// 1. RI = 8 i-particles per lane, so one j record that reaches the registers is
//    used for eight distance tests.
// 2. The distance FILTER runs in packed FP32 for both working precisions:
//    FFMA2 (fma.rn.f32x2, new on sm_100) takes two j per instruction from an
//    SoA tile, FMNMX3 keeps a running minimum per i, and one compare per i per
//    eight j decides whether the group needs the exact path.  The filter is
//    conservative by construction (margins below), and every pair it lets
//    through is re-tested and evaluated in the WORKING precision with the
//    reference's own operations, so the forces are the same as a pure
//    FP64 / FP32 evaluation; the FP64 pipe is left to the in-range pairs only.
// 3. Warps are autonomous: the 4, 8 or 16 warps of a block own the same 32 RI
//    i-particles but different runs of j tiles, each staging its own 32-record
//    tile (one record per lane) behind __syncwarp — no block barrier in the hot
//    loop.  Tiles are dealt round-robin, so the exact-path work (in-range pairs
//    sit in a few tiles near the i-tile) is shared evenly.  Thread-block clusters split j further; partial sums meet in
//    (distributed) shared memory and are added in a fixed order, so the result
//    is run-to-run deterministic.  The split is chosen so that every scheduler
//    of every SM holds about seven warps.
//
// Filter margins.  With float-rounded coordinates xf the tile holds
// (-2 xf_j, w_j) and a lane holds (xf_i, cut_i):
//     est  = w_j + xf_i . (-2 xf_j)          ~ |xf_i - xf_j|^2 - |xf_i|^2
//     w_j  = |xf_j|^2 (1 - k) - s_j ,  cut_i = r_cut^2 (1 + d) - (|xf_i|^2 (1 - k) - s_i)
// k = 32 u32 dominates the rounding of the three FMAs (<= 8 u32 (|xf_i|^2 + |xf_j|^2));
// s = 16 u32 r_cut (|xf| + 1) dominates what rounding the coordinates to float
// can move a distance that is near r_cut (<= 2 r_cut sqrt(3) u32 (|x_i| + |x_j|)).
// A pair with true r < r_cut therefore always has est < cut_i.
#pragma once

#include <cooperative_groups.h>

#include "psim_allpairs.cuh"

namespace psim {

namespace cg = cooperative_groups;

struct FilterRec { float x, y, z, w; };

template <class R>
__device__ __forceinline__ FilterRec filter_record(const Real4<R>& p, const Consts<R>& c, float rcut) {
  constexpr float u32 = 5.9604645e-8f;
  FilterRec f;
  f.x = (float)p.x; f.y = (float)p.y; f.z = (float)p.z;
  const float n2 = f.x * f.x + f.y * f.y + f.z * f.z;
  f.w = n2 * (1.0f - 32.0f * u32) - 16.0f * u32 * rcut * (sqrtf(n2) + 1.0f);
  if (p.w >= c.huge) f.w = 1.0e30f;          // inactive or padding: never a candidate
  return f;
}

// The exact path of one flagged group of U j records for one lane: find which
// of the lane's RI i-particles fired (float state from shared memory, since
// registers cannot be indexed), then evaluate those in the working precision.
// Kept out of line so that its double-precision temporaries do not inflate
// the register allocation of the hot loop.
template <class R, int RI, int U>
__device__ __noinline__ void exact_group(const Real4<R>* __restrict__ pos, const Consts<R>& c, int64_t i_base,
                                         int lane, const float* six, const float* siy, const float* siz,
                                         const float* sic, const float* gx, const float* gy, const float* gz,
                                         const float* gw, int64_t jg0, R* acc_x, R* acc_y, R* acc_z) {
#pragma unroll 1
  for (int r = 0; r < RI; ++r) {
    const int li = r * 32 + lane;
    const float fxi = six[li], fyi = siy[li], fzi = siz[li], fci = sic[li];
    float m = 3.0e38f;
#pragma unroll
    for (int u = 0; u < U; ++u) m = fminf(m, fmaf(fxi, gx[u], fmaf(fyi, gy[u], fmaf(fzi, gz[u], gw[u]))));
    if (!(m < fci)) continue;
    const int64_t ig = i_base + li;
    const Real4<R> pi = pos[ig];
    R fx = (R)0, fy = (R)0, fz = (R)0;
#pragma unroll 1
    for (int u = 0; u < U; ++u) {
      const int64_t jg = jg0 + u;
      if (jg == ig || gw[u] >= 1.0e30f) continue;
      const Real4<R> pj = pos[jg];
      const R dx = r_sub(pi.x, pj.x), dy = r_sub(pi.y, pj.y), dz = r_sub(pi.z, pj.z);
      const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
      if (r2 < c.cut2_pad) pair_force(pi.x, pi.y, pi.z, pj.x, pj.y, pj.z, c, fx, fy, fz);
    }
    // pair_force subtracted from 0: add the group's sum to this warp's running sum
    acc_x[li] = r_add(acc_x[li], fx);
    acc_y[li] = r_add(acc_y[li], fy);
    acc_z[li] = r_add(acc_z[li], fz);
  }
}

// Dynamic shared memory of allpairs_warp_kernel<R, RI, W, MODE>.
template <class R, int RI, int W>
constexpr size_t allpairs_warp_smem() { return sizeof(R) * W * 3 * 32 * RI; }

template <class R, int RI, int W, int MODE>
__global__ void __launch_bounds__(32 * W, 16 / W)
allpairs_warp_kernel(const StepArgs<R> a, const int S) {
  // W warps per block = j-splits inside a block (4, 8 or 16; 16 warps per SM in every case)
  constexpr int T = 32 * W;
  constexpr int G = 32 * RI;                 // i-particles per block (and per cluster)
  constexpr int TJ = 32;                     // j records per warp tile: one per lane
  constexpr int U = 8;                       // j records per filter group
  __shared__ __align__(16) float tx[W][2][TJ], ty[W][2][TJ], tz[W][2][TJ], tw[W][2][TJ];
  __shared__ float six[G], siy[G], siz[G], sic[G];      // float i-state for the exact path's re-filter
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  R (*acc)[3][G] = reinterpret_cast<R (*)[3][G]>(dyn_smem);   // acc[W][3][G]: per-warp partial forces

  cg::cluster_group cluster = cg::this_cluster();
  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int crank = (int)(blockIdx.x % S);
  const int64_t i_base = a.i_lo + (int64_t)(blockIdx.x / S) * G;
  const float rcut = (float)c.r_cut;
  const float cut2f = (float)c.cut2_pad * (1.0f + 1.0e-6f);

  // --- i side: float copies for the filter, in registers and (for the exact path) in shared memory
  for (int li = tid; li < G; li += T) {
    const int64_t i = i_base + li;
    FilterRec f; f.x = f.y = f.z = 0.f; f.w = 1.0e30f;
    if (i < a.i_hi) f = filter_record(a.pos_cur[i], c, rcut);
    six[li] = f.x; siy[li] = f.y; siz[li] = f.z;
    sic[li] = f.w >= 1.0e30f ? -1.0e30f : cut2f - f.w;
#pragma unroll
    for (int w = 0; w < W; ++w) acc[w][0][li] = acc[w][1][li] = acc[w][2][li] = (R)0;
  }
  __syncthreads();
  float2 px[RI], py[RI], pz[RI];
  float cutf[RI];
#pragma unroll
  for (int r = 0; r < RI; ++r) {
    const int li = r * 32 + lane;
    px[r] = make_float2(six[li], six[li]);
    py[r] = make_float2(siy[li], siy[li]);
    pz[r] = make_float2(siz[li], siz[li]);
    cutf[r] = sic[li];
  }

  // --- j side: this warp's run of 32-record tiles ---
  // tiles are dealt round-robin to the S*W warps that share this i-tile, so that the
  // (spatially clustered) in-range pairs, i.e. the exact-path work, spread evenly
  // (32-bit tile arithmetic: n_pad < 2^31, and the hot loop has no registers to spare)
  const int total_tiles = (int)(a.n_pad / TJ);
  const int stride = S * W;
  const int tile0 = crank * W + warp;
  const int ntiles = tile0 < total_tiles ? (total_tiles - tile0 + stride - 1) / stride : 0;
  auto stage = [&](int buf, const Real4<R>& p) {
    const FilterRec f = filter_record(p, c, rcut);
    tx[warp][buf][lane] = -2.0f * f.x; ty[warp][buf][lane] = -2.0f * f.y; tz[warp][buf][lane] = -2.0f * f.z;
    tw[warp][buf][lane] = f.w;
  };
  Real4<R> pre; pre.x = pre.y = pre.z = (R)0; pre.w = c.huge;
  if (ntiles > 0) pre = a.pos_cur[tile0 * TJ + lane];
  stage(0, pre);
  __syncwarp();
  int buf = 0;

  for (int jt = 0; jt < ntiles; ++jt) {
    if (jt + 1 < ntiles) pre = a.pos_cur[(tile0 + (jt + 1) * stride) * TJ + lane];
    const int j_base = (tile0 + jt * stride) * TJ;

#pragma unroll
    for (int jj = 0; jj < TJ; jj += U) {
      const float4 X0 = *reinterpret_cast<const float4*>(&tx[warp][buf][jj]);
      const float4 X1 = *reinterpret_cast<const float4*>(&tx[warp][buf][jj + 4]);
      const float4 Y0 = *reinterpret_cast<const float4*>(&ty[warp][buf][jj]);
      const float4 Y1 = *reinterpret_cast<const float4*>(&ty[warp][buf][jj + 4]);
      const float4 Z0 = *reinterpret_cast<const float4*>(&tz[warp][buf][jj]);
      const float4 Z1 = *reinterpret_cast<const float4*>(&tz[warp][buf][jj + 4]);
      const float4 W0 = *reinterpret_cast<const float4*>(&tw[warp][buf][jj]);
      const float4 W1 = *reinterpret_cast<const float4*>(&tw[warp][buf][jj + 4]);
      bool any = false;
#pragma unroll
      for (int r = 0; r < RI; ++r) {
        const float2 e01 = __ffma2_rn(px[r], make_float2(X0.x, X0.y), __ffma2_rn(py[r], make_float2(Y0.x, Y0.y),
                           __ffma2_rn(pz[r], make_float2(Z0.x, Z0.y), make_float2(W0.x, W0.y))));
        const float2 e23 = __ffma2_rn(px[r], make_float2(X0.z, X0.w), __ffma2_rn(py[r], make_float2(Y0.z, Y0.w),
                           __ffma2_rn(pz[r], make_float2(Z0.z, Z0.w), make_float2(W0.z, W0.w))));
        const float2 e45 = __ffma2_rn(px[r], make_float2(X1.x, X1.y), __ffma2_rn(py[r], make_float2(Y1.x, Y1.y),
                           __ffma2_rn(pz[r], make_float2(Z1.x, Z1.y), make_float2(W1.x, W1.y))));
        const float2 e67 = __ffma2_rn(px[r], make_float2(X1.z, X1.w), __ffma2_rn(py[r], make_float2(Y1.z, Y1.w),
                           __ffma2_rn(pz[r], make_float2(Z1.z, Z1.w), make_float2(W1.z, W1.w))));
        float m = fminf(e01.x, fminf(e01.y, e23.x));                              // FMNMX3
        m = fminf(m, fminf(e23.y, e45.x));
        m = fminf(m, fminf(e45.y, e67.x));
        m = fminf(m, e67.y);
        any |= m < cutf[r];
      }
      if (any)
        exact_group<R, RI, U>(a.pos_cur, c, i_base, lane, six, siy, siz, sic, &tx[warp][buf][jj], &ty[warp][buf][jj],
                              &tz[warp][buf][jj], &tw[warp][buf][jj], j_base + jj, acc[warp][0], acc[warp][1],
                              acc[warp][2]);
    }

    if (jt + 1 < ntiles) stage(buf ^ 1, pre);
    __syncwarp();
    buf ^= 1;
  }

  // --- meet the other j-splits: warps of this block, then blocks of the cluster ---
  cluster.sync();
  R first = (R)0, second = (R)0;
  const int per = (G + S - 1) / S;           // every block finishes its share of the G particles
  const int li_end = min((crank + 1) * per, G);
  for (int li = crank * per + tid; li < li_end; li += T) {
    const int64_t i = i_base + li;
    R sx = (R)0, sy = (R)0, sz = (R)0;
    for (int rr = 0; rr < S; ++rr) {         // fixed order: deterministic
      const R* remote = cluster.map_shared_rank(&acc[0][0][0], rr);
#pragma unroll
      for (int w = 0; w < W; ++w) {
        sx = r_add(sx, remote[(w * 3 + 0) * G + li]);
        sy = r_add(sy, remote[(w * 3 + 1) * G + li]);
        sz = r_add(sz, remote[(w * 3 + 2) * G + li]);
      }
    }
    if (i < a.i_hi) step_epilogue<R, MODE>(a, i, sx, sy, sz, first, second);
  }
  cluster.sync();                            // nobody leaves while its shared memory is still being read
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, T>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

}  // namespace psim
