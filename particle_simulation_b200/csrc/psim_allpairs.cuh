// psim_allpairs.cuh — the tiled all-pairs force pass with the Stoermer-Verlet
// update fused into its tail.
//
// Replaces compute_forces / compute_force_between_two_particles
// (module_stoermer_verlet_method.f08:87-132) and, when MODE > 0, also
// compute_velocities (:191-206) and the next step's compute_locations
// (:136-187).  The reference walks half neighbour lists and updates both
// partners; here every block owns a tile of i-particles, scans ALL j, and only
// ever writes its own i's, so there is no scatter and no atomics, and the sum
// over j runs in ascending j order (run-to-run deterministic).
//
// Hot loop (per ordered pair): a 3-FMA norm-expansion pre-filter
//     est = w_j + x_i*(-2 x_j) + y_i*(-2 y_j) + z_i*(-2 z_j)   ~  r^2 - |x_i|^2
// compared against  cut_i = r_cut^2(1+delta) - |x_i|^2 (1-k).  The weight
// w_j = |x_j|^2 (1-k) carries a relative margin k that covers every rounding
// error of the expansion, so the filter never rejects a pair the exact test
// would accept (DESIGN.md "Pre-filter").  Pairs that pass are re-tested and
// evaluated in the reference's own arithmetic (difference form, sqrt, two
// divisions) by pair_force().
#pragma once

#include "psim_common.cuh"

namespace psim {

// Sharded runs, fused step: where the drift of an owned particle is stored besides this GPU's own buffer — the
// same record of every other rank's copy of the buffer (peer memory over NVLink, plain stores).  The all-gather of
// the position shards thereby happens inside the force kernel, block by block, and overlaps its arithmetic; what is
// left of the collective is a flag per rank (exchange_kernel, psim_integrate.cuh).  world == 0: off.
constexpr int MAX_PEERS = 16;
template <class R>
struct PeerPush {
  int world, rank;
  Real4<R>* pos_nxt[MAX_PEERS];   // [r]: rank r's copy of the buffer this step writes (entry `rank` is not used)
};

template <class R>
struct StepArgs {
  const Real4<R>* pos_cur;   // x_n, all particles           [n_pad]
  Real4<R>* pos_nxt;         // x_{n+1}, written for owned i [n_pad]
  Real4<R>* vel;
  Real4<R>* f_old;
  Real4<R>* f_new;
  Real4<R>* disp;
  R* top2_part;              // [2 * gridDim.x] per-block two largest |disp| (scratch)
  R* rank_pair;              // [2] this rank's two largest |disp|, written by the block that finishes last
  unsigned* ticket;          // block counter of that hand-over (0 between launches)
  int64_t n_pad;             // multiple of the j-tile
  int64_t i_lo, i_hi;        // owned slice of i (multi-GPU shard); i_hi <= n
  Consts<R> c;
  PeerPush<R> peers;         // MODE_FORCE_KICK_DRIFT only
};

// x_{n+1} of owned particle i: this GPU's buffer and every peer's.
template <class R>
__device__ __forceinline__ void store_next(const StepArgs<R>& a, int64_t i, const Real4<R>& x) {
  a.pos_nxt[i] = x;
  if (a.peers.world > 1) {
    for (int r = 0; r < a.peers.world; ++r)
      if (r != a.peers.rank) a.peers.pos_nxt[r][i] = x;
    // No fence here: the flag that tells the peers is raised by exchange_kernel, a later kernel of the same stream,
    // and a grid's stores (peer memory included) are complete when the grid is — the order every stream-ordered
    // peer-to-peer signal relies on.  A system-scope fence per thread cost 50 us per step on 2 GPUs.
  }
}

enum { MODE_FORCE = 0, MODE_FORCE_KICK = 1, MODE_FORCE_KICK_DRIFT = 2 };

// Kick of step n then drift of step n+1 for one particle, in the reference's
// rounding order.  Returns |displacement|.
template <class R>
__device__ __forceinline__ R drift_one(const Consts<R>& c, Real4<R> x, const Real4<R>& v,
                                       R fx, R fy, R fz, Real4<R>& d, Real4<R>& x_out) {
  const R ddx = r_mul(c.delta_t, r_add(v.x, r_mul(c.tmf, fx)));   // stoermer:160
  const R ddy = r_mul(c.delta_t, r_add(v.y, r_mul(c.tmf, fy)));
  const R ddz = r_mul(c.delta_t, r_add(v.z, r_mul(c.tmf, fz)));
  x.x = r_add(x.x, ddx); x.y = r_add(x.y, ddy); x.z = r_add(x.z, ddz);   // :161
  d.x = r_add(d.x, ddx); d.y = r_add(d.y, ddy); d.z = r_add(d.z, ddz);   // :162
  const R dr = r_sqrt(r_add(r_add(r_mul(d.x, d.x), r_mul(d.y, d.y)), r_mul(d.z, d.z)));   // :165
  d.w = dr;
  x.w = filter_weight(x.x, x.y, x.z, c);
  x_out = x;
  return dr;
}

// Block-level two largest values -> part[blockIdx.x]; the block that finishes
// last merges every block's pair into rank_pair (threadfence + ticket), so the
// rebuild trigger needs no reduction kernel of its own.
template <class R, int T>
__device__ __forceinline__ void block_top2_store(R first, R second, R* part, unsigned* ticket, R* rank_pair) {
  __shared__ R s_first[T / 32], s_second[T / 32];
  __shared__ bool s_last;
  top2_warp_reduce(first, second);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_first[warp] = first; s_second[warp] = second; }
  __syncthreads();
  if (warp == 0) {
    first = lane < T / 32 ? s_first[lane] : (R)0;
    second = lane < T / 32 ? s_second[lane] : (R)0;
    top2_warp_reduce(first, second);
    if (lane == 0) {
      part[2 * blockIdx.x] = first; part[2 * blockIdx.x + 1] = second;
      __threadfence();
      s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  first = second = (R)0;
  const volatile R* vp = part;
  for (unsigned b = threadIdx.x; b < gridDim.x; b += T) {
    top2_push(first, second, vp[2 * b]);
    top2_push(first, second, vp[2 * b + 1]);
  }
  top2_warp_reduce(first, second);
  __syncthreads();
  if (lane == 0) { s_first[warp] = first; s_second[warp] = second; }
  __syncthreads();
  if (warp == 0) {
    first = lane < T / 32 ? s_first[lane] : (R)0;
    second = lane < T / 32 ? s_second[lane] : (R)0;
    top2_warp_reduce(first, second);
    if (lane == 0) { rank_pair[0] = first; rank_pair[1] = second; *ticket = 0u; }
  }
}

// What follows the force sum of one owned particle, shared by every force
// kernel: MODE 0 stores the force; MODE 1 also kicks (stoermer:200-201); MODE 2
// also performs the NEXT step's drift (stoermer:155-165) into the other
// position buffer and feeds |displacement| to the block's top-2.
template <class R, int MODE>
__device__ __forceinline__ void step_epilogue(const StepArgs<R>& a, int64_t i, R fx, R fy, R fz,
                                              R& first, R& second) {
  const Consts<R>& c = a.c;
  const Real4<R> p = a.pos_cur[i];
  if (p.w >= c.huge) {                       // inactive: carried along untouched
    if (MODE == MODE_FORCE_KICK_DRIFT) store_next(a, i, p);
    return;
  }
  Real4<R> fn; fn.x = fx; fn.y = fy; fn.z = fz; fn.w = (R)0;
  if (MODE == MODE_FORCE) { a.f_new[i] = fn; return; }
  Real4<R> v = a.vel[i];
  const Real4<R> fo = a.f_old[i];
  v.x = r_add(v.x, r_mul(c.tmf, r_add(fn.x, fo.x)));
  v.y = r_add(v.y, r_mul(c.tmf, r_add(fn.y, fo.y)));
  v.z = r_add(v.z, r_mul(c.tmf, r_add(fn.z, fo.z)));
  a.vel[i] = v;
  if (MODE == MODE_FORCE_KICK) { a.f_new[i] = fn; return; }
  a.f_old[i] = fn;                           // force_old <- force_new
  Real4<R> d = a.disp[i], xn;
  const R dr = drift_one(c, p, v, fn.x, fn.y, fn.z, d, xn);
  a.disp[i] = d;
  store_next(a, i, xn);
  top2_push(first, second, dr);
}

// T threads, RI i-particles per thread (register tile), j-tiles of T records
// double-buffered in shared memory, U j's per filter group.
template <class R, int T, int RI, int MODE, bool EXACT_DIFF>
__global__ void __launch_bounds__(T)
allpairs_step_kernel(const StepArgs<R> a) {
  constexpr int U = 4;
  __shared__ Real4<R> tile[2][T];

  const Consts<R>& c = a.c;
  const int tid = threadIdx.x;
  const int64_t i_base = a.i_lo + (int64_t)blockIdx.x * (T * RI);

  // --- i side: positions and accumulators live in registers ---
  R xi[RI], yi[RI], zi[RI], cut[RI], fx[RI], fy[RI], fz[RI];
  int64_t ig[RI];
#pragma unroll
  for (int r = 0; r < RI; ++r) {
    ig[r] = i_base + r * T + tid;
    fx[r] = fy[r] = fz[r] = (R)0;
    if (ig[r] < a.i_hi) {
      const Real4<R> p = a.pos_cur[ig[r]];
      xi[r] = p.x; yi[r] = p.y; zi[r] = p.z;
      // inactive i: nothing can pass the filter
      cut[r] = (p.w >= c.huge) ? -c.huge : (EXACT_DIFF ? c.cut2_pad : c.cut2_pad - p.w);
    } else {
      xi[r] = yi[r] = zi[r] = (R)0;
      cut[r] = -c.huge;
      ig[r] = -1;
    }
  }

  // --- j side: stage (-2x, -2y, -2z, w) tiles; prefetch the next one in registers ---
  const int64_t ntiles = a.n_pad / T;
  Real4<R> pre = a.pos_cur[tid];
  pre.x *= (R)-2; pre.y *= (R)-2; pre.z *= (R)-2;
  tile[0][tid] = pre;
  __syncthreads();
  int buf = 0;

  for (int64_t jt = 0; jt < ntiles; ++jt) {
    if (jt + 1 < ntiles) pre = a.pos_cur[(jt + 1) * T + tid];
    const Real4<R>* __restrict__ tj = tile[buf];
    const int64_t j_base = jt * T;

#pragma unroll 2
    for (int jj = 0; jj < T; jj += U) {
      bool any = false;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const Real4<R> q = tj[jj + u];
#pragma unroll
        for (int r = 0; r < RI; ++r) {
          R est;
          if (EXACT_DIFF) {
            // difference form: 3 sub, 1 mul, 2 fma (the literal reading of stoermer:117-121)
            const R dx = xi[r] + (R)0.5 * q.x, dy = yi[r] + (R)0.5 * q.y, dz = zi[r] + (R)0.5 * q.z;
            est = dx * dx + dy * dy + dz * dz + (q.w >= c.huge ? c.huge : (R)0);
          } else {
            est = r_fma(xi[r], q.x, r_fma(yi[r], q.y, r_fma(zi[r], q.z, q.w)));
          }
          any |= est < cut[r];
        }
      }
      if (any) {
        // rare: some pair of this group may be in range; redo the group exactly
#pragma unroll 1
        for (int u = 0; u < U; ++u) {
          const Real4<R> q = tj[jj + u];
          if (q.w >= c.huge) continue;
          const R xj = (R)-0.5 * q.x, yj = (R)-0.5 * q.y, zj = (R)-0.5 * q.z;
          const int64_t jg = j_base + jj + u;
#pragma unroll
          for (int r = 0; r < RI; ++r) {
            if (jg == ig[r] || cut[r] <= -c.huge) continue;
            const R dx = r_sub(xi[r], xj), dy = r_sub(yi[r], yj), dz = r_sub(zi[r], zj);
            const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
            if (r2 < c.cut2_pad) pair_force(xi[r], yi[r], zi[r], xj, yj, zj, c, fx[r], fy[r], fz[r]);
          }
        }
      }
    }

    if (jt + 1 < ntiles) {
      pre.x *= (R)-2; pre.y *= (R)-2; pre.z *= (R)-2;
      tile[buf ^ 1][tid] = pre;
    }
    __syncthreads();
    buf ^= 1;
  }

  // --- epilogue: store the force; kick; drift into the other position buffer ---
  R first = (R)0, second = (R)0;
#pragma unroll
  for (int r = 0; r < RI; ++r)
    if (ig[r] >= 0) step_epilogue<R, MODE>(a, ig[r], fx[r], fy[r], fz[r], first, second);
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, T>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

// Ordered in-range pair count with the exact predicate (P_in).
template <class R, int T>
__global__ void __launch_bounds__(T)
count_pairs_kernel(const Real4<R>* __restrict__ pos, int64_t n_pad, int64_t i_lo, int64_t i_hi,
                   Consts<R> c, unsigned long long* out) {
  __shared__ Real4<R> tile[T];
  const int64_t i = i_lo + (int64_t)blockIdx.x * T + threadIdx.x;
  Real4<R> p; p.x = p.y = p.z = 0; p.w = c.huge;
  if (i < i_hi) p = pos[i];
  const bool act = !(p.w >= c.huge);
  unsigned long long cnt = 0;
  for (int64_t jt = 0; jt < n_pad; jt += T) {
    __syncthreads();
    tile[threadIdx.x] = pos[jt + threadIdx.x];
    __syncthreads();
    if (!act) continue;
    for (int jj = 0; jj < T; ++jj) {
      const Real4<R> q = tile[jj];
      if (q.w >= c.huge || jt + jj == i) continue;
      const R dx = r_sub(p.x, q.x), dy = r_sub(p.y, q.y), dz = r_sub(p.z, q.z);
      const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
      if (r2 < c.cut2_pad && r_sqrt(r2) < c.r_cut) ++cnt;
    }
  }
  for (int off = 16; off > 0; off >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(out, cnt);
}

// Kinetic / potential energy and momentum in double (diagnostic; O(N^2)).
// out[0]=K, out[1]=U (each unordered pair once), out[2..4]=P.
template <class R, int T>
__global__ void __launch_bounds__(T)
energy_kernel(const Real4<R>* __restrict__ pos, const Real4<R>* __restrict__ vel, int64_t n_pad,
              int64_t i_lo, int64_t i_hi, Consts<R> c, double mass, double* out) {
  __shared__ Real4<R> tile[T];
  const int64_t i = i_lo + (int64_t)blockIdx.x * T + threadIdx.x;
  Real4<R> p; p.x = p.y = p.z = 0; p.w = c.huge;
  if (i < i_hi) p = pos[i];
  const bool act = !(p.w >= c.huge);
  const double eps = (double)c.epsilon, rmin = (double)c.r_min, rcut = (double)c.r_cut, Frc = (double)c.F_rcut;
  const double sc = pow(rmin / rcut, 6.0);
  const double Cc = -(eps * (sc * sc - 2.0 * sc) - 6.0 * eps * Frc * rcut * rcut);
  double U = 0;
  for (int64_t jt = 0; jt < n_pad; jt += T) {
    __syncthreads();
    tile[threadIdx.x] = pos[jt + threadIdx.x];
    __syncthreads();
    if (!act) continue;
    for (int jj = 0; jj < T; ++jj) {
      const Real4<R> q = tile[jj];
      if (q.w >= c.huge || jt + jj <= i) continue;   // each unordered pair once
      const double dx = (double)p.x - (double)q.x, dy = (double)p.y - (double)q.y, dz = (double)p.z - (double)q.z;
      const double r2 = dx * dx + dy * dy + dz * dz;
      if (r2 < rcut * rcut) {
        const double t = rmin * rmin / r2, s = t * t * t;
        U += eps * (s * s - 2.0 * s) - 6.0 * eps * Frc * r2 + Cc;
      }
    }
  }
  double K = 0, px = 0, py = 0, pz = 0;
  if (act) {
    const Real4<R> v = vel[i];
    K = 0.5 * mass * ((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
    px = mass * v.x; py = mass * v.y; pz = mass * v.z;
  }
  double vals[5] = {K, U, px, py, pz};
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    double v = vals[k];
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) atomicAdd(out + k, v);
  }
}

}  // namespace psim
