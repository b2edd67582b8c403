// psim_integrate.cuh — the O(N) kernels around the force pass: stand-alone drift
// and kick (used un-fused and at the ends of a fused run), the rebuild trigger,
// the rebuild's observable side effects, initial activation and frame packing.
#pragma once

#include "psim_allpairs.cuh"

namespace psim {

// Device-resident control block: lets the rebuild decision stay on the GPU
// (no host round trip inside the time loop).
struct Control {
  int flag[2];                  // rebuild flag of the current / next step (indexed by step parity)
  unsigned long long rebuilds;  // times the trigger fired
  unsigned long long n_active;  // scratch for counting
  unsigned ticket;              // last-block hand-over of the top-2 reduction
  unsigned comm_timeout;        // exchange_kernel gave up waiting for a peer (sharded runs with peer stores)
};

// ---- what is left of the per-step collective when the force kernel stores x_{n+1} into the peers' buffers --------
// (PeerPush, psim_allpairs.cuh): ONE small block launched after the force kernel on the same stream.
//
// publish: this rank's trigger pairs (two largest |displacement| per NB_OF_THREADS slice) go into every peer's
// table, then every peer's flag for this rank takes the step's epoch.  The force kernel's peer stores precede
// this kernel in stream order; the system-scope fence orders them (and the pairs) before the flags.
template <class R>
struct PeerMail {
  int world, rank, nvals;               // nvals = 2 * slices
  const R* mine;                        // this rank's pairs (local)
  R* top2[MAX_PEERS];                   // [r]: rank r's table [world][slices][2]
  unsigned long long* flags[MAX_PEERS]; // [r]: rank r's flag array [world]
};
template <class R>
__global__ void __launch_bounds__(64) exchange_kernel(const PeerMail<R> m, unsigned long long epoch, const unsigned long long* flags,
                                                       Control* ctrl, unsigned long long timeout_ns) {
  // ---- publish ----
  for (int r = 0; r < m.world; ++r) {
    if (r == m.rank) continue;
    for (int k = threadIdx.x; k < m.nvals; k += 64) m.top2[r][(size_t)m.nvals * m.rank + k] = m.mine[k];
  }
  __threadfence_system();
  __syncthreads();
  const int r = threadIdx.x;
  if (r >= m.world || r == m.rank) return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(m.flags[r] + m.rank), "l"(epoch) : "memory");
  // ---- await: peer r's flag in THIS rank's array has reached the epoch: its positions and pairs of the step are
  // in this GPU's memory.  Gives up after timeout_ns (a peer died): the host then reports PSIM_ERR_COMM instead of
  // hanging.  Nothing on this GPU waits for this kernel but the stream's own later work. ----
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + r) : "memory");
    if (v >= epoch) break;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) { ctrl->comm_timeout = 1u; break; }
    __nanosleep(200);
  }
}

// compute_locations (module_stoermer_verlet_method.f08:136-187) for the owned slice.
template <class R, int T>
__global__ void __launch_bounds__(T)
drift_kernel(const Real4<R>* __restrict__ pos_cur, Real4<R>* __restrict__ pos_nxt,
             const Real4<R>* __restrict__ vel, Real4<R>* __restrict__ f_old, Real4<R>* __restrict__ f_new,
             Real4<R>* __restrict__ disp, R* __restrict__ top2_part, unsigned* ticket, R* rank_pair,
             int64_t i_lo, int64_t i_hi, Consts<R> c) {
  const int64_t i = i_lo + (int64_t)blockIdx.x * T + threadIdx.x;
  R first = (R)0, second = (R)0;
  if (i < i_hi) {
    const Real4<R> p = pos_cur[i];
    if (p.w >= c.huge) {
      pos_nxt[i] = p;                       // :151 inactive particles are skipped
    } else {
      const Real4<R> fo = f_new[i];         // :155 force_old = force_new
      f_old[i] = fo;
      Real4<R> z; z.x = z.y = z.z = z.w = (R)0;
      f_new[i] = z;                         // :157
      Real4<R> d = disp[i], xn;
      const R dr = drift_one(c, p, vel[i], fo.x, fo.y, fo.z, d, xn);
      disp[i] = d;
      pos_nxt[i] = xn;
      first = dr;
    }
  }
  block_top2_store<R, T>(first, second, top2_part, ticket, rank_pair);
}

// compute_velocities (module_stoermer_verlet_method.f08:191-206).
template <class R, int T>
__global__ void __launch_bounds__(T)
kick_kernel(const Real4<R>* __restrict__ pos, Real4<R>* __restrict__ vel, const Real4<R>* __restrict__ f_old,
            const Real4<R>* __restrict__ f_new, int64_t i_lo, int64_t i_hi, Consts<R> c) {
  const int64_t i = i_lo + (int64_t)blockIdx.x * T + threadIdx.x;
  if (i >= i_hi) return;
  if (pos[i].w >= c.huge) return;
  Real4<R> v = vel[i];
  const Real4<R> fn = f_new[i], fo = f_old[i];
  v.x = r_add(v.x, r_mul(c.tmf, r_add(fn.x, fo.x)));
  v.y = r_add(v.y, r_mul(c.tmf, r_add(fn.y, fo.y)));
  v.z = r_add(v.z, r_mul(c.tmf, r_add(fn.z, fo.z)));
  vel[i] = v;
}

// NB_OF_THREADS > 1: the reference evaluates the trigger per thread slice
// (program_particle_simulation_openmp.f08:71,86,92; slices :305-323).  One
// block per slice scans |disp| (disp.w) of the part of its slice that this
// handle owns ([i_lo, i_hi): everything on a single GPU, the shard on a
// sharded handle) and writes the two largest values to out[2 * slice .. + 1]
// (zeros when the intersection is empty).  The housekeeping kernel merges the
// ranks' pairs slice by slice and ORs the verdicts.
template <class R>
__global__ void __launch_bounds__(256)
slice_top2_kernel(const Real4<R>* __restrict__ pos, const Real4<R>* __restrict__ disp, int64_t n,
                  int64_t i_lo, int64_t i_hi, R* __restrict__ out, Consts<R> c) {
  __shared__ R s_first[8], s_second[8];
  const int64_t per = n / c.slices;
  int64_t lo = per * blockIdx.x;
  int64_t hi = (blockIdx.x == (unsigned)c.slices - 1) ? n : per * (blockIdx.x + 1);
  if (lo < i_lo) lo = i_lo;
  if (hi > i_hi) hi = i_hi;
  R first = (R)0, second = (R)0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += 256)
    if (!(pos[i].w >= c.huge)) top2_push(first, second, disp[i].w);
  top2_warp_reduce(first, second);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_first[warp] = first; s_second[warp] = second; }
  __syncthreads();
  if (warp == 0) {
    first = lane < 8 ? s_first[lane] : (R)0;
    second = lane < 8 ? s_second[lane] : (R)0;
    top2_warp_reduce(first, second);
    if (lane == 0) { out[2 * blockIdx.x] = first; out[2 * blockIdx.x + 1] = second; }
  }
}

// initialze_cell_space + update_active_particles at start-up
// (module_particle_management.f08:82-92, 95, 534-552): a particle outside
// [0, area] or whose cell index is outside the grid starts inactive.
template <class R, int T>
__global__ void __launch_bounds__(T)
init_positions_kernel(Real4<R>* __restrict__ pos_a, Real4<R>* __restrict__ pos_b, const R* __restrict__ xyz,
                      const unsigned char* __restrict__ active_in, int64_t n, int64_t n_pad, Consts<R> c) {
  const int64_t i = (int64_t)blockIdx.x * T + threadIdx.x;
  if (i >= n_pad) return;
  Real4<R> p; p.x = p.y = p.z = (R)0; p.w = c.huge;
  if (i < n) {
    p.x = xyz[3 * i]; p.y = xyz[3 * i + 1]; p.z = xyz[3 * i + 2];
    bool ok = !(p.x > c.area[0] || p.y > c.area[1] || p.z > c.area[2] || p.x < (R)0 || p.y < (R)0 || p.z < (R)0);
    ok = ok && inside_grid(p.x, p.y, p.z, c);
    if (active_in) ok = active_in[i] != 0;     // operator form: the caller decides
    p.w = ok ? filter_weight(p.x, p.y, p.z, c) : c.huge;
  }
  pos_a[i] = p;
  pos_b[i] = p;
}

template <class R, int T>
__global__ void __launch_bounds__(T)
unpack3_kernel(Real4<R>* __restrict__ dst, const R* __restrict__ xyz, int64_t n, int64_t n_pad) {
  const int64_t i = (int64_t)blockIdx.x * T + threadIdx.x;
  if (i >= n_pad) return;
  Real4<R> v; v.x = v.y = v.z = v.w = (R)0;
  if (i < n && xyz) { v.x = xyz[3 * i]; v.y = xyz[3 * i + 1]; v.z = xyz[3 * i + 2]; }
  dst[i] = v;
}

// Pack a record array into xyz triples (frame output / diagnostics).
template <class R, int T>
__global__ void __launch_bounds__(T)
pack3_kernel(R* __restrict__ xyz, const Real4<R>* __restrict__ src, int64_t lo, int64_t hi) {
  const int64_t i = lo + (int64_t)blockIdx.x * T + threadIdx.x;
  if (i >= hi) return;
  const Real4<R> v = src[i];
  xyz[3 * i] = v.x; xyz[3 * i + 1] = v.y; xyz[3 * i + 2] = v.z;
}

template <class R, int T>
__global__ void __launch_bounds__(T)
pack_active_kernel(unsigned char* __restrict__ active, const Real4<R>* __restrict__ pos, int64_t n, Consts<R> c,
                   Control* ctrl) {
  const int64_t i = (int64_t)blockIdx.x * T + threadIdx.x;
  bool act = false;
  if (i < n) {
    act = !(pos[i].w >= c.huge);
    if (active) active[i] = act ? 1 : 0;
  }
  const unsigned m = __ballot_sync(0xffffffffu, act);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(&ctrl->n_active, (unsigned long long)__popc(m));
}

// Dependent-free FMA loop: ILP independent chains per thread.
template <class R, int ILP>
__global__ void __launch_bounds__(256)
fma_peak_kernel(R* out, int iters, R a, R b, long long* cycles) {
  R acc[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) acc[k] = (R)(threadIdx.x + k);
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) acc[k] = r_fma(acc[k], a, b);
  }
  const long long t1 = clock64();
  R s = (R)0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) s += acc[k];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

}  // namespace psim
