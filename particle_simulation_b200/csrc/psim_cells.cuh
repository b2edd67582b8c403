// psim_cells.cuh — cell-binned force path and the per-step
// housekeeping kernel.
//
// What the reference achieves with linked cells and per-particle neighbour
// lists (module_particle_management.f08:52-108, 128-137, 179-339), done the GPU
// way.  Whenever the reference would rebuild its lists (the displacement
// trigger of module_stoermer_verlet_method.f08:177-185 fired), all particles
// are re-binned on the reference's own grid (nint(area/cell) cells, index
// ceiling(x/cell), pm:62, 95): perm lists the particles cell by cell (x
// fastest), so the 27-cell neighbourhood of a particle is nine contiguous runs
// of slots.  Between rebuilds the binning is kept: a cell is at least
// r_cut + dr_max_tolerable wide (stoermer:75), the same skin argument that
// keeps the reference's lists valid.  Every step a cell-ordered copy of the
// current positions (posj) is gathered through perm.
//
// Ownership is by PARTICLE INDEX (the i-sharding of the all-pairs path), so a
// sharded run needs no communication beyond the per-step position all-gather:
// every rank bins all particles redundantly (same data, deterministic order,
// same result) and walks its own particles in cell order through own_slots.
//
// Everything is conditional ON THE DEVICE: the housekeeping kernel evaluates
// the trigger itself and only then runs its grid-synchronised re-binning
// phases, so the host never synchronises to learn whether a rebuild happened,
// and a step is two launches: force (+kick+drift) and housekeeping.
#pragma once

#include <cooperative_groups.h>

#include "psim_allpairs.cuh"
#include "psim_allpairs2.cuh"
#include "psim_integrate.cuh"

namespace psim {

template <class R>
__device__ __forceinline__ int clamp_cell(R x, R cell, int nc) {
  int c = (int)r_ceil(r_div(x, cell));
  return c < 1 ? 1 : (c > nc ? nc : c);
}

// The force pass over the binned state.
//
// A WARP owns 32 consecutive slots, i.e. the particles of one or two cells
// cmin..cmax (one particle per lane).  For each of the nine (dy, dz) offsets the
// union of its lanes' neighbour cells is ONE contiguous run of slots (cells
// cmin-1+off .. cmax+1+off in linear cell order; a run that wraps around a grid
// edge only adds far-away particles, which the filter rejects).  The warp walks
// that run in 32-record tiles, one record per lane, staged as float SoA behind
// __syncwarp — no block barrier, no per-lane loop bounds, broadcast reads.
//
// Phase 1 is the packed-FP32 conservative filter of psim_allpairs2.cuh (FFMA2:
// two j per instruction); a lane appends the slots that pass to its private
// queue in shared memory.  In-range pairs are ~6 % of the neighbourhood: far too
// dense for a divergent exact evaluation per hit, cheap enough to queue.
// Phase 2 walks the queue with the exact pair term in the working precision.
// Sums run in ascending slot order: deterministic.
//
// The particle state stays in particle (array) order.  posj is a cell-ordered
// COPY of all positions, perm maps slot -> particle, and the block walks
// own_slots, the cell-ordered list of the slots whose particle this rank owns
// (all of them on a single GPU).  The force sum works on slots, the kick/drift
// epilogue on the particle index.
template <class R, int MODE, int QCAP, int T>
__global__ void __launch_bounds__(T)
cell_force_kernel(const StepArgs<R> a, const Real4<R>* __restrict__ posj, const int32_t* __restrict__ cell_of,
                  const int32_t* __restrict__ cell_start, const int32_t* __restrict__ own_slots,
                  const int32_t* __restrict__ own_start, const int32_t* __restrict__ perm, int ncell,
                  unsigned long long* __restrict__ tested) {
  constexpr int W = T / 32, TJ = 32;
  __shared__ int32_t queue[QCAP * T];
  __shared__ __align__(16) float tx[W][2][TJ], ty[W][2][TJ], tz[W][2][TJ], tw[W][2][TJ];
  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t idx = (int64_t)blockIdx.x * T + tid;
  int64_t k = -1;              // slot in cell order (j-side numbering)
  int64_t ip = -1;             // particle index in the state arrays (epilogue numbering)
  if (idx < own_start[ncell]) { k = own_slots[idx]; ip = perm[k]; }
  const float rcut = (float)c.r_cut;
  const float cut2f = (float)c.cut2_pad * (1.0f + 1.0e-6f);

  R xi = 0, yi = 0, zi = 0, fx = 0, fy = 0, fz = 0;
  float fxi = 0.f, fyi = 0.f, fzi = 0.f, cutf = -1.0e30f;
  int cell = ncell;
  if (k >= 0) {
    cell = cell_of[k];
    const Real4<R> p = posj[k];
    if (cell < ncell && !(p.w >= c.huge)) {
      xi = p.x; yi = p.y; zi = p.z;
      const FilterRec f = filter_record(p, c, rcut);
      fxi = f.x; fyi = f.y; fzi = f.z; cutf = cut2f - f.w;
    } else {
      cell = ncell;
    }
  }
  // cells spanned by this warp's active slots
  int cmin = cell < ncell ? cell : INT_MAX, cmax = cell < ncell ? cell : -1;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    cmin = min(cmin, __shfl_xor_sync(0xffffffffu, cmin, off));
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, off));
  }

  int cnt = 0;
  unsigned ntest = 0;          // distance tests of this warp (active lanes x valid records), reported when asked for
  const unsigned nactive = __popc(__ballot_sync(0xffffffffu, cell < ncell));
  // Two queue entries per trip: independent chains (sqrt, two divisions each) that the
  // scheduler interleaves; the additions keep queue order, so the sum stays deterministic.
  auto flush = [&]() {
    if (cnt == 0) return;
    // software pipeline: the next two records are in flight while two pair terms are evaluated
    int ja = queue[tid], jb = cnt > 1 ? queue[T + tid] : ja;
    Real4<R> qa = posj[ja], qb = posj[jb];
    int e = 0;
    for (; e + 1 < cnt; e += 2) {
      const int j0 = ja, j1 = jb;
      const Real4<R> q0 = qa, q1 = qb;
      if (e + 2 < cnt) { ja = queue[(e + 2) * T + tid]; qa = posj[ja]; }
      if (e + 3 < cnt) { jb = queue[(e + 3) * T + tid]; qb = posj[jb]; }
      R ax0 = (R)0, ay0 = (R)0, az0 = (R)0, ax1 = (R)0, ay1 = (R)0, az1 = (R)0;
      if (j0 != (int)k) pair_force(xi, yi, zi, q0.x, q0.y, q0.z, c, ax0, ay0, az0);
      if (j1 != (int)k) pair_force(xi, yi, zi, q1.x, q1.y, q1.z, c, ax1, ay1, az1);
      // pair_force subtracted each force from 0; apply them in order
      fx = r_add(r_add(fx, ax0), ax1); fy = r_add(r_add(fy, ay0), ay1); fz = r_add(r_add(fz, az0), az1);
    }
    if (e < cnt && ja != (int)k) pair_force(xi, yi, zi, qa.x, qa.y, qa.z, c, fx, fy, fz);
    cnt = 0;
  };

  if (cmax >= 0) {                       // warp-uniform: the warp has at least one active particle
    const int nx = c.nc[0], ny = c.nc[1];
    auto fetch = [&](int j, int tb) {      // global load only: the conversion waits until the data is needed
      Real4<R> p; p.x = p.y = p.z = (R)0; p.w = c.huge;
      if (j < tb) p = posj[j];
      return p;
    };
    auto stage = [&](int buf, const Real4<R>& p) {
      const FilterRec f = filter_record(p, c, rcut);
      tx[warp][buf][lane] = -2.0f * f.x; ty[warp][buf][lane] = -2.0f * f.y; tz[warp][buf][lane] = -2.0f * f.z;
      tw[warp][buf][lane] = f.w;
    };
    int next_cell = 0;                   // runs are visited in ascending cell order: never scan a cell twice
    int qidx = tid;                      // queue slot of this lane's next entry: cnt * T + tid
    for (int dz = -1; dz <= 1; ++dz) {
      for (int dy = -1; dy <= 1; ++dy) {
        const int off = nx * (dy + ny * dz);
        const int blo = max(max(cmin - 1 + off, 0), next_cell), bhi = min(cmax + 1 + off, ncell - 1);
        if (bhi < blo) continue;         // warp-uniform
        next_cell = bhi + 1;
        const int ta = cell_start[blo], tb = cell_start[bhi + 1];
        if (tb <= ta) continue;
        int buf = 0;
        stage(0, fetch(ta + lane, tb));
        __syncwarp();
        for (int t0 = ta; t0 < tb; t0 += TJ) {
          // room for a whole tile in every lane's queue, decided for the warp as one
          if (__any_sync(0xffffffffu, cnt > QCAP - TJ)) { flush(); qidx = tid; }
          const bool more = t0 + TJ < tb;
          ntest += nactive * (unsigned)min(TJ, tb - t0);
          Real4<R> pre;
          if (more) pre = fetch(t0 + TJ + lane, tb);   // in flight during this tile's filter
          const float2 px = make_float2(fxi, fxi), py = make_float2(fyi, fyi), pz = make_float2(fzi, fzi);
          unsigned hits = 0u;              // bit jj: record t0 + jj passed the filter for this lane
#pragma unroll
          for (int jj = 0; jj < TJ; jj += 4) {
            const float4 X = *reinterpret_cast<const float4*>(&tx[warp][buf][jj]);
            const float4 Y = *reinterpret_cast<const float4*>(&ty[warp][buf][jj]);
            const float4 Z = *reinterpret_cast<const float4*>(&tz[warp][buf][jj]);
            const float4 Wt = *reinterpret_cast<const float4*>(&tw[warp][buf][jj]);
            const float2 e01 = __ffma2_rn(px, make_float2(X.x, X.y), __ffma2_rn(py, make_float2(Y.x, Y.y),
                               __ffma2_rn(pz, make_float2(Z.x, Z.y), make_float2(Wt.x, Wt.y))));
            const float2 e23 = __ffma2_rn(px, make_float2(X.z, X.w), __ffma2_rn(py, make_float2(Y.z, Y.w),
                               __ffma2_rn(pz, make_float2(Z.z, Z.w), make_float2(Wt.z, Wt.w))));
            if (e01.x < cutf) hits |= 1u << jj;
            if (e01.y < cutf) hits |= 2u << jj;
            if (e23.x < cutf) hits |= 4u << jj;
            if (e23.y < cutf) hits |= 8u << jj;
          }
          // expand the tile's hit mask into this lane's queue (a few entries per tile)
          while (hits) {
            const int b = __ffs(hits) - 1;
            hits &= hits - 1;
            queue[qidx] = t0 + b;
            qidx += T;
          }
          cnt = (qidx - tid) / T;
          if (more) stage(buf ^ 1, pre);
          __syncwarp();
          buf ^= 1;
        }
      }
    }
    flush();
    if (tested && lane == 0) atomicAdd(tested, (unsigned long long)ntest);
  }

  R first = (R)0, second = (R)0;
  if (ip >= 0) step_epilogue<R, MODE>(a, ip, fx, fy, fz, first, second);
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, T>(first, second, a.top2_part, a.ticket, a.rank_pair);
}


// ---- housekeeping: trigger, rebuild side effects, re-binning, position copy --------

template <class R>
struct HouseArgs {
  Real4<R>* pos_new;      // the buffer the drift just filled (the current one from here on)
  Real4<R>* pos_other;    // the other buffer: next step's target
  Real4<R>* disp;
  const R* rank_pairs;    // [world][slices][2] every rank's two largest |disp| per trigger slice (null: no trigger this call)
  int world;
  Control* ctrl;
  int parity;
  int64_t n, i_lo, i_hi;
  Consts<R> c;
  // cell path (cells == 0: all-pairs, nothing to bin)
  int cells, force_rebin, ncell;
  int carry_all;          // no cell tables on this rank (it owns nothing): still carry every inactive record over
  int32_t *key, *perm, *cell_start, *own_start, *cell_of, *own_slots;
  unsigned *count, *own_count, *fill;
  Real4<R>* posj;
};

// exclusive scan of cnt[0..nb) into start[0..nb] by ONE block of 256 threads
__device__ __forceinline__ void block_scan_256(const unsigned* cnt, int32_t* start, unsigned* zero_me, int nb,
                                               unsigned* s_sum) {
  const int tid = threadIdx.x;
  const int per = (nb + 255) / 256;
  const int lo = tid * per, hi = min(nb, lo + per);
  unsigned sum = 0;
  for (int b = lo; b < hi; ++b) sum += cnt[b];
  s_sum[tid] = sum;
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {
    const unsigned v = tid >= off ? s_sum[tid - off] : 0u;
    __syncthreads();
    s_sum[tid] += v;
    __syncthreads();
  }
  unsigned run = s_sum[tid] - sum;
  for (int b = lo; b < hi; ++b) {
    start[b] = (int32_t)run;
    run += cnt[b];
    if (zero_me) zero_me[b] = 0u;
  }
  if (tid == 255) start[nb] = (int32_t)s_sum[255];
  __syncthreads();
}

// One cooperative launch per step (256 threads per block, grid = what is co-resident).
//  1. trigger: dr_max_first + dr_max_second > dr_max_tolerable (stoermer:168-185) per NB_OF_THREADS slice
//     (program:71,86,92) over the merged per-rank pairs; every block computes it for itself.
//  2. if it fired: the observable side effects of update_neighbor_list_and_cell_space
//     (pm:112-124): displacement reset (pm:158), deactivation of particles whose cell index left
//     the grid (pm:440-441, 503-504, 534-552);
//  3. cell path, if it fired (or on demand): histogram -> scan -> scatter -> per-cell order,
//     separated by grid-wide barriers;
//  4. cell path, always: posj = positions in cell order; owned inactive records are carried
//     into the other position buffer (they are not in own_slots, so no force block does it).
template <class R>
__global__ void __launch_bounds__(256)
housekeeping_kernel(const HouseArgs<R> h) {
  namespace cgn = cooperative_groups;
  cgn::grid_group grid = cgn::this_grid();
  __shared__ unsigned s_sum[256];
  const Consts<R>& c = h.c;
  const int64_t gtid = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const int64_t gsz = (int64_t)gridDim.x * 256;

  // The reference's trigger, slice by slice (program:71,86,92): the two largest |disp| of a slice are the
  // two largest among the ranks' pairs for that slice; any slice over the tolerance fires the rebuild.
  int mine = 0;
  if (h.rank_pairs) {
    for (int k = threadIdx.x; k < c.slices; k += 256) {
      R first = (R)0, second = (R)0;
      for (int r = 0; r < h.world; ++r) {
        const R* pr = h.rank_pairs + 2 * ((size_t)r * c.slices + k);
        top2_push(first, second, pr[0]);
        top2_push(first, second, pr[1]);
      }
      if (r_add(first, second) > c.dr_max_tolerable) mine = 1;
    }
  }
  const int flag = __syncthreads_or(mine);
  if (gtid == 0) {
    h.ctrl->flag[h.parity] = flag;
    if (flag) h.ctrl->rebuilds += 1;
  }
  if (gtid == 1 || (gtid == 0 && gsz == 1)) h.ctrl->flag[h.parity ^ 1] = 0;
  const bool rebin = h.cells && (flag || h.force_rebin);

  if (flag || rebin) {
    for (int64_t i = gtid; i < h.n; i += gsz) {
      Real4<R> p = h.pos_new[i];
      bool active = !(p.w >= c.huge);
      if (flag && active) {
        if (inside_grid(p.x, p.y, p.z, c)) {
          if (i >= h.i_lo && i < h.i_hi) { Real4<R> z; z.x = z.y = z.z = z.w = (R)0; h.disp[i] = z; }
        } else {
          p.w = c.huge;
          h.pos_new[i] = p;
          active = false;
        }
      }
      if (rebin) {
        int cell = h.ncell;
        if (active) {
          const int cx = clamp_cell(p.x, c.cell[0], c.nc[0]);
          const int cy = clamp_cell(p.y, c.cell[1], c.nc[1]);
          const int cz = clamp_cell(p.z, c.cell[2], c.nc[2]);
          cell = (cx - 1) + c.nc[0] * ((cy - 1) + c.nc[1] * (cz - 1));
        }
        h.key[i] = cell;
        atomicAdd(&h.count[cell], 1u);
        if (i >= h.i_lo && i < h.i_hi && cell < h.ncell) atomicAdd(&h.own_count[cell], 1u);
      }
    }
  }

  if (rebin) {
    grid.sync();
    if (blockIdx.x == 0) {
      block_scan_256(h.count, h.cell_start, h.fill, h.ncell + 1, s_sum);
      block_scan_256(h.own_count, h.own_start, nullptr, h.ncell + 1, s_sum);
    }
    grid.sync();
    for (int64_t i = gtid; i < h.n; i += gsz) {
      const int cell = h.key[i];
      h.perm[h.cell_start[cell] + atomicAdd(&h.fill[cell], 1u)] = (int32_t)i;
    }
    grid.sync();
    // per cell: ascending particle order (deterministic), cell of every slot, owned slots in cell order
    for (int64_t cell = gtid; cell <= h.ncell; cell += gsz) {
      h.count[cell] = 0u;
      h.own_count[cell] = 0u;
      const int lo = h.cell_start[cell], hi = h.cell_start[cell + 1];
      if (cell == h.ncell) {                       // inactive tail: never read through the cell tables
        for (int a = lo; a < hi; ++a) h.cell_of[a] = h.ncell;
        continue;
      }
      for (int a = lo + 1; a < hi; ++a) {
        const int32_t v = h.perm[a];
        int b = a - 1;
        while (b >= lo && h.perm[b] > v) { h.perm[b + 1] = h.perm[b]; --b; }
        h.perm[b + 1] = v;
      }
      int m = h.own_start[cell];
      for (int a = lo; a < hi; ++a) {
        h.cell_of[a] = (int32_t)cell;
        const int32_t i = h.perm[a];
        if (i >= h.i_lo && i < h.i_hi) h.own_slots[m++] = a;
      }
    }
    grid.sync();
  }

  if (h.cells) {
    // (every inactive record, not only the owned ones: when positions travel by peer stores no all-gather
    // spreads a rank's copies, and every rank knows all of them anyway)
    for (int64_t k = gtid; k < h.n; k += gsz) {
      const int32_t i = h.perm[k];
      const Real4<R> p = h.pos_new[i];
      h.posj[k] = p;
      if (p.w >= c.huge) h.pos_other[i] = p;
    }
  } else if (h.carry_all) {
    for (int64_t i = gtid; i < h.n; i += gsz) {
      const Real4<R> p = h.pos_new[i];
      if (p.w >= c.huge) h.pos_other[i] = p;
    }
  }
}

template <class R>
struct CellPath {
  static constexpr int TF = 64;          // threads of the force kernel (two autonomous warps)
  static constexpr int QCAP = 96;        // queue entries per lane
  int64_t launches = 0;
  int64_t n = 0, n_pad = 0;
  int ncell = 0;
  bool binned = false;                   // perm / posj are valid for the current positions
  int32_t *key = nullptr, *perm = nullptr, *cell_start = nullptr, *own_start = nullptr;
  int32_t *cell_of = nullptr, *own_slots = nullptr;
  unsigned *count = nullptr, *own_count = nullptr, *fill = nullptr;
  Real4<R>* posj = nullptr;
  unsigned long long* tested = nullptr;  // distance tests of the last counted force launch (PSIM_OPT_KERNEL_TIMING)

  bool allocated() const { return key != nullptr; }

  void destroy() {
    cudaFree(key); cudaFree(perm); cudaFree(cell_start); cudaFree(own_start); cudaFree(cell_of);
    cudaFree(own_slots); cudaFree(count); cudaFree(own_count); cudaFree(fill); cudaFree(posj); cudaFree(tested);
    tested = nullptr;
    key = perm = cell_start = own_start = cell_of = own_slots = nullptr;
    count = own_count = fill = nullptr; posj = nullptr;
    binned = false;
  }

  // positions were replaced wholesale (or the algorithm was switched): re-bin before the next force pass
  void invalidate() { binned = false; }

  template <class S>
  cudaError_t ensure(S& sim) {
    if (allocated()) return cudaSuccess;
    n = sim.n; n_pad = sim.n_pad;
    ncell = sim.c.nc[0] * sim.c.nc[1] * sim.c.nc[2];
    cudaError_t e;
#define PSIM_TRY(x) if ((e = (x)) != cudaSuccess) return e
    PSIM_TRY(cudaMalloc(&key, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&perm, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&cell_start, sizeof(int32_t) * (ncell + 2)));
    PSIM_TRY(cudaMalloc(&own_start, sizeof(int32_t) * (ncell + 2)));
    PSIM_TRY(cudaMalloc(&count, sizeof(unsigned) * (ncell + 1)));
    PSIM_TRY(cudaMalloc(&own_count, sizeof(unsigned) * (ncell + 1)));
    PSIM_TRY(cudaMalloc(&fill, sizeof(unsigned) * (ncell + 1)));
    PSIM_TRY(cudaMalloc(&cell_of, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&own_slots, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&posj, sizeof(Real4<R>) * n_pad));
    PSIM_TRY(cudaMalloc(&tested, sizeof *tested));
    PSIM_TRY(cudaMemsetAsync(tested, 0, sizeof *tested, sim.stream));
    PSIM_TRY(cudaMemsetAsync(count, 0, sizeof(unsigned) * (ncell + 1), sim.stream));
    PSIM_TRY(cudaMemsetAsync(own_count, 0, sizeof(unsigned) * (ncell + 1), sim.stream));
    PSIM_TRY(cudaMemsetAsync(own_start, 0, sizeof(int32_t) * (ncell + 2), sim.stream));
#undef PSIM_TRY
    binned = false;
    return cudaSuccess;
  }

  void fill_args(HouseArgs<R>& h) const {
    h.ncell = ncell;
    h.key = key; h.perm = perm; h.cell_start = cell_start; h.own_start = own_start; h.cell_of = cell_of;
    h.own_slots = own_slots; h.count = count; h.own_count = own_count; h.fill = fill; h.posj = posj;
  }

  template <class S>
  psim_status force(S& sim, int mode) {
    if (ensure(sim) != cudaSuccess) return PSIM_ERR_CUDA;
    if (!binned) {                         // first use, new positions, or the algorithm was just selected
      psim_status st = sim.launch_housekeeping(sim.pos[sim.cur], sim.pos[sim.cur ^ 1], false, true);
      if (st != PSIM_OK) return st;
    }
    const StepArgs<R> a = sim.step_args();
    const int64_t n_own = sim.i_hi - sim.i_lo;
    if (n_own <= 0) return PSIM_OK;
    const unsigned grid = (unsigned)((n_own + TF - 1) / TF);
    unsigned long long* cnt = sim.timing ? tested : nullptr;      // counting costs a memset per launch: only on request
    if (cnt && cudaMemsetAsync(cnt, 0, sizeof *cnt, sim.stream) != cudaSuccess) return PSIM_ERR_CUDA;
    if (mode == MODE_FORCE)
      cell_force_kernel<R, MODE_FORCE, QCAP, TF><<<grid, TF, 0, sim.stream>>>(
          a, posj, cell_of, cell_start, own_slots, own_start, perm, ncell, cnt);
    else if (mode == MODE_FORCE_KICK)
      cell_force_kernel<R, MODE_FORCE_KICK, QCAP, TF><<<grid, TF, 0, sim.stream>>>(
          a, posj, cell_of, cell_start, own_slots, own_start, perm, ncell, cnt);
    else
      cell_force_kernel<R, MODE_FORCE_KICK_DRIFT, QCAP, TF><<<grid, TF, 0, sim.stream>>>(
          a, posj, cell_of, cell_start, own_slots, own_start, perm, ncell, cnt);
    ++launches;
    return PSIM_OK;
  }
};

}  // namespace psim
