// psim_cells.cuh — cell-binned force path.
//
// What the reference achieves with linked cells and per-particle neighbour
// lists (module_particle_management.f08:52-108, 128-137, 179-339), done the GPU
// way: whenever the reference would rebuild its lists (the displacement trigger
// of module_stoermer_verlet_method.f08:177-185 fired), the particle STATE is
// physically re-ordered by cell (x fastest, the reference's own grid:
// nint(area/cell) cells, index ceiling(x/cell), pm:62, 95), so that the 27-cell
// neighbourhood of a particle is nine contiguous runs of slots.  Between
// rebuilds the order is kept: a cell is at least r_cut + dr_max_tolerable wide
// (stoermer:75), the same skin argument that keeps the reference's lists valid.
//
// The force kernel gives one lane one particle and one warp 32 consecutive
// slots; see cell_force_kernel below.
//
// Everything here is conditional ON THE DEVICE: the re-binning kernels read
// the rebuild flag from the control block and return at once when it is clear,
// so the host never has to synchronise to learn whether a rebuild happened.
#pragma once

#include "psim_allpairs.cuh"
#include "psim_allpairs2.cuh"
#include "psim_integrate.cuh"

namespace psim {

template <class R>
__device__ __forceinline__ int clamp_cell(R x, R cell, int nc) {
  int c = (int)r_ceil(r_div(x, cell));
  return c < 1 ? 1 : (c > nc ? nc : c);
}

// K1: binned cell of every slot (inactive -> tail bucket ncell) + histogram.
template <class R, int T>
__global__ void __launch_bounds__(T)
cells_key_kernel(const Real4<R>* __restrict__ pos, int32_t* __restrict__ key, unsigned* __restrict__ count,
                 int64_t n, int ncell, Consts<R> c, const int* flag, int force) {
  if (!force && !*flag) return;
  const int64_t k = (int64_t)blockIdx.x * T + threadIdx.x;
  if (k >= n) return;
  const Real4<R> p = pos[k];
  int cell = ncell;
  if (!(p.w >= c.huge)) {
    const int cx = clamp_cell(p.x, c.cell[0], c.nc[0]);
    const int cy = clamp_cell(p.y, c.cell[1], c.nc[1]);
    const int cz = clamp_cell(p.z, c.cell[2], c.nc[2]);
    cell = (cx - 1) + c.nc[0] * ((cy - 1) + c.nc[1] * (cz - 1));
  }
  key[k] = cell;
  atomicAdd(&count[cell], 1u);
}

// K2: exclusive scan of the histogram (one block; ncell+1 buckets).
__global__ void __launch_bounds__(1024)
cells_scan_kernel(const unsigned* __restrict__ count, int32_t* __restrict__ cell_start, unsigned* __restrict__ fill,
                  int nbuckets, const int* flag, int force) {
  if (!force && !*flag) return;
  __shared__ unsigned s_sum[1024];
  const int tid = threadIdx.x;
  const int per = (nbuckets + 1023) / 1024;
  const int lo = tid * per, hi = min(nbuckets, lo + per);
  unsigned sum = 0;
  for (int b = lo; b < hi; ++b) sum += count[b];
  s_sum[tid] = sum;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {      // Hillis-Steele inclusive scan
    unsigned v = tid >= off ? s_sum[tid - off] : 0u;
    __syncthreads();
    s_sum[tid] += v;
    __syncthreads();
  }
  unsigned run = s_sum[tid] - sum;                // exclusive prefix of this thread's chunk
  for (int b = lo; b < hi; ++b) {
    cell_start[b] = (int32_t)run;
    run += count[b];
    fill[b] = 0u;
  }
  if (tid == 1023) cell_start[nbuckets] = (int32_t)s_sum[1023];
}

// K3: scatter slots into their cell segment (order inside a cell fixed by K4).
template <int T>
__global__ void __launch_bounds__(T)
cells_scatter_kernel(const int32_t* __restrict__ key, const int32_t* __restrict__ cell_start,
                     unsigned* __restrict__ fill, int32_t* __restrict__ src, int64_t n, const int* flag, int force) {
  if (!force && !*flag) return;
  const int64_t k = (int64_t)blockIdx.x * T + threadIdx.x;
  if (k >= n) return;
  const int cell = key[k];
  const unsigned at = atomicAdd(&fill[cell], 1u);
  src[cell_start[cell] + at] = (int32_t)k;
}

// K4: make the order inside every cell deterministic (ascending previous slot)
// and clear the histogram for the next rebuild.  One thread per bucket; the
// tail bucket of inactive particles is left unsorted (it never contributes).
template <int T>
__global__ void __launch_bounds__(T)
cells_order_kernel(const int32_t* __restrict__ cell_start, int32_t* __restrict__ src, unsigned* __restrict__ count,
                   int ncell, const int* flag, int force) {
  if (!force && !*flag) return;
  const int cell = blockIdx.x * T + threadIdx.x;
  if (cell > ncell) return;
  count[cell] = 0u;
  if (cell == ncell) return;
  const int lo = cell_start[cell], hi = cell_start[cell + 1];
  for (int a = lo + 1; a < hi; ++a) {
    const int32_t v = src[a];
    int b = a - 1;
    while (b >= lo && src[b] > v) { src[b + 1] = src[b]; --b; }
    src[b + 1] = v;
  }
}

template <class R>
struct CellArrays {
  Real4<R>* a[5];     // pos, vel, f_old, f_new, disp
  int32_t* orig;
  int32_t* cell_of;
};

// K5: gather the state into the spare set in the new order; K6: copy it back.
template <class R, int T>
__global__ void __launch_bounds__(T)
cells_gather_kernel(CellArrays<R> from, CellArrays<R> to, const int32_t* __restrict__ src,
                    const int32_t* __restrict__ key, int64_t n, const int* flag, int force) {
  if (!force && !*flag) return;
  const int64_t k = (int64_t)blockIdx.x * T + threadIdx.x;
  if (k >= n) return;
  const int32_t s = src[k];
#pragma unroll
  for (int q = 0; q < 5; ++q) to.a[q][k] = from.a[q][s];
  to.orig[k] = from.orig[s];
  to.cell_of[k] = key[s];
}

template <class R, int T>
__global__ void __launch_bounds__(T)
cells_copyback_kernel(CellArrays<R> from, CellArrays<R> to, int64_t n, const int* flag, int force) {
  if (!force && !*flag) return;
  const int64_t k = (int64_t)blockIdx.x * T + threadIdx.x;
  if (k >= n) return;
#pragma unroll
  for (int q = 0; q < 5; ++q) to.a[q][k] = from.a[q][k];
  to.orig[k] = from.orig[k];
  to.cell_of[k] = from.cell_of[k];
}

__global__ void iota_kernel(int32_t* p, int64_t n) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) p[k] = (int32_t)k;
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
// INDEXED = false: the state itself is in cell order (single GPU): slot k holds
// particle k of the permuted arrays and posj is a.pos_cur.
// INDEXED = true (sharded runs): the state stays in particle order; posj is a
// cell-ordered COPY of all positions, perm maps slot -> particle, and the block
// walks own_slots, the cell-ordered list of the slots whose particle this rank
// owns.  The force sum uses slots, the kick/drift epilogue the particle index.
template <class R, int MODE, int QCAP, int T, bool INDEXED>
__global__ void __launch_bounds__(T)
cell_force_kernel(const StepArgs<R> a, const Real4<R>* __restrict__ posj, const int32_t* __restrict__ cell_of,
                  const int32_t* __restrict__ cell_start, const int32_t* __restrict__ own_slots,
                  const int32_t* __restrict__ own_start, const int32_t* __restrict__ perm, int ncell, int64_t n) {
  constexpr int W = T / 32, TJ = 32;
  __shared__ int32_t queue[QCAP * T];
  __shared__ __align__(16) float tx[W][2][TJ], ty[W][2][TJ], tz[W][2][TJ], tw[W][2][TJ];
  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t idx = (int64_t)blockIdx.x * T + tid;
  int64_t k = -1;              // slot in cell order (j-side numbering)
  int64_t ip = -1;             // particle index in the state arrays (epilogue numbering)
  if (INDEXED) {
    if (idx < own_start[ncell]) { k = own_slots[idx]; ip = perm[k]; }
  } else if (idx < n) {
    k = idx; ip = idx;
  }
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
  }

  R first = (R)0, second = (R)0;
  if (ip >= 0) step_epilogue<R, MODE>(a, ip, fx, fy, fz, first, second);
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, T>(first, second, a.top2_part);
}

// ---- sharded runs: binning by index, state stays in particle order ------------

// S1: cell of every PARTICLE (inactive -> tail bucket) + histograms of all and of owned particles.
template <class R, int T>
__global__ void __launch_bounds__(T)
cellsx_key_kernel(const Real4<R>* __restrict__ pos, int32_t* __restrict__ key, unsigned* __restrict__ count,
                  unsigned* __restrict__ own_count, int64_t n, int64_t i_lo, int64_t i_hi, int ncell, Consts<R> c,
                  const int* flag, int force) {
  if (!force && !*flag) return;
  const int64_t i = (int64_t)blockIdx.x * T + threadIdx.x;
  if (i >= n) return;
  const Real4<R> p = pos[i];
  int cell = ncell;
  if (!(p.w >= c.huge)) {
    const int cx = clamp_cell(p.x, c.cell[0], c.nc[0]);
    const int cy = clamp_cell(p.y, c.cell[1], c.nc[1]);
    const int cz = clamp_cell(p.z, c.cell[2], c.nc[2]);
    cell = (cx - 1) + c.nc[0] * ((cy - 1) + c.nc[1] * (cz - 1));
  }
  key[i] = cell;
  atomicAdd(&count[cell], 1u);
  if (i >= i_lo && i < i_hi && cell < ncell) atomicAdd(&own_count[cell], 1u);
}

// S4: per cell, put the slot -> particle map in ascending particle order (deterministic),
// record the cell of every slot, and list the slots of the owned particles in cell order.
template <int T>
__global__ void __launch_bounds__(T)
cellsx_order_kernel(const int32_t* __restrict__ cell_start, const int32_t* __restrict__ own_start,
                    int32_t* __restrict__ perm, int32_t* __restrict__ cell_of, int32_t* __restrict__ own_slots,
                    unsigned* __restrict__ count, unsigned* __restrict__ own_count, int64_t i_lo, int64_t i_hi,
                    int ncell, const int* flag, int force) {
  if (!force && !*flag) return;
  const int cell = blockIdx.x * T + threadIdx.x;
  if (cell > ncell) return;
  count[cell] = 0u;
  own_count[cell] = 0u;
  const int lo = cell_start[cell], hi = cell_start[cell + 1];
  if (cell == ncell) {                     // inactive tail: never read through the cell tables
    for (int a = lo; a < hi; ++a) cell_of[a] = ncell;
    return;
  }
  for (int a = lo + 1; a < hi; ++a) {
    const int32_t v = perm[a];
    int b = a - 1;
    while (b >= lo && perm[b] > v) { perm[b + 1] = perm[b]; --b; }
    perm[b + 1] = v;
  }
  int m = own_start[cell];
  for (int a = lo; a < hi; ++a) {
    cell_of[a] = cell;
    const int32_t i = perm[a];
    if (i >= i_lo && i < i_hi) own_slots[m++] = a;
  }
}

// every step: the cell-ordered copy of all positions
template <class R, int T>
__global__ void __launch_bounds__(T)
cellsx_gather_kernel(Real4<R>* __restrict__ posj, const Real4<R>* __restrict__ pos,
                     const int32_t* __restrict__ perm, int64_t n) {
  const int64_t k = (int64_t)blockIdx.x * T + threadIdx.x;
  if (k < n) posj[k] = pos[perm[k]];
}

// every fused step of a sharded run: owned INACTIVE particles are not in own_slots, so their
// (frozen) record is carried into the next position buffer here
template <class R, int T>
__global__ void __launch_bounds__(T)
cellsx_carry_inactive_kernel(const Real4<R>* __restrict__ pos_cur, Real4<R>* __restrict__ pos_nxt,
                             int64_t i_lo, int64_t i_hi, Consts<R> c) {
  const int64_t i = i_lo + (int64_t)blockIdx.x * T + threadIdx.x;
  if (i >= i_hi) return;
  const Real4<R> p = pos_cur[i];
  if (p.w >= c.huge) pos_nxt[i] = p;
}

template <class R>
struct CellPath {
  static constexpr int T = 128;          // threads of the re-binning kernels
  static constexpr int TF = 64;          // threads of the force kernel (two autonomous warps)
  static constexpr int QCAP = 96;        // queue entries per lane
  int64_t launches = 0;
  int64_t n = 0, n_pad = 0;
  int ncell = 0;
  bool binned = false;
  int32_t *key = nullptr, *src = nullptr, *cell_start = nullptr;
  unsigned *count = nullptr, *fill = nullptr;
  int32_t *orig = nullptr, *cell_of = nullptr;
  CellArrays<R> spare = {};
  // indexed (sharded) mode: state stays in particle order
  bool indexed = false;
  Real4<R>* posj = nullptr;
  int32_t *own_slots = nullptr, *own_start = nullptr;
  unsigned* own_count = nullptr;

  bool allocated() const { return key != nullptr; }
  const int32_t* orig_or_null() const { return indexed ? nullptr : orig; }

  void destroy() {
    cudaFree(key); cudaFree(src); cudaFree(cell_start); cudaFree(count); cudaFree(fill);
    cudaFree(orig); cudaFree(cell_of);
    for (int q = 0; q < 5; ++q) cudaFree(spare.a[q]);
    cudaFree(spare.orig); cudaFree(spare.cell_of);
    cudaFree(posj); cudaFree(own_slots); cudaFree(own_start); cudaFree(own_count);
    key = src = cell_start = nullptr; count = fill = nullptr; orig = cell_of = nullptr;
    posj = nullptr; own_slots = own_start = nullptr; own_count = nullptr;
    spare = CellArrays<R>();
    binned = false;
  }

  // positions were replaced wholesale: the next force pass must re-bin
  void invalidate() { binned = false; }

  template <class S>
  cudaError_t ensure(S& sim) {
    if (allocated()) return cudaSuccess;
    n = sim.n; n_pad = sim.n_pad;
    ncell = sim.c.nc[0] * sim.c.nc[1] * sim.c.nc[2];
    indexed = sim.world > 1 || sim.cells_indexed;
    cudaError_t e;
#define PSIM_TRY(x) if ((e = (x)) != cudaSuccess) return e
    if (indexed) {
      PSIM_TRY(cudaMalloc(&key, sizeof(int32_t) * n_pad));
      PSIM_TRY(cudaMalloc(&src, sizeof(int32_t) * n_pad));            // perm: slot -> particle
      PSIM_TRY(cudaMalloc(&cell_start, sizeof(int32_t) * (ncell + 2)));
      PSIM_TRY(cudaMalloc(&own_start, sizeof(int32_t) * (ncell + 2)));
      PSIM_TRY(cudaMalloc(&count, sizeof(unsigned) * (ncell + 1)));
      PSIM_TRY(cudaMalloc(&own_count, sizeof(unsigned) * (ncell + 1)));
      PSIM_TRY(cudaMalloc(&fill, sizeof(unsigned) * (ncell + 1)));
      PSIM_TRY(cudaMalloc(&cell_of, sizeof(int32_t) * n_pad));
      PSIM_TRY(cudaMalloc(&own_slots, sizeof(int32_t) * n_pad));
      PSIM_TRY(cudaMalloc(&posj, sizeof(Real4<R>) * n_pad));
      PSIM_TRY(cudaMemsetAsync(count, 0, sizeof(unsigned) * (ncell + 1), sim.stream));
      PSIM_TRY(cudaMemsetAsync(own_count, 0, sizeof(unsigned) * (ncell + 1), sim.stream));
      binned = false;
      return cudaSuccess;
    }
    PSIM_TRY(cudaMalloc(&key, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&src, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&cell_start, sizeof(int32_t) * (ncell + 2)));
    PSIM_TRY(cudaMalloc(&count, sizeof(unsigned) * (ncell + 1)));
    PSIM_TRY(cudaMalloc(&fill, sizeof(unsigned) * (ncell + 1)));
    PSIM_TRY(cudaMalloc(&orig, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&cell_of, sizeof(int32_t) * n_pad));
    for (int q = 0; q < 5; ++q) PSIM_TRY(cudaMalloc(&spare.a[q], sizeof(Real4<R>) * n_pad));
    PSIM_TRY(cudaMalloc(&spare.orig, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMalloc(&spare.cell_of, sizeof(int32_t) * n_pad));
    PSIM_TRY(cudaMemsetAsync(count, 0, sizeof(unsigned) * (ncell + 1), sim.stream));
    PSIM_TRY(cudaMemsetAsync(cell_of, 0, sizeof(int32_t) * n_pad, sim.stream));
#undef PSIM_TRY
    iota_kernel<<<(unsigned)((n_pad + 255) / 256), 256, 0, sim.stream>>>(orig, n_pad);
    ++launches;
    binned = false;
    return cudaSuccess;
  }

  // Re-order the state held in (posbuf, vel, f_old, f_new, disp) by cell, if
  // the rebuild flag of this step is set (or unconditionally with force).
  template <class S>
  void rebin(S& sim, Real4<R>* posbuf, int parity, bool force) {
    const int* flag = &sim.ctrl->flag[parity];
    const int f = force ? 1 : 0;
    const unsigned gn = (unsigned)((n + T - 1) / T);
    cudaStream_t st = sim.stream;
    if (indexed) {
      cellsx_key_kernel<R, T><<<gn, T, 0, st>>>(posbuf, key, count, own_count, n, sim.i_lo, sim.i_hi, ncell, sim.c, flag, f);
      cells_scan_kernel<<<1, 1024, 0, st>>>(count, cell_start, fill, ncell + 1, flag, f);
      cells_scan_kernel<<<1, 1024, 0, st>>>(own_count, own_start, fill, ncell + 1, flag, f);
      cells_scatter_kernel<T><<<gn, T, 0, st>>>(key, cell_start, fill, src, n, flag, f);
      cellsx_order_kernel<T><<<(unsigned)((ncell + 1 + T - 1) / T), T, 0, st>>>(
          cell_start, own_start, src, cell_of, own_slots, count, own_count, sim.i_lo, sim.i_hi, ncell, flag, f);
      launches += 5;
      if (force) binned = true;
      return;
    }
    cells_key_kernel<R, T><<<gn, T, 0, st>>>(posbuf, key, count, n, ncell, sim.c, flag, f);
    cells_scan_kernel<<<1, 1024, 0, st>>>(count, cell_start, fill, ncell + 1, flag, f);
    cells_scatter_kernel<T><<<gn, T, 0, st>>>(key, cell_start, fill, src, n, flag, f);
    cells_order_kernel<T><<<(unsigned)((ncell + 1 + T - 1) / T), T, 0, st>>>(cell_start, src, count, ncell, flag, f);
    CellArrays<R> live;
    live.a[0] = posbuf; live.a[1] = sim.vel; live.a[2] = sim.f_old; live.a[3] = sim.f_new; live.a[4] = sim.disp;
    live.orig = orig; live.cell_of = cell_of;
    cells_gather_kernel<R, T><<<gn, T, 0, st>>>(live, spare, src, key, n, flag, f);
    cells_copyback_kernel<R, T><<<gn, T, 0, st>>>(spare, live, n, flag, f);
    launches += 6;
    if (force) binned = true;
  }

  template <class S>
  psim_status force(S& sim, int mode) {
    if (ensure(sim) != cudaSuccess) return PSIM_ERR_CUDA;
    if (!binned) rebin(sim, sim.pos[sim.cur], 0, true);
    const StepArgs<R> a = sim.step_args();
    if (indexed) {
      const int64_t n_own = sim.i_hi - sim.i_lo;
      if (n_own <= 0) { if (mode == MODE_FORCE_KICK_DRIFT) sim.top2_count = 0; return PSIM_OK; }
      const unsigned gn = (unsigned)((n + T - 1) / T);
      cellsx_gather_kernel<R, T><<<gn, T, 0, sim.stream>>>(posj, a.pos_cur, src, n);
      const unsigned grid = (unsigned)((n_own + TF - 1) / TF);
      if (mode == MODE_FORCE)
        cell_force_kernel<R, MODE_FORCE, QCAP, TF, true><<<grid, TF, 0, sim.stream>>>(
            a, posj, cell_of, cell_start, own_slots, own_start, src, ncell, n);
      else if (mode == MODE_FORCE_KICK)
        cell_force_kernel<R, MODE_FORCE_KICK, QCAP, TF, true><<<grid, TF, 0, sim.stream>>>(
            a, posj, cell_of, cell_start, own_slots, own_start, src, ncell, n);
      else {
        cellsx_carry_inactive_kernel<R, T><<<(unsigned)((n_own + T - 1) / T), T, 0, sim.stream>>>(
            a.pos_cur, a.pos_nxt, sim.i_lo, sim.i_hi, sim.c);
        cell_force_kernel<R, MODE_FORCE_KICK_DRIFT, QCAP, TF, true><<<grid, TF, 0, sim.stream>>>(
            a, posj, cell_of, cell_start, own_slots, own_start, src, ncell, n);
        sim.top2_count = grid;
        ++launches;
      }
      launches += 2;
      return PSIM_OK;
    }
    const unsigned grid = (unsigned)((n + TF - 1) / TF);
    if (mode == MODE_FORCE)
      cell_force_kernel<R, MODE_FORCE, QCAP, TF, false><<<grid, TF, 0, sim.stream>>>(
          a, a.pos_cur, cell_of, cell_start, nullptr, nullptr, nullptr, ncell, n);
    else if (mode == MODE_FORCE_KICK)
      cell_force_kernel<R, MODE_FORCE_KICK, QCAP, TF, false><<<grid, TF, 0, sim.stream>>>(
          a, a.pos_cur, cell_of, cell_start, nullptr, nullptr, nullptr, ncell, n);
    else {
      cell_force_kernel<R, MODE_FORCE_KICK_DRIFT, QCAP, TF, false><<<grid, TF, 0, sim.stream>>>(
          a, a.pos_cur, cell_of, cell_start, nullptr, nullptr, nullptr, ncell, n);
      sim.top2_count = grid;
    }
    ++launches;
    return PSIM_OK;
  }
};

}  // namespace psim
