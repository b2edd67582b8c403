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
// The force kernel gives one thread one particle.  Phase 1 scans the nine runs
// with the 3-FMA pre-filter and appends the few survivors to a per-thread queue
// in shared memory (in-range pairs are ~6 % of the neighbourhood, far too dense
// for a divergent branch per pair); phase 2 walks the queue with the exact pair
// term.  Sums run in ascending slot order: deterministic.
//
// Everything here is conditional ON THE DEVICE: the re-binning kernels read
// the rebuild flag from the control block and return at once when it is clear,
// so the host never has to synchronise to learn whether a rebuild happened.
#pragma once

#include "psim_allpairs.cuh"
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
// A block owns T consecutive slots, i.e. the particles of a short run of cells
// cmin..cmax.  For each of the nine (dy, dz) offsets the union of its threads'
// neighbour cells is ONE contiguous run of slots (cells cmin-1+off ..
// cmax+1+off in linear cell order; a run that wraps around a grid edge only
// adds far-away particles, which the filter rejects).  The block stages that
// run into shared memory in chunks of CAP records and every thread scans the
// part of it that belongs to its own three cells.
template <class R, int T, int MODE, int QCAP, int CAP>
__global__ void __launch_bounds__(T)
cell_force_kernel(const StepArgs<R> a, const int32_t* __restrict__ cell_of,
                  const int32_t* __restrict__ cell_start, int ncell, int64_t n) {
  __shared__ int32_t queue[QCAP * T];
  __shared__ Real4<R> tile[CAP];
  __shared__ int s_cmin, s_cmax;
  const Consts<R>& c = a.c;
  const int tid = threadIdx.x;
  const int64_t k = (int64_t)blockIdx.x * T + tid;

  R xi = 0, yi = 0, zi = 0, ax = 0, ay = 0, az = 0, cut = -c.huge;
  R fx = 0, fy = 0, fz = 0;
  int cell = ncell;
  if (k < n) {
    cell = cell_of[k];
    const Real4<R> p = a.pos_cur[k];
    if (cell < ncell && !(p.w >= c.huge)) {
      xi = p.x; yi = p.y; zi = p.z;
      ax = (R)-2 * xi; ay = (R)-2 * yi; az = (R)-2 * zi;
      cut = c.cut2_pad - p.w;
    } else {
      cell = ncell;
    }
  }
  if (tid == 0) { s_cmin = ncell; s_cmax = -1; }
  __syncthreads();
  if (cell < ncell) { atomicMin(&s_cmin, cell); atomicMax(&s_cmax, cell); }
  __syncthreads();
  const int cmin = s_cmin, cmax = s_cmax;

  int cnt = 0;
  auto flush = [&]() {
    if (cnt == 0) return;
    Real4<R> q = a.pos_cur[queue[tid]];
    for (int e = 0; e < cnt; ++e) {
      const Real4<R> qc = q;
      if (e + 1 < cnt) q = a.pos_cur[queue[(e + 1) * T + tid]];     // next record in flight
      pair_force(xi, yi, zi, qc.x, qc.y, qc.z, c, fx, fy, fz);
    }
    cnt = 0;
  };

  if (cmax >= 0) {                       // the block has at least one active particle
    const int nx = c.nc[0], ny = c.nc[1], nz = c.nc[2];
    const int cx = cell % nx, cy = (cell / nx) % ny, cz = cell / (nx * ny);   // 0-based (unused if inactive)
    const int x0 = max(cx - 1, 0), x1 = min(cx + 1, nx - 1);
    for (int dz = -1; dz <= 1; ++dz) {
      for (int dy = -1; dy <= 1; ++dy) {
        const int off = nx * (dy + ny * dz);
        const int blo = max(cmin - 1 + off, 0), bhi = min(cmax + 1 + off, ncell - 1);
        if (bhi < blo) continue;         // block-uniform
        const int ta_all = cell_start[blo], tb_all = cell_start[bhi + 1];
        // this thread's own run inside the block's run
        int ja = 0, jb = 0;
        const int y = cy + dy, z = cz + dz;
        if (cell < ncell && y >= 0 && y < ny && z >= 0 && z < nz) {
          const int row = nx * (y + ny * z);
          ja = cell_start[row + x0];
          jb = cell_start[row + x1 + 1];
        }
        for (int ta = ta_all; ta < tb_all; ta += CAP) {
          const int tb = min(ta + CAP, tb_all);
          __syncthreads();               // previous chunk fully consumed
          for (int t = ta + tid; t < tb; t += T) tile[t - ta] = a.pos_cur[t];
          __syncthreads();
          const int lo = max(ja, ta), hi = min(jb, tb);
          for (int j = lo; j < hi; j += 4) {
            Real4<R> q[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) q[u] = tile[min(j + u, hi - 1) - ta];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const R est = r_fma(ax, q[u].x, r_fma(ay, q[u].y, r_fma(az, q[u].z, q[u].w)));
              if (est < cut && j + u < hi && j + u != (int)k) {
                queue[cnt * T + tid] = j + u;
                if (++cnt == QCAP) flush();
              }
            }
          }
        }
      }
    }
    flush();
  }

  R first = (R)0, second = (R)0;
  if (k < n) step_epilogue<R, MODE>(a, k, fx, fy, fz, first, second);
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, T>(first, second, a.top2_part);
}

template <class R>
struct CellPath {
  static constexpr int T = 128;
  static constexpr int QCAP = 64;
  static constexpr int CAP = 256;
  int64_t launches = 0;
  int64_t n = 0, n_pad = 0;
  int ncell = 0;
  bool binned = false;
  int32_t *key = nullptr, *src = nullptr, *cell_start = nullptr;
  unsigned *count = nullptr, *fill = nullptr;
  int32_t *orig = nullptr, *cell_of = nullptr;
  CellArrays<R> spare = {};

  bool allocated() const { return orig != nullptr; }
  const int32_t* orig_or_null() const { return orig; }

  void destroy() {
    cudaFree(key); cudaFree(src); cudaFree(cell_start); cudaFree(count); cudaFree(fill);
    cudaFree(orig); cudaFree(cell_of);
    for (int q = 0; q < 5; ++q) cudaFree(spare.a[q]);
    cudaFree(spare.orig); cudaFree(spare.cell_of);
    key = src = cell_start = nullptr; count = fill = nullptr; orig = cell_of = nullptr;
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
    cudaError_t e;
#define PSIM_TRY(x) if ((e = (x)) != cudaSuccess) return e
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
    const unsigned grid = (unsigned)((n + T - 1) / T);
    if (mode == MODE_FORCE)
      cell_force_kernel<R, T, MODE_FORCE, QCAP, CAP><<<grid, T, 0, sim.stream>>>(a, cell_of, cell_start, ncell, n);
    else if (mode == MODE_FORCE_KICK)
      cell_force_kernel<R, T, MODE_FORCE_KICK, QCAP, CAP><<<grid, T, 0, sim.stream>>>(a, cell_of, cell_start, ncell, n);
    else {
      cell_force_kernel<R, T, MODE_FORCE_KICK_DRIFT, QCAP, CAP><<<grid, T, 0, sim.stream>>>(a, cell_of, cell_start, ncell, n);
      sim.top2_count = grid;
    }
    ++launches;
    return PSIM_OK;
  }
};

}  // namespace psim
