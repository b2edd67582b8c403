// psim_static_scratch.cu — psim_compute_locations_static: the drift step
// (compute_locations) for callers that must not allocate device memory.
//
// Who this is for.  Every other entry point of this library sets up a psim_sim
// handle, and that handle obtains its working buffers from cudaMalloc.  Some
// hosts cannot live with that:
//   * a driver that has already handed the whole GPU to its own memory plan
//     (an OpenACC / CUDA Fortran host with its particle arrays resident on the
//     device, a job running under a fixed per-process memory budget) and
//     refuses hidden allocations from a library it links;
//   * a latency-critical stepping loop, where cudaMalloc / cudaFree are
//     unwelcome because they synchronise the whole device.
// For them this file offers ONE entry point that works purely on arrays the
// caller already owns on the device.  The only extra memory it needs — the
// per-block partial results of the "two largest displacements" reduction and a
// block counter — lives in a fixed array that is part of the library image.
//
// Cost and limits of that choice:
//   * the array is PSIM_STATIC_SCRATCH_BYTES (about 64 KiB) and every GPU that
//     loads this library keeps it until the process exits;
//   * at most PSIM_STATIC_MAX_PARTICLES (1,048,576) particles per call; larger
//     requests are refused with PSIM_ERR_INVALID_ARGUMENT, nothing is launched;
//   * all calls on one GPU share the array.  Two calls whose kernels can overlap
//     on that GPU (different streams, different host threads) race on it;
//     issue them on one stream or order them yourself.  The array keeps
//     whatever the previous call left; the part that has to start from zero
//     (the block counter) is cleared on the stream at the start of every call.
//   * the first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress) and caches it.  Make that first call during
//     set-up, before any stream capture is begun.
//   * no cudaMalloc, cudaFree or other allocating routine is called, ever.

#include <cuda_runtime.h>

#include <mutex>

#include "psim_allpairs.cuh"

namespace psim {
namespace {

constexpr int kThreads = 256;
constexpr long long kMaxParticles = PSIM_STATIC_MAX_PARTICLES;
constexpr int kMaxBlocks = (int)(kMaxParticles / kThreads);             // 4096
constexpr int kCtrlBytes = 256;                                         // block counter lives here
constexpr int kScratchBytes = kCtrlBytes + kMaxBlocks * 2 * (int)sizeof(double);
static_assert(kScratchBytes == PSIM_STATIC_SCRATCH_BYTES, "header and implementation disagree");

// The scratch itself.  In HOST code this name is only a handle for the runtime
// (it is not a device address): host code must go through
// cudaGetSymbolAddress, see scratch_for_device().
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {nullptr};

// Device address of g_static_scratch on the CURRENT device, resolved once per device.
cudaError_t scratch_for_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_ptr[device];
  return cudaSuccess;
}

// One drift step on packed xyz arrays (module_stoermer_verlet_method.f08:136-187),
// same arithmetic as drift_kernel; the per-block two largest |displacement|
// go to `scratch`, and the block that finishes last merges them and writes the
// rebuild verdict (:177-185) to *flag_out.
template <class R>
__global__ void __launch_bounds__(kThreads)
compute_locations_static_kernel(R* __restrict__ pos, const R* __restrict__ vel, R* __restrict__ f_old,
                                R* __restrict__ f_new, R* __restrict__ disp,
                                const unsigned char* __restrict__ active, long long n, Consts<R> c,
                                unsigned char* scratch, int* flag_out) {
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
  R* part = reinterpret_cast<R*>(scratch + kCtrlBytes);
  __shared__ bool s_last;
  __shared__ R s_first[kThreads / 32], s_second[kThreads / 32];

  const long long i = (long long)blockIdx.x * kThreads + threadIdx.x;
  R first = (R)0, second = (R)0;
  if (i < n && (!active || active[i])) {
    Real4<R> x, v, d, xn;
    x.x = pos[3 * i]; x.y = pos[3 * i + 1]; x.z = pos[3 * i + 2]; x.w = (R)0;
    v.x = vel[3 * i]; v.y = vel[3 * i + 1]; v.z = vel[3 * i + 2]; v.w = (R)0;
    d.x = disp[3 * i]; d.y = disp[3 * i + 1]; d.z = disp[3 * i + 2]; d.w = (R)0;
    const R fx = f_new[3 * i], fy = f_new[3 * i + 1], fz = f_new[3 * i + 2];
    f_old[3 * i] = fx; f_old[3 * i + 1] = fy; f_old[3 * i + 2] = fz;          // :155
    f_new[3 * i] = (R)0; f_new[3 * i + 1] = (R)0; f_new[3 * i + 2] = (R)0;    // :157
    first = drift_one(c, x, v, fx, fy, fz, d, xn);
    pos[3 * i] = xn.x; pos[3 * i + 1] = xn.y; pos[3 * i + 2] = xn.z;
    disp[3 * i] = d.x; disp[3 * i + 1] = d.y; disp[3 * i + 2] = d.z;
  }
  // block-level two largest values
  top2_warp_reduce(first, second);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_first[warp] = first; s_second[warp] = second; }
  __syncthreads();
  if (warp == 0) {
    first = lane < kThreads / 32 ? s_first[lane] : (R)0;
    second = lane < kThreads / 32 ? s_second[lane] : (R)0;
    top2_warp_reduce(first, second);
    if (lane == 0) {
      part[2 * blockIdx.x] = first;
      part[2 * blockIdx.x + 1] = second;
      __threadfence();
      const unsigned int ticket = atomicAdd(counter, 1u);
      s_last = (ticket == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (!s_last) return;
  // last block: every partial is visible now
  __threadfence();
  first = second = (R)0;
  for (unsigned int b = threadIdx.x; b < gridDim.x; b += kThreads) {
    const volatile R* vp = part;
    top2_push(first, second, vp[2 * b]);
    top2_push(first, second, vp[2 * b + 1]);
  }
  top2_warp_reduce(first, second);
  __syncthreads();
  if (lane == 0) { s_first[warp] = first; s_second[warp] = second; }
  __syncthreads();
  if (warp == 0) {
    first = lane < kThreads / 32 ? s_first[lane] : (R)0;
    second = lane < kThreads / 32 ? s_second[lane] : (R)0;
    top2_warp_reduce(first, second);
    if (lane == 0) {
      *flag_out = (r_add(first, second) > c.dr_max_tolerable) ? 1 : 0;
      *counter = 0u;   // leave the counter clean for the next call
    }
  }
}

// The integrator constants this kernel needs (stoermer:68-81), in R.
template <class R>
Consts<R> locations_consts(const psim_params& p) {
  Consts<R> c = {};
  const R sigma = (R)p.sigma, mass = (R)p.mass;
  c.delta_t = (R)p.delta_t;
  c.r_cut = (R)p.r_cut_factor * sigma;
  for (int d = 0; d < 3; ++d) { c.area[d] = (R)p.length_of_area[d]; c.cell[d] = (R)p.length_of_cell[d]; }
  c.min_cell = c.cell[0] < c.cell[1] ? c.cell[0] : c.cell[1];
  if (c.cell[2] < c.min_cell) c.min_cell = c.cell[2];
  c.dr_max_tolerable = c.min_cell - c.r_cut;
  c.tmf = (R)0.5 * c.delta_t / mass;
  c.one_minus_k = (R)1;
  c.huge = RealTraits<R>::huge;
  c.slices = 1;
  return c;
}

template <class R>
psim_status run(const psim_params& p, long long n, void* pos, const void* vel, void* f_old, void* f_new,
                void* disp, const unsigned char* active, int* flag, cudaStream_t stream,
                unsigned char* scratch) {
  const Consts<R> c = locations_consts<R>(p);
  cudaError_t e = cudaMemsetAsync(scratch, 0, kCtrlBytes, stream);   // the block counter must start at zero
  if (e != cudaSuccess) { cudaGetLastError(); return PSIM_ERR_CUDA; }
  const unsigned int grid = (unsigned int)((n + kThreads - 1) / kThreads);
  compute_locations_static_kernel<R><<<grid, kThreads, 0, stream>>>(
      (R*)pos, (const R*)vel, (R*)f_old, (R*)f_new, (R*)disp, active, n, c, scratch, flag);
  e = cudaGetLastError();
  if (e != cudaSuccess) return PSIM_ERR_CUDA;
  return PSIM_OK;
}

}  // namespace
}  // namespace psim

extern "C" psim_status psim_compute_locations_static(const psim_params* params, int precision, int64_t n,
                                                     void* pos_dev, const void* vel_dev, void* force_old_dev,
                                                     void* force_new_dev, void* displacement_dev,
                                                     const uint8_t* active_dev, int32_t* rebuild_flag_dev,
                                                     void* cuda_stream) {
  using namespace psim;
  if (!params || !pos_dev || !vel_dev || !force_old_dev || !force_new_dev || !displacement_dev || !rebuild_flag_dev)
    return PSIM_ERR_INVALID_ARGUMENT;
  if (precision != 4 && precision != 8) return PSIM_ERR_INVALID_ARGUMENT;
  if (n <= 0) return PSIM_ERR_NO_PARTICLES;
  if (n > kMaxParticles) return PSIM_ERR_INVALID_ARGUMENT;   // beyond what the fixed scratch can serve
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) { cudaGetLastError(); return PSIM_ERR_NO_DEVICE; }
  if (device < 0 || device >= kMaxDevices) return PSIM_ERR_UNSUPPORTED;
  unsigned char* scratch = nullptr;
  e = scratch_for_device(device, &scratch);
  if (e != cudaSuccess) { cudaGetLastError(); return PSIM_ERR_CUDA; }
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  if (precision == 4)
    return run<float>(*params, n, pos_dev, vel_dev, force_old_dev, force_new_dev, displacement_dev, active_dev,
                      rebuild_flag_dev, stream, scratch);
  return run<double>(*params, n, pos_dev, vel_dev, force_old_dev, force_new_dev, displacement_dev, active_dev,
                     rebuild_flag_dev, stream, scratch);
}
