// tc_rtt.cu — the accumulator round trip of the tensor-core all-pairs kernel (csrc/psim_allpairs_tc.cuh), in
// isolation: MMA thread -> tensor pipe -> commit -> consumers wake -> tcgen05.ld -> release -> MMA thread wakes.
// Operands sit in shared memory (zeros; no TMA), so what is measured is the pipeline's own latency and how the
// number of accumulators, the tile width, the consumers per tile and the reduction change the cycles per tile.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tc_rtt tools/microbench/tc_rtt.cu && ./tc_rtt
//
// Part E: latency probe (one accumulator): the stages of one round trip, averaged.
// Part F: cycles per 128 x 128 pair tests for the kernel's structure and for variations of it.
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>

static constexpr int M = 128;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long make_desc(unsigned addr) {
  const unsigned long long lbo = 128 >> 4, sbo = 256 >> 4;
  return (unsigned long long)((addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
}
__device__ __forceinline__ void mbar_init(unsigned a, unsigned n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(a), "r"(n)); }
__device__ __forceinline__ void mbar_arrive(unsigned a) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(a) : "memory"); }
__device__ __forceinline__ void mbar_wait(unsigned addr, unsigned parity) {
  unsigned done = 0;
  while (!done)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
}
// WAIT 0: every lane polls (the kernel of round 2 so far); 1: lane 0 polls, the vote releases the warp;
// 2: every lane, with a long suspend-time hint; 3: lane 0 with the hint.
template <int WAIT>
__device__ __forceinline__ void mbar_wait_warp(unsigned addr, unsigned parity) {
  unsigned done;
  const bool poller = (WAIT & 1) ? (threadIdx.x & 31) == 0 : true;
  do {
    done = 1u;
    if (poller) {
      if (WAIT & 2)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(addr), "r"(parity), "r"(1000000u) : "memory");
      else
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    }
  } while (!__all_sync(0xffffffffu, done));
}
#define LD32P(r, taddr)                                                                                            \
  asm volatile(                                                                                                    \
      "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "                                                          \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                    \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"                  \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),            \
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),      \
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),    \
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])     \
      : "r"(taddr))

__device__ __forceinline__ unsigned or_tree(const unsigned (&r)[32]) {
  unsigned g0 = r[0], g1 = r[8], g2 = r[16], g3 = r[24];
#pragma unroll
  for (int i = 1; i < 8; ++i) { g0 |= r[i]; g1 |= r[8 + i]; g2 |= r[16 + i]; g3 |= r[24 + i]; }
  return g0 | g1 | g2 | g3;
}

// NB accumulators of NT columns (NB * NT <= 512); CW consumer warps, CPT of them per tile (4, 8 or 16), each its
// 32 rows and NT / (CPT / 4) columns (64 or 128); CW / CPT consumer groups, group g owns the tiles t = g (mod
// groups), and NB is a multiple of the group count, so an accumulator always meets the same group (every waiter
// sees every phase).  NMMA issuing threads (NB a multiple): thread w issues the tiles t = w (mod NMMA), in order.
// REDUCE = 1: OR tree + sign test after the release, as in the kernel.  PROBE = 1 (NB = 1): stage latencies.
template <int NB, int NT, int CW, int CPT, int NMMA, int REDUCE, int PROBE, int WAIT>
__global__ void __launch_bounds__(32 * (CW + NMMA), 1)
rtt_kernel(long long* cycles, unsigned* sink, int ntiles, long long* lat) {
  constexpr int G = CW / CPT, PARTS = CPT / 4, WIDTH = NT / PARTS;
  constexpr unsigned TCOLS = NB * NT <= 32 ? 32u : NB * NT <= 64 ? 64u : NB * NT <= 128 ? 128u : NB * NT <= 256 ? 256u : 512u;
  static_assert(NB * NT <= 512 && NB % G == 0 && NB % NMMA == 0 && (WIDTH == 64 || WIDTH == 128), "configuration");
  __shared__ __align__(128) unsigned char sA[M * 32];
  __shared__ __align__(128) unsigned char sB[NT * 32];
  __shared__ __align__(8) unsigned long long d_full[NB], d_empty[NB];
  __shared__ unsigned tmem_base;
  __shared__ volatile long long t_issue[2], t_rel[CW];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < M * 32 / 4; e += blockDim.x) reinterpret_cast<unsigned*>(sA)[e] = 0u;
  for (int e = tid; e < NT * 32 / 4; e += blockDim.x) reinterpret_cast<unsigned*>(sB)[e] = 0u;
  if (tid == 0) {
    for (int b = 0; b < NB; ++b) { mbar_init(smem_u32(&d_full[b]), 1); mbar_init(smem_u32(&d_empty[b]), CPT); }
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == CW) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base)), "r"(TCOLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  const long long t0 = clock64();
  long long s_seen = 0, s_ld = 0, s_rel = 0, s_wake = 0, s_issue = 0;

  if (warp >= CW) {
    if (lane == 0) {
      const int w = warp - CW;
      const unsigned idesc = ((unsigned)(NT >> 3) << 17) | ((unsigned)(M >> 4) << 24);       // F16 in, F16 accumulate
      const unsigned long long da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
      for (int t = w; t < ntiles; t += NMMA) {
        const int b = t % NB, ph = (t / NB) & 1;
        long long now = 0, c_a = 0, c_b = 0, c_c = 0;
        if (PROBE == 2) c_a = clock64();
        mbar_wait(smem_u32(&d_empty[b]), (unsigned)(ph ^ 1));
        if (PROBE == 1) {
          now = clock64();
          if (t > 0) { long long m = 0; for (int c = 0; c < CPT; ++c) m = t_rel[c] > m ? t_rel[c] : m; s_wake += now - m; }
          now = clock64();
        }
        if (PROBE == 2) c_b = clock64();
        asm volatile("tcgen05.fence::after_thread_sync;");
        if (PROBE == 1) t_issue[t & 1] = now;
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
            :: "r"(tm + (unsigned)(b * NT)), "l"(da), "l"(db), "r"(idesc), "r"(0u), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
        if (PROBE == 2) c_c = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&d_full[b])) : "memory");
        if (PROBE == 1) s_issue += clock64() - now;
        if (PROBE == 2) { const long long c_d = clock64(); s_wake += c_b - c_a; s_issue += c_c - c_b; s_seen += c_d - c_c; }
      }
    }
  } else {
    const int g = warp / CPT, part = (warp % CPT) >> 2;
    const unsigned lane_base = tm + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)(part * WIDTH);
    unsigned r[32], r2[32];
    unsigned acc = 0;
    for (int t = g; t < ntiles; t += G) {
      const int b = t % NB, ph = (t / NB) & 1;
      mbar_wait_warp<WAIT>(smem_u32(&d_full[b]), (unsigned)ph);
      long long c0 = 0, c1 = 0;
      if (PROBE == 1) { c0 = clock64(); s_seen += c0 - t_issue[t & 1]; }
      asm volatile("tcgen05.fence::after_thread_sync;");
      LD32P(r, lane_base + (unsigned)(b * NT));
      if (WIDTH == 128) LD32P(r2, lane_base + (unsigned)(b * NT + 64));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (PROBE == 1) { c1 = clock64(); s_ld += c1 - c0; }
      asm volatile("tcgen05.fence::before_thread_sync;");
      __syncwarp();
      if (PROBE == 1 && lane == 0) t_rel[warp] = clock64();
      if (lane == 0) mbar_arrive(smem_u32(&d_empty[b]));
      if (PROBE == 1) s_rel += clock64() - c1;
      asm volatile("" : "+r"(r[0]), "+r"(r[8]), "+r"(r[16]), "+r"(r[24]));
      if (REDUCE) {
        unsigned m = or_tree(r);
        if (WIDTH == 128) m |= or_tree(r2);
        if (m & 0x80008000u) acc += m;
      } else {
        acc |= r[0] ^ r[31];
        if (WIDTH == 128) acc |= r2[0];
      }
    }
    if (acc == 0x12345u) sink[blockIdx.x * 1024 + tid] = acc;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid == 0) cycles[blockIdx.x] = clock64() - t0;
  if (PROBE == 1 && blockIdx.x == 0) {
    if (tid == 0) { lat[0] = s_seen; lat[1] = s_ld; lat[2] = s_rel; }
    if (tid == CW * 32) { lat[3] = s_wake; lat[4] = s_issue; }
  }
  if (PROBE == 2 && blockIdx.x == 0 && tid == CW * 32) { lat[0] = s_wake; lat[1] = s_issue; lat[2] = s_seen; }
  if (warp == CW) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tm), "r"(TCOLS));
  }
}

// The consumers issue: 4 accumulators of 128 columns, 16 warps, group g (4 warps, one per lane quadrant, 128
// columns each) serves accumulator g alone.  After the group's loads completed (named barrier of the group's 128
// threads) one lane of its first warp issues the next MMA into the accumulator itself — no release barrier, no
// issuing warp to wake.  ISSUER: which warp of the group issues (0: the first; 1: rotate over the four).
#define LD64P(r, taddr)                                                                                            \
  asm volatile(                                                                                                    \
      "tcgen05.ld.sync.aligned.32x32b.x64.pack::16b.b32 "                                                          \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                    \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "                           \
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "                           \
      "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];\n"                  \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),            \
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),      \
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),    \
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]),    \
        "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]),    \
        "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]),    \
        "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]),    \
        "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])     \
      : "r"(taddr))

template <int REDUCE, int LD64, int ISSUER>
__global__ void __launch_bounds__(512, 1)
self_issue_kernel(long long* cycles, unsigned* sink, int ntiles) {
  constexpr int NB = 4, NT = 128;
  __shared__ __align__(128) unsigned char sA[M * 32];
  __shared__ __align__(128) unsigned char sB[NT * 32];
  __shared__ __align__(8) unsigned long long d_full[NB];
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < M * 32 / 4; e += blockDim.x) reinterpret_cast<unsigned*>(sA)[e] = 0u;
  for (int e = tid; e < NT * 32 / 4; e += blockDim.x) reinterpret_cast<unsigned*>(sB)[e] = 0u;
  if (tid == 0) {
    for (int b = 0; b < NB; ++b) mbar_init(smem_u32(&d_full[b]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  const long long t0 = clock64();
  const int g = warp >> 2, quad = warp & 3;
  const unsigned idesc = ((unsigned)(NT >> 3) << 17) | ((unsigned)(M >> 4) << 24);
  const unsigned long long da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
  const unsigned acc_addr = tm + (unsigned)(g * NT), full = smem_u32(&d_full[g]);
  const unsigned lane_base = tm + ((unsigned)(quad * 32) << 16) + (unsigned)(g * NT);
  auto issue = [&]() {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
        :: "r"(acc_addr), "l"(da), "l"(db), "r"(idesc), "r"(0u), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(full) : "memory");
  };
  if (quad == 0 && lane == 0 && g < ntiles) issue();
  unsigned r[64];
  unsigned acc = 0;
  int k = 0;
  for (int t = g; t < ntiles; t += NB, ++k) {
    mbar_wait_warp<0>(full, (unsigned)(k & 1));
    asm volatile("tcgen05.fence::after_thread_sync;");
    if (LD64) { LD64P(r, lane_base); }
    else { unsigned (&ra)[32] = *reinterpret_cast<unsigned (*)[32]>(&r[0]); unsigned (&rb)[32] = *reinterpret_cast<unsigned (*)[32]>(&r[32]);
           LD32P(ra, lane_base); LD32P(rb, lane_base + 64u); }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    asm volatile("bar.sync %0, 128;" :: "r"(1 + g) : "memory");
    const int iw = ISSUER ? (k & 3) : 0;
    if (quad == iw && t + NB < ntiles) {
      asm volatile("tcgen05.fence::after_thread_sync;");
      if (lane == 0) issue();
      __syncwarp();
    }
    asm volatile("" : "+r"(r[0]), "+r"(r[8]), "+r"(r[16]), "+r"(r[24]));
    if (REDUCE) {
      unsigned m0 = r[0], m1 = r[16], m2 = r[32], m3 = r[48];
#pragma unroll
      for (int i = 1; i < 16; ++i) { m0 |= r[i]; m1 |= r[16 + i]; m2 |= r[32 + i]; m3 |= r[48 + i]; }
      const unsigned m = m0 | m1 | m2 | m3;
      if (m & 0x80008000u) acc += m;
    } else {
      acc |= r[0] ^ r[63];
    }
  }
  if (acc == 0x12345u) sink[blockIdx.x * 1024 + tid] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (tid == 0) cycles[blockIdx.x] = clock64() - t0;
  if (warp == 0) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
  }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } } while (0)

template <int NB, int NT, int CW, int CPT, int NMMA, int REDUCE, int PROBE, int WAIT = 0>
static int run(const char* name, int nsm, long long* dcyc, unsigned* dsink, long long* dlat, int per_sm = 1) {
  const int ntiles = 4096;
  const int nsm_real = nsm;
  nsm *= per_sm;                      // per_sm blocks share an SM (each with its own part of TMEM)
  CK(cudaMemset(dlat, 0, 64));
  rtt_kernel<NB, NT, CW, CPT, NMMA, REDUCE, PROBE, WAIT><<<nsm, 32 * (CW + NMMA)>>>(dcyc, dsink, ntiles, dlat);   // warm-up
  rtt_kernel<NB, NT, CW, CPT, NMMA, REDUCE, PROBE, WAIT><<<nsm, 32 * (CW + NMMA)>>>(dcyc, dsink, ntiles, dlat);
  CK(cudaDeviceSynchronize());
  std::vector<long long> cyc(nsm); long long lat[8];
  CK(cudaMemcpy(cyc.data(), dcyc, nsm * sizeof(long long), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(lat, dlat, 64, cudaMemcpyDeviceToHost));
  double mean = 0; long long mx = 0;
  for (long long c : cyc) { mean += (double)c / nsm; mx = c > mx ? c : mx; }
  (void)nsm_real;
  const double per_tile = mean / ntiles / per_sm, per_128 = per_tile * 128.0 / NT;
  printf("  %-44s %8.1f cycles per tile of %3d columns = %6.1f per 128 x 128 pair tests (max SM %.1f)\n", name, per_tile, NT,
         per_128, (double)mx / ntiles / per_sm * 128.0 / NT);
  if (PROBE == 2)
    printf("      issuing thread 0 (block 0, mean cycles per own tile): in the d_empty wait %.0f, fence + MMA %.0f, commit %.0f\n",
           (double)lat[0] * NMMA / ntiles, (double)lat[1] * NMMA / ntiles, (double)lat[2] * NMMA / ntiles);
  if (PROBE == 1)
    printf("      stages (block 0, mean cycles): wake of the MMA thread after the last release %.0f, issue + commit %.0f, "
           "issue -> consumer sees d_full %.0f, load %.0f, fence + release %.0f\n",
           (double)lat[3] / (ntiles - 1), (double)lat[4] / ntiles, (double)lat[0] / ntiles, (double)lat[1] / ntiles, (double)lat[2] / ntiles);
  return 0;
}

template <int REDUCE, int LD64, int ISSUER>
static int run_self(const char* name, int nsm, long long* dcyc, unsigned* dsink) {
  const int ntiles = 4096;
  self_issue_kernel<REDUCE, LD64, ISSUER><<<nsm, 512>>>(dcyc, dsink, ntiles);
  self_issue_kernel<REDUCE, LD64, ISSUER><<<nsm, 512>>>(dcyc, dsink, ntiles);
  CK(cudaDeviceSynchronize());
  std::vector<long long> cyc(nsm);
  CK(cudaMemcpy(cyc.data(), dcyc, nsm * sizeof(long long), cudaMemcpyDeviceToHost));
  double mean = 0; long long mx = 0;
  for (long long c : cyc) { mean += (double)c / nsm; mx = c > mx ? c : mx; }
  printf("  %-44s %8.1f cycles per 128 x 128 pair tests (max SM %.1f)\n", name, mean / ntiles, (double)mx / ntiles);
  return 0;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int nsm = p.multiProcessorCount;
  long long *dcyc, *dlat; unsigned* dsink;
  CK(cudaMalloc(&dcyc, 2 * nsm * sizeof(long long))); CK(cudaMalloc(&dlat, 64)); CK(cudaMalloc(&dsink, (size_t)2 * nsm * 1024 * 4));
  int rc = 0;
  printf("SMs %d\n\n== E. one accumulator: the round trip itself (no other work on the SM) ==\n", nsm);
  //        NB   NT  CW CPT NMMA RED PROBE
  rc |= run<1, 128,  8,  8,  1,  0,  1>("128 columns, 8 warps x 64 columns", nsm, dcyc, dsink, dlat);
  rc |= run<1, 128,  4,  4,  1,  0,  1>("128 columns, 4 warps x 128 columns", nsm, dcyc, dsink, dlat);
  rc |= run<1, 256, 16, 16,  1,  0,  1>("256 columns, 16 warps x 64 columns", nsm, dcyc, dsink, dlat);
  rc |= run<1,  64,  4,  4,  1,  0,  1>("64 columns, 4 warps x 64 columns", nsm, dcyc, dsink, dlat);
  printf("\n== F. all of TMEM in flight: cycles per 128 x 128 pair tests ==\n");
  rc |= run<4, 128, 16,  8,  4,  1,  0>("kernel: 4 x 128, 16 warps, 8/tile, 4 issuers", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  8,  4,  0,  0>("  same, no reduction", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  8,  1,  1,  0>("  same, one issuing thread", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  4,  4,  1,  0>("4 x 128, 16 warps, 4/tile (128 col each)", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  4,  1,  1,  0>("  same, one issuing thread", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128,  8,  4,  4,  1,  0>("4 x 128, 8 warps, 4/tile", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128,  8,  8,  4,  1,  0>("4 x 128, 8 warps, 8/tile (every tile)", nsm, dcyc, dsink, dlat);
  rc |= run<2, 256, 16, 16,  2,  1,  0>("2 x 256, 16 warps, 16/tile, 2 issuers", nsm, dcyc, dsink, dlat);
  rc |= run<2, 256, 16,  8,  2,  1,  0>("2 x 256, 16 warps, 8/tile (128 col each)", nsm, dcyc, dsink, dlat);
  rc |= run<8,  64, 16,  4,  4,  1,  0>("8 x 64, 16 warps, 4/tile, 4 issuers", nsm, dcyc, dsink, dlat);
  rc |= run<8,  64, 16,  4,  1,  1,  0>("  same, one issuing thread", nsm, dcyc, dsink, dlat);
  rc |= run<8,  64, 16,  4,  8,  1,  0>("  same, 8 issuers", nsm, dcyc, dsink, dlat);
  rc |= run<8,  64,  8,  4,  4,  1,  0>("8 x 64, 8 warps, 4/tile, 4 issuers", nsm, dcyc, dsink, dlat);
  printf("\n== H. where the issuing thread spends its time ==\n");
  rc |= run<4, 128, 16,  4,  4,  1,  2>("4 x 128, 16 warps, 4/tile, 4 issuers", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  4,  1,  1,  2>("4 x 128, 16 warps, 4/tile, 1 issuer", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  4,  2,  1,  2>("4 x 128, 16 warps, 4/tile, 2 issuers", nsm, dcyc, dsink, dlat);
  rc |= run<4, 128, 16,  8,  4,  1,  2>("4 x 128, 16 warps, 8/tile, 4 issuers", nsm, dcyc, dsink, dlat);
  printf("\n== J. two blocks per SM, each with half of TMEM (cycles per tile per SM) ==\n");
  rc |= run<2, 128,  8,  4,  2,  1,  0>("2 blocks x (2 x 128, 8 warps, 4/tile, 2 issuers)", nsm, dcyc, dsink, dlat, 2);
  rc |= run<2, 128,  8,  4,  2,  0,  0>("  same, no reduction", nsm, dcyc, dsink, dlat, 2);
  rc |= run<2, 128,  8,  4,  2,  1,  0>("1 block  x (2 x 128, 8 warps, 4/tile, 2 issuers)", nsm, dcyc, dsink, dlat, 1);
  rc |= run<4,  64,  8,  4,  4,  1,  0>("2 blocks x (4 x 64, 8 warps, 4/tile, 4 issuers)", nsm, dcyc, dsink, dlat, 2);
  rc |= run<2, 128, 16,  8,  2,  1,  0>("2 blocks x (2 x 128, 16 warps, 8/tile, 2 issuers)", nsm, dcyc, dsink, dlat, 2);
  printf("\n== I. the consumers issue the next MMA themselves (4 x 128, 16 warps, 4/tile) ==\n");
  rc |= run_self<1, 0, 0>("two x32 loads, first warp issues", nsm, dcyc, dsink);
  rc |= run_self<1, 1, 0>("one x64 load, first warp issues", nsm, dcyc, dsink);
  rc |= run_self<1, 0, 1>("two x32 loads, issuer rotates", nsm, dcyc, dsink);
  rc |= run_self<1, 1, 1>("one x64 load, issuer rotates", nsm, dcyc, dsink);
  rc |= run_self<0, 1, 0>("one x64 load, first warp issues, no reduction", nsm, dcyc, dsink);
  printf("tc_rtt exit %d\n", rc);
  return rc;
}
