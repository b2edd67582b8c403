// Microbenchmarks behind the round-2 all-pairs kernel (csrc/psim_allpairs_tc.cuh).  One binary, four questions:
//
//  A. numerics + TMEM layout of tcgen05.mma kind::f16 with FP16 operands, K = 16 in ONE instruction, with the
//     accumulator kept as FP32 and as FP16: is the sum of the 16 exact products formed at (about) FP32 width and
//     rounded once, or at FP16 width?  The distance filter needs the former: its terms are ~|x|^2 and cancel
//     down to ~r_cut^2.  Also: where do FP16 accumulator entries live in TMEM, and what does
//     tcgen05.ld ... .pack::16b return?
//  B. TMEM -> register throughput of tcgen05.ld 32x32b.x32 with and without .pack::16b (values per clock per SM).
//  C. issue rate of the reduction candidates: FMNMX3 (fp32 three-input min), VHMNMX (half2 three-input min),
//     LOP3 (three-input OR), HMNMX2, per SM sub-partition.
//  D. dense tcgen05.mma rate of kind::tf32 (K = 8) and kind::f16 (K = 16), M128 N256, one issuing thread per SM
//     on all SMs: the measured tensor peak that bench.py's roofline uses for this kernel family.
//
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 tc_f16.cu -o _build/tc_f16
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

static constexpr int M = 128, N = 128, K = 16;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// rows x 16 halves (32 B per row): 8-row groups 256 B apart (SBO), the two 16-byte K-chunks of a row 128 B apart
// (LBO), rows of a group 16 B apart — the canonical K-major SWIZZLE_NONE layout, byte-identical to the TF32 K = 8 case.
__host__ __device__ __forceinline__ unsigned tile_off16(int r, int k) {
  return (unsigned)((r >> 3) * 256 + (k >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2);
}
__device__ __forceinline__ unsigned long long make_desc(unsigned addr) {
  const unsigned long long lbo = 128 >> 4, sbo = 256 >> 4;
  return (unsigned long long)((addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
}
__device__ __forceinline__ void mbar_wait(unsigned addr, unsigned parity) {
  unsigned done = 0;
  while (!done)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
}

#define LD32(r, taddr, PACK)                                                                                       \
  asm volatile(                                                                                                    \
      "tcgen05.ld.sync.aligned.32x32b.x32" PACK ".b32 "                                                            \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                    \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"                  \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),            \
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),      \
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),    \
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])     \
      : "r"(taddr))

// ---------------------------------------------------------------- A. numerics and layout
// acc_f16 = 0: D is FP32 (c_format 1); 1: D is FP16 (c_format 0).
// out_raw[row][0..127]   : 128 TMEM columns of the row read WITHOUT pack (raw 32-bit words)
// out_pack[row][0..63]   : the same 128 columns read WITH .pack::16b (two x32 loads of 64 columns each)
__global__ void __launch_bounds__(128) gram16_kernel(const __half* A, const __half* B, unsigned* out_raw, unsigned* out_pack, int acc_f16) {
  __shared__ __align__(128) unsigned char sA[M * K * 2];
  __shared__ __align__(128) unsigned char sB[N * K * 2];
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < M * K; e += 128) *(__half*)(sA + tile_off16(e / K, e % K)) = A[e];
  for (int e = tid; e < N * K; e += 128) *(__half*)(sB + tile_off16(e / K, e % K)) = B[e];
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  if (tid == 0) {
    // c_format at bit 4 (0 = F16, 1 = F32), a/b format F16 = 0, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    const unsigned idesc = ((acc_f16 ? 0u : 1u) << 4) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
    const unsigned long long da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
        :: "r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(0u), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar)) : "memory");
  }
  mbar_wait(smem_u32(&mbar), 0u);
  asm volatile("tcgen05.fence::after_thread_sync;");
  const int row = warp * 32 + lane;
  const unsigned lane_base = tm + ((unsigned)(warp * 32) << 16);
  for (int c = 0; c < N; c += 32) {
    unsigned r[32];
    LD32(r, lane_base + c, "");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int k = 0; k < 32; ++k) out_raw[row * N + c + k] = r[k];
  }
  for (int c = 0; c < N; c += 64) {
    unsigned r[32];
    LD32(r, lane_base + c, ".pack::16b");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int k = 0; k < 32; ++k) out_pack[row * (N / 2) + c / 2 + k] = r[k];
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tm));
}

// ---------------------------------------------------------------- B. TMEM read throughput
template <int PACK>
__global__ void __launch_bounds__(512) ld_rate_kernel(long long* cycles, unsigned* sink, int iters) {
  __shared__ unsigned tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned base = tmem_base + ((unsigned)((warp & 3) * 32) << 16);
  unsigned acc = 0;
  const int span = PACK ? 64 : 32;                    // TMEM columns one x32 load covers
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {                     // 8 loads per iteration, two in flight
      unsigned r[32], q[32];
      const unsigned c0 = (unsigned)(((2 * c) * span) & 511), c1 = (unsigned)(((2 * c + 1) * span) & 511);
      if (PACK) { LD32(r, base + c0, ".pack::16b"); LD32(q, base + c1, ".pack::16b"); }
      else      { LD32(r, base + c0, "");           LD32(q, base + c1, ""); }
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      acc ^= r[0] ^ r[31] ^ q[0] ^ q[31];
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem_base));
}

// ---------------------------------------------------------------- C. reduction instruction rates
// OP 0: FMNMX3, 1: VHMNMX (half2 three-input min), 2: LOP3 (a | b | c), 3: HMNMX2 (two-input), 4: HSETP2 chain,
// 5: HFMA2 (the FMA pipe, for reference), 6: four HSETP2 beside four LOP3 (do the two pipes overlap?)
template <int OP>
__global__ void __launch_bounds__(512) alu_rate_kernel(long long* cycles, unsigned* sink, const unsigned* seed, int iters) {
  unsigned v[8], s1 = seed[threadIdx.x & 31], s2 = seed[32 + (threadIdx.x & 31)];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = seed[(threadIdx.x + k) & 63];
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      unsigned& nb = v[(k + 1) & 7];                   // a neighbouring chain as the second operand: three live inputs
      if (OP == 0) v[k] = __float_as_uint(fminf(__uint_as_float(v[k]), fminf(__uint_as_float(nb), __uint_as_float(s1))));
      if (OP == 1) { __half2 a = *reinterpret_cast<__half2*>(&v[k]), b = *reinterpret_cast<__half2*>(&nb), c = *reinterpret_cast<__half2*>(&s1);
                     a = __hmin2(a, __hmin2(b, c)); v[k] = *reinterpret_cast<unsigned*>(&a); }
      if (OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0xFE;" : "+r"(v[k]) : "r"(nb), "r"(s1));
      if (OP == 3) { __half2 a = *reinterpret_cast<__half2*>(&v[k]), b = *reinterpret_cast<__half2*>(&nb);
                     a = __hmin2(a, b); v[k] = *reinterpret_cast<unsigned*>(&a); }
      if (OP == 5) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(v[k]) : "r"(nb), "r"(s1));
    }
    if (OP == 4 || OP == 6) {
      // 8 (OP 4) or 4 (OP 6, beside 4 LOP3) packed-half compares "any entry negative", chained through a predicate
      unsigned any;
      if (OP == 4)
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %9, 0;\n\t"
                     "setp.lt.or.f16x2 p|q, %1, %10, p;\n\tsetp.lt.or.f16x2 p|q, %2, %10, p;\n\t"
                     "setp.lt.or.f16x2 p|q, %3, %10, p;\n\tsetp.lt.or.f16x2 p|q, %4, %10, p;\n\t"
                     "setp.lt.or.f16x2 p|q, %5, %10, p;\n\tsetp.lt.or.f16x2 p|q, %6, %10, p;\n\t"
                     "setp.lt.or.f16x2 p|q, %7, %10, p;\n\tsetp.lt.or.f16x2 p|q, %8, %10, p;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(any) : "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(s2 & 1u), "r"(0u));
      else
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 p, %9, 0;\n\t"
                     "setp.lt.or.f16x2 p|q, %1, %10, p;\n\tlop3.b32 %5, %5, %6, %11, 0xFE;\n\t"
                     "setp.lt.or.f16x2 p|q, %2, %10, p;\n\tlop3.b32 %6, %6, %7, %11, 0xFE;\n\t"
                     "setp.lt.or.f16x2 p|q, %3, %10, p;\n\tlop3.b32 %7, %7, %8, %11, 0xFE;\n\t"
                     "setp.lt.or.f16x2 p|q, %4, %10, p;\n\tlop3.b32 %8, %8, %5, %11, 0xFE;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(any), "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7])
                     : "r"(s2 & 1u), "r"(0u), "r"(s1));
      v[0] ^= any;
    }
    s1 += s2;                                         // keeps the loop from collapsing
  }
  const long long t1 = clock64();
  unsigned x = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) x ^= v[k];
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = x;
}

// ---------------------------------------------------------------- D. dense MMA rate
// KIND 0: kind::tf32 (K = 8 per instruction, 32 B rows), 1: kind::f16 (K = 16 per instruction, 32 B rows).
// M128 N256; one thread per SM issues `iters` MMAs into alternating accumulators and commits once at the end.
template <int KIND>
__global__ void __launch_bounds__(32) mma_rate_kernel(long long* cycles, int iters) {
  constexpr int NN = 256;
  __shared__ __align__(128) unsigned char sA[M * 32];
  __shared__ __align__(128) unsigned char sB[NN * 32];
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x;
  for (int e = tid; e < M * 32 / 4; e += 32) reinterpret_cast<unsigned*>(sA)[e] = 0u;
  for (int e = tid; e < NN * 32 / 4; e += 32) reinterpret_cast<unsigned*>(sB)[e] = 0u;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  if (tid == 0) {
    const unsigned fmt = KIND == 0 ? ((2u << 7) | (2u << 10)) : 0u;
    const unsigned idesc = (1u << 4) | fmt | ((unsigned)(NN >> 3) << 17) | ((unsigned)(M >> 4) << 24);
    const unsigned long long da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const unsigned acc = tm + (unsigned)((it & 1) * NN);
      if (KIND == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
                     :: "r"(acc), "l"(da), "l"(db), "r"(idesc), "r"(1u), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
      else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
                     :: "r"(acc), "l"(da), "l"(db), "r"(idesc), "r"(1u), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar)) : "memory");
    mbar_wait(smem_u32(&mbar), 0u);
    cycles[blockIdx.x] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
}

// ---------------------------------------------------------------- host
static float h2f(__half h) { return __half2float(h); }
static void split2h(double x, __half& a, __half& b) { a = __float2half_rn((float)x); b = __float2half_rn((float)(x - (double)h2f(a))); }
static void split3h(double x, __half& a, __half& b, __half& c) {
  a = __float2half_rn((float)x); const double r = x - (double)h2f(a);
  b = __float2half_rn((float)r); c = __float2half_rn((float)(r - (double)h2f(b)));
}
static double urand() { return rand() / (RAND_MAX + 1.0); }

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } } while (0)

int main() {
  int nsm = 0;
  CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  printf("SMs %d, nominal clock %d MHz\n", nsm, khz / 1000);

  // ---- A: the filter's operands for a 198 x 198 x 99 box, centred, scaled by 2^-3 ----
  const double S = 0.125, half_box[3] = {99.0, 99.0, 49.5}, rcut = 2.5;
  const double thr = rcut * rcut * S * S;                      // scaled threshold: 0.09765625
  std::vector<double> xi(M * 3), xj(N * 3);
  srand(11);
  for (int i = 0; i < M; ++i) for (int d = 0; d < 3; ++d) xi[3 * i + d] = (2 * urand() - 1) * half_box[d];
  for (int j = 0; j < N; ++j) {
    if (j < M && (j % 2 == 0)) {                               // partner of row j at a distance close to the cut-off
      double u[3], nrm = 0; for (int d = 0; d < 3; ++d) { u[d] = 2 * urand() - 1; nrm += u[d] * u[d]; }
      const double r = 2.3 + 0.4 * urand();
      for (int d = 0; d < 3; ++d) xj[3 * j + d] = xi[3 * j + d] + u[d] / sqrt(nrm) * r;
    } else {
      for (int d = 0; d < 3; ++d) xj[3 * j + d] = (2 * urand() - 1) * half_box[d];
    }
  }
  std::vector<__half> A(M * K), B(N * K);
  const __half one = __float2half_rn(1.0f), zero = __float2half_rn(0.0f);
  for (int i = 0; i < M; ++i) {
    __half p[3][2]; double n2 = 0;
    for (int d = 0; d < 3; ++d) { const double v = xi[3 * i + d] * S; split2h(v, p[d][0], p[d][1]); n2 += v * v; }
    __half c1, c2, c3; split3h(-(thr - n2), c1, c2, c3);       // acc = w_j - 2 x_i.x_j - (thr - |x_i|^2) = r^2 - thr
    __half* a = &A[i * K];
    for (int d = 0; d < 3; ++d) { a[3 * d] = p[d][0]; a[3 * d + 1] = p[d][0]; a[3 * d + 2] = p[d][1]; }
    a[9] = a[10] = a[11] = one; a[12] = c1; a[13] = c2; a[14] = c3; a[15] = zero;
  }
  for (int j = 0; j < N; ++j) {
    __half p[3][2]; double n2 = 0;
    for (int d = 0; d < 3; ++d) { const double v = xj[3 * j + d] * S; split2h(-2.0 * v, p[d][0], p[d][1]); n2 += v * v; }
    __half w1, w2, w3; split3h(n2, w1, w2, w3);
    __half* b = &B[j * K];
    for (int d = 0; d < 3; ++d) { b[3 * d] = p[d][0]; b[3 * d + 1] = p[d][1]; b[3 * d + 2] = p[d][0]; }
    b[9] = w1; b[10] = w2; b[11] = w3; b[12] = b[13] = b[14] = one; b[15] = zero;
  }
  std::vector<double> ref(M * N), truth(M * N);
  for (int i = 0; i < M; ++i) for (int j = 0; j < N; ++j) {
    double s = 0; for (int k = 0; k < K; ++k) s += (double)h2f(A[i * K + k]) * (double)h2f(B[j * K + k]);
    ref[i * N + j] = s;                                        // what an exact tensor core would return
    double r2 = 0; for (int d = 0; d < 3; ++d) { const double t = (xi[3 * i + d] - xj[3 * j + d]) * S; r2 += t * t; }
    truth[i * N + j] = r2 - thr;                               // what the filter wants to know the sign of
  }
  __half *dA, *dB; unsigned *dRaw, *dPack;
  CK(cudaMalloc(&dA, A.size() * 2)); CK(cudaMalloc(&dB, B.size() * 2));
  CK(cudaMalloc(&dRaw, M * N * 4)); CK(cudaMalloc(&dPack, M * N * 2));
  CK(cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice));
  std::vector<unsigned> raw(M * N), pack(M * N / 2);
  for (int acc_f16 = 0; acc_f16 < 2; ++acc_f16) {
    CK(cudaMemset(dRaw, 0xCD, M * N * 4)); CK(cudaMemset(dPack, 0xCD, M * N * 2));
    gram16_kernel<<<1, 128>>>(dA, dB, dRaw, dPack, acc_f16);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(raw.data(), dRaw, M * N * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(pack.data(), dPack, M * N * 2, cudaMemcpyDeviceToHost));
    printf("\n== A. kind::f16, accumulator %s ==\n", acc_f16 ? "FP16" : "FP32");
    for (int row : {0, 37}) {
      printf("row %d raw[0..7]  :", row); for (int k = 0; k < 8; ++k) printf(" %08x", raw[row * N + k]); printf("\n");
      printf("row %d pack[0..7] :", row); for (int k = 0; k < 8; ++k) printf(" %08x", pack[row * (N / 2) + k]); printf("\n");
      printf("row %d expect f32 :", row); for (int k = 0; k < 8; ++k) { float f = (float)ref[row * N + k]; unsigned u; memcpy(&u, &f, 4); printf(" %08x", u); } printf("\n");
      printf("row %d expect f16 :", row); for (int k = 0; k < 8; ++k) { __half h = __float2half_rn((float)ref[row * N + k]); unsigned short u; memcpy(&u, &h, 2); printf("     %04x", u); } printf("\n");
    }
    auto as_f32 = [&](int i, int j) { float f; memcpy(&f, &raw[i * N + j], 4); return (double)f; };
    auto lo16 = [&](unsigned w) { __half h; unsigned short u = (unsigned short)(w & 0xFFFF); memcpy(&h, &u, 2); return (double)h2f(h); };
    auto hi16 = [&](unsigned w) { __half h; unsigned short u = (unsigned short)(w >> 16); memcpy(&h, &u, 2); return (double)h2f(h); };
    auto report = [&](const char* name, auto get) {
      double maxerr = 0, maxerr_near = 0; int sign_bad = 0, sign_bad_margin = 0, near = 0, nan = 0;
      for (int i = 0; i < M; ++i) for (int j = 0; j < N; ++j) {
        const double v = get(i, j), r = ref[i * N + j], t = truth[i * N + j];
        if (!(v == v)) { ++nan; continue; }
        double e = fabs(v - r);
        if (std::isinf(v)) e = (r > 6.0e4) ? 0 : 1e30;                 // FP16 overflow of a far pair is fine
        if (e > maxerr) maxerr = e;
        if (fabs(r) < 4 * thr) { ++near; if (e > maxerr_near) maxerr_near = e; }
        if ((v < 0) != (t < 0)) { ++sign_bad; if (fabs(t) > 0.02 * thr) ++sign_bad_margin; }
      }
      printf("  [%s] max|D-exact| = %.3g (%.3g thr), among %d entries within 4 thr of 0: %.3g (%.3g thr); wrong sign vs true r^2-thr: %d, of which |true| > 2%% thr: %d; NaN %d\n",
             name, maxerr, maxerr / thr, near, maxerr_near, maxerr_near / thr, sign_bad, sign_bad_margin, nan);
    };
    if (!acc_f16) {
      report("raw word = f32 D[row][col]", [&](int i, int j) { return as_f32(i, j); });
    } else {
      report("raw word low half = D[row][col]", [&](int i, int j) { return lo16(raw[i * N + j]); });
      report("raw word = (D[row][2c], D[row][2c+1]) for c < 64", [&](int i, int j) { const unsigned w = raw[i * N + j / 2]; return (j & 1) ? hi16(w) : lo16(w); });
      report("pack word = (D[row][2c] lo, D[row][2c+1] hi)", [&](int i, int j) { const unsigned w = pack[i * (N / 2) + j / 2]; return (j & 1) ? hi16(w) : lo16(w); });
    }
  }

  long long* dcyc; unsigned* dsink;
  CK(cudaMalloc(&dcyc, sizeof(long long) * nsm)); CK(cudaMalloc(&dsink, sizeof(unsigned) * nsm * 512));
  cudaEvent_t ea, eb; cudaEventCreate(&ea); cudaEventCreate(&eb);
  auto timed = [&](auto launch, float* ms, long long* cyc) -> cudaError_t {
    launch(); cudaDeviceSynchronize();
    cudaEventRecord(ea); launch(); cudaEventRecord(eb);
    cudaError_t e = cudaEventSynchronize(eb);
    cudaEventElapsedTime(ms, ea, eb);
    cudaMemcpy(cyc, dcyc, sizeof(long long), cudaMemcpyDeviceToHost);
    return e;
  };

  printf("\n== B. tcgen05.ld 32x32b.x32 throughput (two loads in flight per warp) ==\n");
  for (int warps : {4, 8, 16}) {
    for (int packed = 0; packed < 2; ++packed) {
      const int iters = 2000; float ms = 0; long long cyc = 0;
      cudaError_t e = packed ? timed([&] { ld_rate_kernel<1><<<nsm, 32 * warps>>>(dcyc, dsink, iters); }, &ms, &cyc)
                             : timed([&] { ld_rate_kernel<0><<<nsm, 32 * warps>>>(dcyc, dsink, iters); }, &ms, &cyc);
      const double regs = (double)iters * 8 * 32 * 32 * warps;             // 32-bit registers written per SM
      const double vals = regs * (packed ? 2 : 1);
      printf("  %2d warps/SM, %s: %s, %.3f ms, %lld cycles: %.1f registers/clk/SM, %.1f accumulator entries/clk/SM\n", warps,
             packed ? ".pack::16b" : "unpacked  ", cudaGetErrorString(e), ms, cyc, regs / cyc, vals / cyc);
    }
  }

  printf("\n== C. reduction instruction rates, 16 warps/SM, 8 chains per thread ==\n");
  {
    std::vector<unsigned> seed(64); for (auto& s : seed) s = 0x3c003c00u + (rand() & 0x03ff03ff);
    unsigned* dseed; CK(cudaMalloc(&dseed, 256)); CK(cudaMemcpy(dseed, seed.data(), 256, cudaMemcpyHostToDevice));
    const char* names[7] = {"FMNMX3 (min.f32 x3)", "VHMNMX (min.f16x2 x3)", "LOP3 (a|b|c)", "HMNMX2 (min.f16x2 x2)",
                            "HSETP2 (lt.or.f16x2) x8", "HFMA2 (fma.f16x2)", "4 HSETP2 + 4 LOP3"};
    for (int op = 0; op < 7; ++op) {
      const int iters = 4000; float ms = 0; long long cyc = 0; cudaError_t e = cudaSuccess;
      if (op == 0) e = timed([&] { alu_rate_kernel<0><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      if (op == 1) e = timed([&] { alu_rate_kernel<1><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      if (op == 2) e = timed([&] { alu_rate_kernel<2><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      if (op == 3) e = timed([&] { alu_rate_kernel<3><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      if (op == 4) e = timed([&] { alu_rate_kernel<4><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      if (op == 5) e = timed([&] { alu_rate_kernel<5><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      if (op == 6) e = timed([&] { alu_rate_kernel<6><<<nsm, 512>>>(dcyc, dsink, dseed, iters); }, &ms, &cyc);
      const double winstr = (double)iters * 8 * 16;                        // warp instructions per SM (+ ~3 of loop overhead per 8)
      printf("  %-22s %s, %lld cycles: %.2f warp-instr/clk/SM = %.2f per SMSP (reciprocal %.2f clk)\n", names[op], cudaGetErrorString(e), cyc,
             winstr / cyc, winstr / cyc / 4, 4 * cyc / winstr);
    }
  }

  printf("\n== D. dense tcgen05.mma rate, M128 N256, one issuing thread per SM, all SMs ==\n");
  for (int kind = 0; kind < 2; ++kind) {
    const int iters = 20000; float ms = 0; long long cyc = 0;
    cudaError_t e = kind ? timed([&] { mma_rate_kernel<1><<<nsm, 32>>>(dcyc, iters); }, &ms, &cyc)
                         : timed([&] { mma_rate_kernel<0><<<nsm, 32>>>(dcyc, iters); }, &ms, &cyc);
    const double kk = kind ? 16 : 8;
    const double flops = 2.0 * M * 256 * kk * iters * nsm;
    printf("  kind::%s K=%d: %s, %.3f ms, %lld cycles/SM (%.1f per MMA): %.1f TFLOP/s dense, clock %.0f MHz\n", kind ? "f16 " : "tf32", (int)kk,
           cudaGetErrorString(e), ms, cyc, (double)cyc / iters, flops / (ms * 1e-3) / 1e12, cyc / (ms * 1e-3) / 1e6);
  }
  return 0;
}
