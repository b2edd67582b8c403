// De-risking test for the tensor-core distance filter: D[128 x 256] = A[128 x 16] * B[256 x 16]^T in
// TF32 with tcgen05.mma (operands in shared memory, canonical K-major no-swizzle layout, descriptors
// built by hand), accumulator in TMEM, read back with tcgen05.ld.  Compared with a CPU reference.
// nvcc -gencode arch=compute_100a,code=sm_100a tc_gram.cu -o tc_gram
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

static constexpr int M = 128, N = 256, K = 16;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// rows x 16 tf32, two K-blocks of 8; inside a K-block: 8-row groups 256 B apart, 16-byte K-chunks 128 B apart
__device__ __forceinline__ unsigned tile_off(int rows, int r, int k) {
  return (unsigned)((k >> 3) * rows * 32 + (r >> 3) * 256 + ((k >> 2) & 1) * 128 + (r & 7) * 16 + (k & 3) * 4);
}
__device__ __forceinline__ unsigned long long make_desc(unsigned addr) {
  const unsigned long long lbo = 128 >> 4, sbo = 256 >> 4;
  return (unsigned long long)((addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
}

__global__ void __launch_bounds__(128) gram_kernel(const float* A, const float* B, float* D) {
  __shared__ __align__(128) unsigned char sA[M * K * 4];
  __shared__ __align__(128) unsigned char sB[N * K * 4];
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  for (int e = tid; e < M * K; e += 128) *(float*)(sA + tile_off(M, e / K, e % K)) = A[e];
  for (int e = tid; e < N * K; e += 128) *(float*)(sB + tile_off(N, e / K, e % K)) = B[e];
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> async proxy (tensor core)
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;

  if (tid == 0) {
    // c_format F32 (1<<4), a/b TF32 (2<<7, 2<<10), K-major both, N>>3 at bit 17, M>>4 at bit 24
    const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
    for (int kb = 0; kb < 2; ++kb) {
      const unsigned long long da = make_desc(smem_u32(sA) + kb * M * 32), db = make_desc(smem_u32(sB) + kb * N * 32);
      const unsigned acc = kb;
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
          :: "r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar)) : "memory");
  }
  // everybody waits for the MMAs
  {
    const unsigned addr = smem_u32(&mbar);
    unsigned done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(addr), "r"(0u) : "memory");
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;");
  const int row = warp * 32 + lane;
  for (int c = 0; c < N; c += 32) {
    unsigned r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(tm + ((unsigned)(warp * 32) << 16) + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int k = 0; k < 32; ++k) D[row * N + c + k] = __uint_as_float(r[k]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tm));
}

// Issue throughput of the control path: ONE thread issues, per iteration, 2 MMAs (alternating accumulator),
// `ncommit` commits (each to its own never-waited mbarrier) and `nwait` try_waits on an already-completed
// barrier.  Nothing waits for the MMAs until the end.  Cycles per iteration.
template <int NN>
__global__ void __launch_bounds__(32) issue_kernel(const float* A, const float* B, long long* out, int iters, int ncommit, int nwait) {
  __shared__ __align__(128) unsigned char sA[M * K * 4];
  __shared__ __align__(128) unsigned char sB[N * K * 4];
  __shared__ __align__(8) unsigned long long mbar[4];
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x;
  for (int e = tid; e < M * K; e += 32) *(float*)(sA + tile_off(M, e / K, e % K)) = A[e];
  for (int e = tid; e < NN * K; e += 32) *(float*)(sB + tile_off(NN, e / K, e % K)) = B[e];
  if (tid == 0) {
    for (int b = 0; b < 4; ++b) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar[b])));
    asm volatile("fence.mbarrier_init.release.cluster;");
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(&mbar[3])) : "memory");   // phase 0 of mbar[3] complete
  }
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base)));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(NN >> 3) << 17) | ((unsigned)(M >> 4) << 24);
  long long t0 = 0, t1 = 0, t2 = 0;
  if (tid == 0) {
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int kb = 0; kb < 2; ++kb) {
        const unsigned long long da = make_desc(smem_u32(sA) + kb * M * 32), db = make_desc(smem_u32(sB) + kb * NN * 32);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
                     :: "r"(tm + (unsigned)((it % (512 / NN)) * NN)), "l"(da), "l"(db), "r"(idesc), "r"((unsigned)kb), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
      }
      for (int cmt = 0; cmt < ncommit; ++cmt)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar[1 + cmt])) : "memory");
      for (int w = 0; w < nwait; ++w) {
        unsigned done = 0;
        while (!done)
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                       : "=r"(done) : "r"(smem_u32(&mbar[3])), "r"(0u) : "memory");
      }
    }
    t1 = clock64();
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar[0])) : "memory");
    unsigned done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(smem_u32(&mbar[0])), "r"(0u) : "memory");
    t2 = clock64();
    out[0] = (t1 - t0) / iters; out[1] = (t2 - t0) / iters;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
}

template <int NN>
static void run_issue(const float* dA, const float* dB, long long* dout) {
  for (int nc = 0; nc <= 2; ++nc)
    for (int nw = 0; nw <= 2; nw += 2) {
      issue_kernel<NN><<<1, 32>>>(dA, dB, dout, 4000, nc, nw);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2] = {0, 0}; cudaMemcpy(h, dout, 16, cudaMemcpyDeviceToHost);
      printf("issue throughput N=%d: 2 MMAs + %d commits + %d try_waits: %lld cycles/iter issued, %lld cycles/iter completed (%s)\n",
             NN, nc, nw, h[0], h[1], cudaGetErrorString(e));
    }
}

// Latency of the accumulator chain: thread 0 issues the two MMAs and commits; everybody waits on the
// mbarrier; optionally every thread then reads its 256 columns back.  Cycles per iteration.
__global__ void __launch_bounds__(128) chain_kernel(const float* A, const float* B, long long* out, int iters, int do_ld) {
  __shared__ __align__(128) unsigned char sA[M * K * 4];
  __shared__ __align__(128) unsigned char sB[N * K * 4];
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int e = tid; e < M * K; e += 128) *(float*)(sA + tile_off(M, e / K, e % K)) = A[e];
  for (int e = tid; e < N * K; e += 128) *(float*)(sB + tile_off(N, e / K, e % K)) = B[e];
  if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
  unsigned sink = 0;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (tid == 0) {
      for (int kb = 0; kb < 2; ++kb) {
        const unsigned long long da = make_desc(smem_u32(sA) + kb * M * 32), db = make_desc(smem_u32(sB) + kb * N * 32);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
                     :: "r"(tm), "l"(da), "l"(db), "r"(idesc), "r"((unsigned)kb), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&mbar)) : "memory");
    }
    unsigned done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(done) : "r"(smem_u32(&mbar)), "r"((unsigned)(it & 1)) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;");
    if (do_ld) {
      unsigned r[32];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
            "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
            "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(tm + ((unsigned)(warp * 32) << 16)));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      sink += r[it & 31];
      asm volatile("tcgen05.fence::before_thread_sync;");
      __syncthreads();
    }
  }
  const long long t1 = clock64();
  if (tid == 0) out[0] = (t1 - t0) / iters;
  if (sink == 0x12345678u) out[1] = sink;
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tm));
}

static float tf32_trunc(float x) { unsigned u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }

int main() {
  std::vector<float> A(M * K), B(N * K), D(M * N), ref(M * N);
  srand(7);
  for (auto& v : A) v = tf32_trunc((rand() / (float)RAND_MAX - 0.5f) * 200.0f);
  for (auto& v : B) v = tf32_trunc((rand() / (float)RAND_MAX - 0.5f) * 200.0f);
  for (int i = 0; i < M; ++i) for (int j = 0; j < N; ++j) {
    double s = 0; for (int k = 0; k < K; ++k) s += (double)A[i * K + k] * B[j * K + k];
    ref[i * N + j] = (float)s;
  }
  float *dA, *dB, *dD;
  cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0, D.size() * 4);
  gram_kernel<<<1, 128>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0, maxref = 0; int bad = 0;
  for (int i = 0; i < M * N; ++i) { double d = fabs((double)D[i] - ref[i]); if (d > maxerr) maxerr = d; if (fabs(ref[i]) > maxref) maxref = fabs(ref[i]); if (d > 1e-2) ++bad; }
  printf("%s: max |D - ref| = %.4g (max |ref| = %.4g), %d of %d entries off by > 1e-2; D[0]=%.3f ref %.3f D[last]=%.3f ref %.3f\n",
         cudaGetErrorString(e), maxerr, maxref, bad, M * N, D[0], ref[0], D[M * N - 1], ref[M * N - 1]);
  long long* dout; cudaMalloc(&dout, 16);
  for (int do_ld = 0; do_ld < 2; ++do_ld) {
    chain_kernel<<<1, 128>>>(dA, dB, dout, 2000, do_ld);
    cudaError_t e2 = cudaDeviceSynchronize();
    long long h = 0; cudaMemcpy(&h, dout, 8, cudaMemcpyDeviceToHost);
    printf("chain latency (2 MMAs M128 N256 K8 + commit + mbarrier wait%s): %lld cycles per iteration (%s)\n",
           do_ld ? " + tcgen05.ld x32 + wait + barrier" : "", h, cudaGetErrorString(e2));
  }
  run_issue<256>(dA, dB, dout);
  run_issue<128>(dA, dB, dout);
  return bad != 0;
}
