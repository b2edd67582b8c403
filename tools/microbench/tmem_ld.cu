// Microbenchmark: TMEM -> register read throughput of tcgen05.ld (32x32b.x32) on sm_100a.
// Decides whether a tensor-core formulation of the distance filter (Gram matrix in TMEM,
// min-reduction epilogue) can beat the FFMA2 filter.  nvcc -arch=sm_100a tmem_ld.cu -o tmem_ld
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(128) tmem_ld_kernel(long long* cycles, float* sink, int iters, int do_min) {
  __shared__ unsigned tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n"
                 :: "l"((unsigned long long)__cvta_generic_to_shared(&tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
  }
  asm volatile("tcgen05.fence::before_thread_sync;\n");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n");
  const unsigned base = tmem_base + ((unsigned)(warp * 32) << 16);
  float m = 3.0e38f;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < 16; c += 2) {
      unsigned r[32], q[32];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
          : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7]),
            "=r"(q[8]), "=r"(q[9]), "=r"(q[10]), "=r"(q[11]), "=r"(q[12]), "=r"(q[13]), "=r"(q[14]), "=r"(q[15]),
            "=r"(q[16]), "=r"(q[17]), "=r"(q[18]), "=r"(q[19]), "=r"(q[20]), "=r"(q[21]), "=r"(q[22]), "=r"(q[23]),
            "=r"(q[24]), "=r"(q[25]), "=r"(q[26]), "=r"(q[27]), "=r"(q[28]), "=r"(q[29]), "=r"(q[30]), "=r"(q[31])
          : "r"(base + (c + 1) * 32));
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
            "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
            "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(base + c * 32));
      asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
      if (do_min) {
#pragma unroll
        for (int k = 0; k < 32; k += 2) m = fminf(m, fminf(__uint_as_float(r[k]), __uint_as_float(r[k + 1])));
#pragma unroll
        for (int k = 0; k < 32; k += 2) m = fminf(m, fminf(__uint_as_float(q[k]), __uint_as_float(q[k + 1])));
      } else {
        m = fminf(m, fminf(__uint_as_float(r[0]), __uint_as_float(q[31])));
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * 128 + threadIdx.x] = m;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" :: "r"(tmem_base));
}

int main() {
  int nsm = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  long long* cyc; float* sink;
  cudaMalloc(&cyc, sizeof(long long) * nsm);
  cudaMalloc(&sink, sizeof(float) * nsm * 128);
  for (int do_min = 0; do_min < 2; ++do_min) {
    const int iters = 2000;
    tmem_ld_kernel<<<nsm, 128>>>(cyc, sink, 10, do_min);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    tmem_ld_kernel<<<nsm, 128>>>(cyc, sink, iters, do_min);
    cudaEventRecord(b);
    cudaError_t e = cudaEventSynchronize(b);
    float ms = 0; cudaEventElapsedTime(&ms, a, b);
    long long h = 0; cudaMemcpy(&h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    const double bytes_per_sm = (double)iters * 16 * 32 * 4 * 128;       // 16 loads x 32 cols x 4 B x 128 lanes
    printf("do_min=%d: %s, %.3f ms, %lld cycles/SM, %.1f B/clk/SM, %.1f fp32/clk/SM, %.2f TB/s chip\n", do_min,
           cudaGetErrorString(e), ms, h, bytes_per_sm / h, bytes_per_sm / h / 4, bytes_per_sm * nsm / (ms * 1e-3) / 1e12);
  }
  return 0;
}
