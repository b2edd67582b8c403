// psim_allpairs_tc.cuh — all-pairs force pass with the distance filter on the
// 5th-generation tensor cores (tcgen05 + TMEM).
//
// The conservative filter of psim_allpairs2.cuh,
//     est(i, j) = w_j + x_i (-2 x_j) + y_i (-2 y_j) + z_i (-2 z_j)  <  cut_i ,
// is one entry of a Gram matrix: row i of A = (x_i, y_i, z_i, 1), row j of
// B = (-2 x_j, -2 y_j, -2 z_j, w_j).  A block owns 128 i-particles (the M = 128
// rows of the accumulator, one TMEM lane each) and walks all j in tiles of
// N = 256: one elected thread issues tcgen05.mma (kind::tf32, operands in
// shared memory in the canonical K-major no-swizzle layout, descriptors built
// by hand), the 128 x 256 FP32 result lands in TMEM, and the block reads it
// back with tcgen05.ld — every thread owns one row and one half of the columns —
// keeping only a running minimum against cut_i.  TMEM is double-buffered
// (2 x 256 columns), so the MMA of tile t+1 runs while tile t is reduced.
//
// Precision.  TF32 keeps 11 significand bits, far too few for coordinates of
// ~100 and a cutoff of 2.5.  Each float is therefore split into two TF32 pieces
// (x = x1 + x2 + r, |r| <= 2^-22 |x|) and the K dimension carries the three
// largest cross products per coordinate:
//     A row: x1 x1 x2 | y1 y1 y2 | z1 z1 z2 | 1  1  1  | 0 0 0 0      (K = 16)
//     B row: X1 X2 X1 | Y1 Y2 Y1 | Z1 Z2 Z1 | w1 w2 w3 | 0 0 0 0      (X = -2 x ...)
// All products are exact in FP32 (11 x 11 bits) and the tensor core accumulates
// in FP32, so est differs from the FFMA2 filter's value by the dropped terms
// (<= 3 * 2^-21 |x_i||x_j| per pair) plus accumulation rounding; the weight's
// relative margin is widened from 32 u to 256 u to cover both.  As before the
// filter only ever passes EXTRA pairs; every pair that passes is decided and
// evaluated exactly, in the working precision, by pair_force().
//
// Measured inputs to this design (tools/microbench/): tcgen05.ld delivers
// >= 518 B/clk/SM (128 FP32 per clock per SM) with two loads in flight, about
// six times the pair rate of the FFMA2 kernel; the hand-built descriptors are
// validated against a CPU Gram matrix in tools/microbench/tc_gram.cu.
#pragma once

#include "psim_allpairs.cuh"
#include "psim_allpairs2.cuh"

namespace psim {
namespace tc {

constexpr int M = 128;        // i-particles per block = accumulator rows = TMEM lanes
constexpr int N = 256;        // j records per tile = accumulator columns
constexpr int K = 16;         // split-TF32 inner dimension (two MMAs of K = 8)
constexpr int T = 256;        // threads: warp w reads TMEM lanes 32 (w % 4) .., columns 128 (w / 4) ..

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// rows x 16 tf32 as two K-blocks of 8; inside a K-block 8-row groups are 256 B apart (SBO) and the
// two 16-byte K-chunks of a row 128 B apart (LBO); rows of a group 16 B apart: the canonical
// K-major SWIZZLE_NONE layout of a UMMA shared-memory descriptor.
__device__ __forceinline__ unsigned chunk_off(int rows, int r, int chunk) {   // chunk = k / 4, 0..3
  return (unsigned)((chunk >> 1) * rows * 32 + (r >> 3) * 256 + (chunk & 1) * 128 + (r & 7) * 16);
}
__device__ __forceinline__ unsigned long long make_desc(unsigned addr) {
  const unsigned long long lbo = 128 >> 4, sbo = 256 >> 4;
  return (unsigned long long)((addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);   // version 1 (sm_100)
}

__device__ __forceinline__ float tf32_hi(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ void split2(float x, float& a, float& b) { a = tf32_hi(x); b = tf32_hi(x - a); }
__device__ __forceinline__ void split3(float x, float& a, float& b, float& c) {
  a = tf32_hi(x); const float r = x - a; b = tf32_hi(r); c = tf32_hi(r - b);
}

// filter record with the wider margin of the tensor-core path
template <class R>
__device__ __forceinline__ FilterRec filter_record_tc(const Real4<R>& p, const Consts<R>& c, float rcut) {
  constexpr float u32 = 5.9604645e-8f;
  FilterRec f;
  f.x = (float)p.x; f.y = (float)p.y; f.z = (float)p.z;
  const float n2 = f.x * f.x + f.y * f.y + f.z * f.z;
  f.w = n2 * (1.0f - 256.0f * u32) - 16.0f * u32 * rcut * (sqrtf(n2) + 1.0f);
  if (p.w >= c.huge) f.w = 1.0e30f;
  return f;
}

__device__ __forceinline__ void mbar_wait(unsigned addr, unsigned parity) {
  unsigned done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  }
}

#define PSIM_TMEM_LD32(r, taddr)                                                                                   \
  asm volatile(                                                                                                    \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                    \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                    \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"                  \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),            \
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),      \
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),    \
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])     \
      : "r"(taddr))

// exact decision and evaluation of one pair, out of line (rare)
template <class R>
__device__ __noinline__ void exact_one(const Real4<R>* __restrict__ pos, const Consts<R>& c, int64_t ig, int64_t jg,
                                       R xi, R yi, R zi, R* f) {
  if (jg == ig) return;
  const Real4<R> pj = pos[jg];
  if (pj.w >= c.huge) return;
  const R dx = r_sub(xi, pj.x), dy = r_sub(yi, pj.y), dz = r_sub(zi, pj.z);
  const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
  if (r2 < c.cut2_pad) pair_force(xi, yi, zi, pj.x, pj.y, pj.z, c, f[0], f[1], f[2]);
}

// one 32-column chunk of this thread's row: running minima, then the rare descent to single pairs
template <class R>
__device__ __forceinline__ void reduce_chunk(const unsigned (&r)[32], float cut, const Real4<R>* pos,
                                             const Consts<R>& c, int64_t ig, int64_t j0, R xi, R yi, R zi, R* f) {
  float m8[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const float a = fminf(__uint_as_float(r[8 * g]), fminf(__uint_as_float(r[8 * g + 1]), __uint_as_float(r[8 * g + 2])));
    const float b = fminf(__uint_as_float(r[8 * g + 3]), fminf(__uint_as_float(r[8 * g + 4]), __uint_as_float(r[8 * g + 5])));
    m8[g] = fminf(fminf(a, b), fminf(__uint_as_float(r[8 * g + 6]), __uint_as_float(r[8 * g + 7])));
  }
  const float m = fminf(fminf(m8[0], m8[1]), fminf(m8[2], m8[3]));
  if (m < cut) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (m8[g] < cut) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
          if (__uint_as_float(r[8 * g + u]) < cut) exact_one(pos, c, ig, j0 + 8 * g + u, xi, yi, zi, f);
      }
    }
  }
}

template <class R, int MODE>
__global__ void __launch_bounds__(T, 1)
allpairs_tc_kernel(const StepArgs<R> a) {
  __shared__ __align__(128) unsigned char sA[M * K * 4];          //  8 KB
  __shared__ __align__(128) unsigned char sB[2][N * K * 4];       // 32 KB
  __shared__ __align__(8) unsigned long long mbar[2];
  __shared__ unsigned tmem_base_s;
  __shared__ R fhalf[3][M];

  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row = (warp & 3) * 32 + lane;          // accumulator row = TMEM lane of this thread
  const int half = warp >> 2;                      // which 128 columns of every tile
  const int64_t i_base = a.i_lo + (int64_t)blockIdx.x * M;
  const int64_t ig = i_base + row;
  const float rcut = (float)c.r_cut;
  const float cut2f = (float)c.cut2_pad * (1.0f + 1.0e-6f);

  // --- A tile (this block's i-particles), cut value and exact position of this thread's row ---
  float cut = -1.0e30f;
  R xi = (R)0, yi = (R)0, zi = (R)0;
  {
    FilterRec f; f.x = f.y = f.z = 0.f; f.w = 1.0e30f;
    if (ig < a.i_hi) {
      const Real4<R> p = a.pos_cur[ig];
      xi = p.x; yi = p.y; zi = p.z;
      f = filter_record_tc(p, c, rcut);
    }
    if (f.w < 1.0e30f) cut = cut2f - f.w;
    if (half == 0) {
      float x1, x2, y1, y2, z1, z2;
      split2(f.x, x1, x2); split2(f.y, y1, y2); split2(f.z, z1, z2);
      *reinterpret_cast<float4*>(sA + chunk_off(M, row, 0)) = make_float4(x1, x1, x2, y1);
      *reinterpret_cast<float4*>(sA + chunk_off(M, row, 1)) = make_float4(y1, y2, z1, z1);
      *reinterpret_cast<float4*>(sA + chunk_off(M, row, 2)) = make_float4(z2, 1.f, 1.f, 1.f);
      *reinterpret_cast<float4*>(sA + chunk_off(M, row, 3)) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  auto stage_b = [&](int buf, const Real4<R>& pj) {   // thread tid stages its record of the tile
    const FilterRec f = filter_record_tc(pj, c, rcut);
    float x1, x2, y1, y2, z1, z2, w1, w2, w3;
    split2(-2.0f * f.x, x1, x2); split2(-2.0f * f.y, y1, y2); split2(-2.0f * f.z, z1, z2);
    split3(f.w, w1, w2, w3);
    unsigned char* b = sB[buf];
    *reinterpret_cast<float4*>(b + chunk_off(N, tid, 0)) = make_float4(x1, x2, x1, y1);
    *reinterpret_cast<float4*>(b + chunk_off(N, tid, 1)) = make_float4(y2, y1, z1, z2);
    *reinterpret_cast<float4*>(b + chunk_off(N, tid, 2)) = make_float4(z1, w1, w2, w3);
    *reinterpret_cast<float4*>(b + chunk_off(N, tid, 3)) = make_float4(0.f, 0.f, 0.f, 0.f);
  };

  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&mbar[1])));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  const int nt = (int)(a.n_pad / N);
  stage_b(0, a.pos_cur[tid]);
  // the record of tile t+1 is loaded one iteration before it is staged: its latency hides behind
  // the reduction of tile t-1
  Real4<R> pre = a.pos_cur[(nt > 1 ? N : 0) + tid];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base_s;
  // c_format F32, a/b TF32, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
  const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
  auto issue_mma = [&](int buf) {                    // one thread: two K = 8 MMAs into TMEM buffer `buf`, then commit
#pragma unroll
    for (int kb = 0; kb < 2; ++kb) {
      const unsigned long long da = make_desc(smem_u32(sA) + kb * M * 32);
      const unsigned long long db = make_desc(smem_u32(sB[buf]) + kb * N * 32);
      asm volatile(
          "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
          :: "r"(tm + (unsigned)(buf * N)), "l"(da), "l"(db), "r"(idesc), "r"((unsigned)kb), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 :: "r"(smem_u32(&mbar[buf])) : "memory");
  };
  if (tid == 0) issue_mma(0);

  R f[3] = {(R)0, (R)0, (R)0};
  const unsigned lane_base = tm + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)(half * 128);
  for (int t = 0; t < nt; ++t) {
    const int buf = t & 1;
    // stage the next tile (its buffer was read by MMA t-1, whose completion every thread has seen)
    if (t + 1 < nt) stage_b(buf ^ 1, pre);
    if (t + 2 < nt) pre = a.pos_cur[(int64_t)(t + 2) * N + tid];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");          // orders this thread's tcgen05.ld of tile t-1
    __syncthreads();
    if (t + 1 < nt && tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;");
      issue_mma(buf ^ 1);                                          // runs while tile t is reduced below
    }
    mbar_wait(smem_u32(&mbar[buf]), (unsigned)((t >> 1) & 1));
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tb = lane_base + (unsigned)(buf * N);
    const int64_t j0 = (int64_t)t * N + half * 128;
#pragma unroll
    for (int cc = 0; cc < 4; cc += 2) {
      unsigned r0[32], r1[32];
      PSIM_TMEM_LD32(r0, tb + cc * 32);
      PSIM_TMEM_LD32(r1, tb + (cc + 1) * 32);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      reduce_chunk<R>(r0, cut, a.pos_cur, c, ig, j0 + cc * 32, xi, yi, zi, f);
      reduce_chunk<R>(r1, cut, a.pos_cur, c, ig, j0 + (cc + 1) * 32, xi, yi, zi, f);
    }
  }

  // --- the two column halves of a row meet; epilogue by the thread of half 0 ---
  asm volatile("tcgen05.fence::before_thread_sync;");
  if (half == 1) { fhalf[0][row] = f[0]; fhalf[1][row] = f[1]; fhalf[2][row] = f[2]; }
  __syncthreads();
  R first = (R)0, second = (R)0;
  if (half == 0 && ig < a.i_hi)
    step_epilogue<R, MODE>(a, ig, r_add(f[0], fhalf[0][row]), r_add(f[1], fhalf[1][row]), r_add(f[2], fhalf[2][row]),
                           first, second);
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, T>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

}  // namespace tc
}  // namespace psim
