// psim_allpairs_tc.cuh — the default all-pairs force pass: the distance filter of every ordered pair on the
// 5th-generation tensor cores (tcgen05 + TMEM), the exact pair arithmetic on the CUDA cores for the few
// pairs that pass.
//
// Mathematics.  The conservative filter of psim_allpairs2.cuh,
//     est(i, j) = w_j + y_i . (-2 y_j)  <  cut_i ,      y = position relative to the box centre,
// is one entry of a Gram matrix: row i of A = (y_i, 1), row j of B = (-2 y_j, w_j).  A block owns M = 128
// i-particles (the rows of the accumulator, one TMEM lane each) and walks the j records in tiles of N2 = 128.
// TF32 keeps 11 significand bits, far too few for coordinates of ~100 and a cutoff of 2.5, so each float is
// split into two TF32 pieces (x = x1 + x2 + r, |r| < 2^-20 |x|) and K carries the three largest cross
// products per coordinate:
//     A row: x1 x1 x2 | y1 y1 y2 | z1 z1 z2 | 1  1  1  | 0 0 0 0      (K = 16 = two MMAs of K = 8)
//     B row: X1 X2 X1 | Y1 Y2 Y1 | Z1 Z2 Z1 | w1 w2 w3 | 0 0 0 0      (X = -2 x ...)
// All products are exact (11 x 11 bits) and the tensor core accumulates in FP32, so est differs from the
// packed-FP32 filter's value by the dropped terms (< 2^-18 (|y_i|^2 + |y_j|^2)) plus accumulation rounding;
// the weight's relative margin is 256 u (u = 2^-24) to cover both, and an absolute term 16 u r_cut (|y| + 1)
// covers the rounding of the coordinates themselves (tests/test_filter_math.py restates this in numpy and
// checks the margin under a worst-case accumulation model).  Coordinates are taken relative to the box
// centre: the margins scale with |y|^2, so this quarters the false positives of a corner particle.  The
// filter only ever passes EXTRA pairs; every pair that passes is decided and evaluated exactly, in the
// working precision, with the reference's own arithmetic (pair_force, psim_common.cuh) — and the kernel
// counts the ordered in-range pairs it evaluated, which the tests compare with the exact O(N^2) count.
//
// Organisation (one block of 21 warps per SM; all 512 TMEM columns as four 128-column accumulators):
//   * tc_prepare_kernel builds the B operand ONCE per force pass, in global memory, already in the UMMA
//     shared-memory tile layout (8 KB per tile of 128 j records);
//   * a producer warp streams those tiles into a 4-slot shared-memory ring with TMA bulk copies
//     (cp.async.bulk, completion on an mbarrier, 16 KB = two tiles per copy);
//   * four MMA warps (one elected lane each, one accumulator each) issue the tcgen05.mma pairs — one
//     thread doing all of it sustains only ~1 tile per 950 cycles (measured, tools/microbench/tc_gram.cu);
//   * sixteen consumer warps read their 32 rows x 32 columns of every tile with one tcgen05.ld.32x32b.x32,
//     hand the accumulator back, reduce the 32 values with 16 three-input minima and compare with cut_i.
//     The consumer tile loop is a function of its own (consume_tiles) so that ptxas gives it a register
//     allocation with two sets of 32 values: the load of tile t + 1 is in flight while tile t is reduced.
//     Rare hits go to a per-thread 16-bit queue in shared memory; the exact path runs between calls of
//     consume_tiles, never while an accumulator is held;
//   * rings of mbarriers (B full/empty, accumulator full/empty) replace every block-wide barrier;
//   * j-split: when a launch has too few i-blocks to fill the SMs evenly (a shard of a multi-GPU run, a
//     small N) the j range is cut into S segments, one block each; the partial forces of an i-block meet
//     in global scratch and the block that arrives last adds them in segment order (fixed order, so the
//     result is run-to-run identical) and runs the kick/drift epilogue.
#pragma once

#include "psim_allpairs.cuh"
#include "psim_allpairs2.cuh"

namespace psim {
namespace tc {

constexpr int M = 128;                          // i-particles per block = accumulator rows = TMEM lanes
constexpr int K = 16;                           // split-TF32 inner dimension (two MMAs of K = 8)
constexpr int N2 = 128;                         // j records per tile = accumulator columns per buffer
constexpr int NBUF = 512 / N2;                  // TMEM ring: four accumulators of 128 columns
constexpr int NCTL = NBUF;                      // MMA warps: warp w owns accumulator w, tiles t = w (mod NCTL)
constexpr int NPAIR = 4;                        // shared-memory ring: 4 slots of two tiles (one 16 KB bulk copy each)
// NPAIR must be EVEN.  Tile pair P lives in slot P % NPAIR and is consumed by MMA warps 0,1 (P even) or 2,3
// (P odd).  With an odd ring (round 1's experimental kernel had 3) consecutive phases of a slot's b_full
// barrier belong to ALTERNATING warp sets: a warp waits on phase n and next on phase n + 2, the same parity.
// mbarrier.try_wait.parity(p) only says "the phase with parity p is over", so whenever the copy of phase
// n + 1 had not landed yet the wait for n + 2 passed at once: the warp multiplied a stale tile and arrived on
// b_empty a phase early (round 1's intermittent launch failure / hang).  With an even ring a slot is always
// used by the same warp set, phase after phase.
static_assert(NPAIR % 2 == 0, "odd ring: phase-parity aliasing on b_full (see above)");
constexpr int STAGES = 2 * NPAIR;               // 8 x 8 KB
constexpr int TILE_BYTES = N2 * K * 4;          // 8 KB
constexpr int NQ = 4;                           // column groups: warp w reads TMEM lanes 32 (w % 4).., columns 32 (w / 4)..
constexpr int CW = 4 * NQ;                      // consumer warps
constexpr int TW = 32 * (CW + NCTL + 1);        // + the MMA warps + the TMA producer warp
constexpr int QC = 96;                          // logical capacity of a consumer thread's candidate queue ...
constexpr int QPHYS = QC + 32;                  // ... + room for one tile's unbounded first pass (queue_chunk): 128 KB in all
constexpr int QMAX_TILE = 5;                    // logical entries one tile can add: up to 5 columns, or 1 dense chunk
constexpr int QSTRIDE = 32 * CW;                // entry e of thread tid at q[e * QSTRIDE + tid] (16-bit): conflict-free
constexpr int SEG = 1024;                       // tiles per consume_tiles call: a queue entry holds a 10-bit tile offset
constexpr int RING_BYTES = STAGES * TILE_BYTES;
constexpr int SMEM_BYTES = RING_BYTES + QPHYS * QSTRIDE * 2;
constexpr unsigned DENSE = 0x8000u;             // queue entry: [dense:1][tile - t_begin:10][column:5]; dense = all 32 columns
constexpr int MAX_JSPLIT = 8;                   // j segments per i-block (launch-time choice)

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

// Filter record of one particle: float coordinates RELATIVE TO THE BOX CENTRE and the weight with the
// tensor-core path's margins (header).  The subtraction is done in the working precision.
template <class R>
__device__ __forceinline__ FilterRec filter_record_tc(const Real4<R>& p, const Consts<R>& c, float rcut) {
  constexpr float u32 = 5.9604645e-8f;
  FilterRec f;
  f.x = (float)r_sub(p.x, r_mul((R)0.5, c.area[0]));
  f.y = (float)r_sub(p.y, r_mul((R)0.5, c.area[1]));
  f.z = (float)r_sub(p.z, r_mul((R)0.5, c.area[2]));
  const float n2 = f.x * f.x + f.y * f.y + f.z * f.z;
  f.w = n2 * (1.0f - 256.0f * u32) - 16.0f * u32 * rcut * (sqrtf(n2) + 1.0f);
  if (p.w >= c.huge) f.w = 1.0e30f;          // inactive or padding: never a candidate
  return f;
}

__device__ __forceinline__ void mbar_init(unsigned addr, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(addr), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(addr) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned addr, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned addr, unsigned parity) {
  unsigned done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  }
}
// Wait for an mbarrier phase with the whole warp and leave the spin loop in the SAME iteration in all lanes:
// a .sync.aligned tcgen05 instruction follows (the vote is a real warp-synchronous instruction; a bare
// __syncwarp() after mbar_wait is elided by ptxas when the loop state is uniform).
__device__ __forceinline__ void mbar_wait_warp(unsigned addr, unsigned parity) {
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  } while (!__all_sync(0xffffffffu, done));
}
// TMA bulk copy global -> shared (this GPU's memory only), completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_copy_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned mbar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(mbar) : "memory");
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

// B operand rows of all particles, tile by tile, in the canonical layout of chunk_off(); also clears the
// pair counter of the force kernel that follows.
template <class R>
__global__ void __launch_bounds__(256)
tc_prepare_kernel(const Real4<R>* __restrict__ pos, unsigned char* __restrict__ bops, int64_t n_pad, Consts<R> c,
                  unsigned long long* __restrict__ pair_count) {
  const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (j == 0) *pair_count = 0ull;                  // the force kernel that follows counts what its exact path evaluates
  if (j >= n_pad) return;
  const FilterRec f = filter_record_tc(pos[j], c, (float)c.r_cut);
  float x1, x2, y1, y2, z1, z2, w1, w2, w3;
  split2(-2.0f * f.x, x1, x2); split2(-2.0f * f.y, y1, y2); split2(-2.0f * f.z, z1, z2);
  split3(f.w, w1, w2, w3);
  unsigned char* b = bops + (j / N2) * TILE_BYTES;
  const int r = (int)(j % N2);
  *reinterpret_cast<float4*>(b + chunk_off(N2, r, 0)) = make_float4(x1, x2, x1, y1);
  *reinterpret_cast<float4*>(b + chunk_off(N2, r, 1)) = make_float4(y2, y1, z1, z2);
  *reinterpret_cast<float4*>(b + chunk_off(N2, r, 2)) = make_float4(z1, w1, w2, w3);
  *reinterpret_cast<float4*>(b + chunk_off(N2, r, 3)) = make_float4(0.f, 0.f, 0.f, 0.f);
}

// The exact path: one queued candidate j of row ig in the working precision, with the reference's own
// arithmetic (pair_force, psim_common.cuh).
template <class R>
__device__ __noinline__ void exact_pair(const Real4<R>* __restrict__ pos, const Consts<R>& c, int64_t ig, const Real4<R>& pi,
                                        int64_t jg, R* f, unsigned* in_range) {
  if (jg == ig) return;
  const Real4<R> pj = pos[jg];
  if (pj.w >= c.huge) return;
  const R dx = r_sub(pi.x, pj.x), dy = r_sub(pi.y, pj.y), dz = r_sub(pi.z, pj.z);
  const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
  if (r2 < c.cut2_pad) {
    pair_force(pi.x, pi.y, pi.z, pj.x, pj.y, pj.z, c, f[0], f[1], f[2]);
    if (r_sqrt(r2) < c.r_cut) ++*in_range;         // the predicate of count_pairs_kernel (psim_allpairs.cuh)
  }
}

// the queue is addressed in the shared window (one register instead of a generic pointer's two)
__device__ __forceinline__ void q_store(unsigned qaddr, int e, unsigned v) {
  asm volatile("st.shared.u16 [%0], %1;" :: "r"(qaddr + (unsigned)(e * (QSTRIDE * 2))), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ unsigned q_load(unsigned qaddr, int e) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(qaddr + (unsigned)(e * (QSTRIDE * 2))) : "memory");
  return v;
}

// What the exact path needs, kept on the kernel's stack so that consume_tiles carries ONE pointer for it.
template <class R>
struct FlushCtx {
  const Real4<R>* pos; const Consts<R>* c; int64_t ig;
  int64_t col0;              // global index of the first column of this thread's group in tile 0 of the block's j segment
  R f[3]; unsigned in_range; // force on row ig; ordered in-range pairs evaluated
};

// seg_begin: the tile (relative to the block's j segment) that queue codes are relative to.
template <class R>
__device__ __noinline__ void flush_queue(FlushCtx<R>* x, unsigned q, int cnt, int seg_begin) {
  if (cnt == 0) return;                            // rows past i_hi / inactive rows never queue anything
  const Real4<R>* __restrict__ pos = x->pos;
  const Consts<R>& c = *x->c;
  const int64_t ig = x->ig;
  const Real4<R> pi = pos[ig];
  for (int e = 0; e < cnt; ++e) {
    const unsigned v = q_load(q, e);
    const int64_t j0 = (int64_t)(seg_begin + (int)((v >> 5) & 1023u)) * N2 + x->col0;
    if (v & DENSE) {
      for (int u = 0; u < 32; ++u) exact_pair(pos, c, ig, pi, j0 + u, x->f, &x->in_range);
    } else {
      exact_pair(pos, c, ig, pi, j0 + (int64_t)(v & 31u), x->f, &x->in_range);
    }
  }
}

__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(a, fminf(b, c)); }   // one FMNMX3

// 32 columns of this thread's row: 16 three-input minima (groups of 9, 9, 9 and 5 columns), one compare.
// The rare hit pushes the columns below cut (tile-relative code `tcode`); more than QMAX_TILE of them are
// replaced by ONE dense entry, so a tile adds at most QMAX_TILE logical entries (what lets consume_tiles
// bound the queue) — the first pass may write up to 32 entries past cnt, hence QPHYS.
__device__ __forceinline__ void queue_chunk(const unsigned (&r)[32], float cut, unsigned tcode, unsigned q, int& cnt) {
  float mg[4];
#pragma unroll
  for (int g = 0; g < 3; ++g) {
#define PSIM_V(k) __uint_as_float(r[9 * g + (k)])
    mg[g] = min3f(min3f(PSIM_V(0), PSIM_V(1), PSIM_V(2)), min3f(PSIM_V(3), PSIM_V(4), PSIM_V(5)), min3f(PSIM_V(6), PSIM_V(7), PSIM_V(8)));
#undef PSIM_V
  }
  mg[3] = min3f(min3f(__uint_as_float(r[27]), __uint_as_float(r[28]), __uint_as_float(r[29])), __uint_as_float(r[30]),
                __uint_as_float(r[31]));
  const float m = fminf(min3f(mg[0], mg[1], mg[2]), mg[3]);
  if (m < cut) {
    const int c0 = cnt;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (mg[g] < cut) {
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          const int u = 9 * g + k;                 // static after unrolling: r[] must stay in registers
          if (u < 32 && __uint_as_float(r[u < 32 ? u : 0]) < cut) {
            q_store(q, cnt, tcode + u);
            ++cnt;
          }
        }
      }
    }
    if (cnt - c0 > QMAX_TILE) {
      q_store(q, c0, tcode | DENSE);
      cnt = c0 + 1;
    }
  }
}

// The consumer tile loop, a function of its own so that ptxas gives it its own register allocation
// (a few scalars + two sets of 32 accumulator values): the tcgen05.ld of tile t + 1 is in flight while
// tile t is reduced.  tcgen05.wait::ld waits for ALL loads of the thread, hence the order
// wait -> hand the accumulator back -> start the next load -> reduce.  Unrolled over the four accumulators
// so that buffer offsets are immediates.
// The loop state must live in UNIFORM registers (measured: with vector loop state ptxas wraps every wait in
// divergence barriers and spills, 1.3x slower).  The tile count and the start tile arrive in vector
// registers (function arguments), so both are passed through a warp reduction (REDUX writes a uniform
// register) — they are the same in all lanes anyway.
// Returns at a group boundary with no load in flight: at the end, every SEG tiles (a queue code holds a
// 10-bit tile offset), and when some lane's queue could overflow during the next two groups (one is already
// committed when the decision is taken).  The caller flushes the queue and calls again.  A flush in
// mid-stream stalls the accumulator ring, so the capacity is sized to make the last kind rare.  The flush
// is NOT called from here: values live across a call must sit above the callee chain's registers, and with
// pair_force<double> below that left too few for this loop.
// Returns (next tile << 32) | queue length.  Queue codes are relative to t_begin.  Tile numbers are relative
// to the block's j segment.
// dbar0: shared address of d_full[0]; d_empty[b] lives 8 * NBUF bytes after d_full[b].  nt: multiple of 8.
__device__ __noinline__ unsigned long long consume_tiles(float cut, unsigned q, unsigned lane_base, unsigned dbar0, int nt_v, int t_begin_v) {
  const int lane = threadIdx.x & 31;
  const int t_begin = (int)__reduce_max_sync(0xffffffffu, (unsigned)t_begin_v);
  const int nt = (int)__reduce_max_sync(0xffffffffu, (unsigned)nt_v);
#define PSIM_TC_ACQUIRE(r, buf, par_)                                                            \
  {                                                                                              \
    mbar_wait_warp(dbar0 + 8u * (buf), (par_));   /* all lanes leave together: .aligned load next */ \
    asm volatile("tcgen05.fence::after_thread_sync;");                                           \
    PSIM_TMEM_LD32(r, lane_base + (unsigned)((buf) * N2));                                       \
  }
#define PSIM_TC_RELEASE(buf)                                                                     \
  {                                                                                              \
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");                                 \
    asm volatile("tcgen05.fence::before_thread_sync;");   /* values are in registers */          \
    __syncwarp();                                                                                \
    if (lane == 0) mbar_arrive(dbar0 + 8u * (NBUF + (buf)));                                     \
  }
  unsigned rA[32], rB[32];
  int cnt = 0, t0 = t_begin;
  PSIM_TC_ACQUIRE(rA, 0, (unsigned)((t0 / NBUF) & 1))
  for (;;) {
    const unsigned par = (unsigned)((t0 / NBUF) & 1);
    const unsigned tc0 = (unsigned)(t0 - t_begin) << 5;
    // this group adds at most 4 QMAX_TILE logical entries, and so does the next one if we go on
    const bool more = (t0 + NBUF < nt) && (t0 + NBUF - t_begin < SEG) &&
                      !__any_sync(0xffffffffu, cnt > QC - 8 * QMAX_TILE);
    PSIM_TC_RELEASE(0) PSIM_TC_ACQUIRE(rB, 1, par)
    queue_chunk(rA, cut, tc0, q, cnt);
    PSIM_TC_RELEASE(1) PSIM_TC_ACQUIRE(rA, 2, par)
    queue_chunk(rB, cut, tc0 + 32u, q, cnt);
    PSIM_TC_RELEASE(2) PSIM_TC_ACQUIRE(rB, 3, par)
    queue_chunk(rA, cut, tc0 + 64u, q, cnt);
    PSIM_TC_RELEASE(3)
    if (more) PSIM_TC_ACQUIRE(rA, 0, par ^ 1u)
    queue_chunk(rB, cut, tc0 + 96u, q, cnt);
    t0 += NBUF;
    if (!more) break;
  }
#undef PSIM_TC_ACQUIRE
#undef PSIM_TC_RELEASE
  return ((unsigned long long)(unsigned)t0 << 32) | (unsigned)cnt;
}

// Scratch of the j-split (S > 1): partial forces of every (i-block, segment) and one arrival counter per
// i-block (zero between launches: the combining block clears it).
template <class R>
struct JSplit {
  int S;                     // segments per i-block (1: no split, scratch unused)
  R* part;                   // [gridDim.x][3][M]
  unsigned* arrived;         // [i-blocks]
};

template <class R, int MODE>
__global__ void __launch_bounds__(TW, 1)
allpairs_tc_kernel(const StepArgs<R> a, const unsigned char* __restrict__ bops, unsigned long long* __restrict__ pair_count,
                   const JSplit<R> js) {
  extern __shared__ __align__(1024) unsigned char sB[];          // ring of B tiles, then the candidate queues
  __shared__ __align__(128) unsigned char sA[M * K * 4];          // 8 KB
  __shared__ __align__(8) unsigned long long b_full[NPAIR], b_empty[NPAIR], d_bar[2 * NBUF];   // d_bar: full[0..3], empty[0..3]
  unsigned long long* const d_full = d_bar;
  unsigned long long* const d_empty = d_bar + NBUF;
  __shared__ unsigned tmem_base_s;
  __shared__ int last_s;
  __shared__ R fpart[NQ][3][M];

  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool consumer = warp < CW;
  const int row = (warp & 3) * 32 + lane;          // accumulator row = TMEM lane (consumers)
  const int half = (warp >> 2) % NQ;               // which 32 columns of every tile (consumers)
  const int S = js.S;
  const int ib = (int)blockIdx.x / S, seg = (int)blockIdx.x - ib * S;
  const int64_t ig = a.i_lo + (int64_t)ib * M + row;
  const float rcut = (float)c.r_cut;
  const float cut2f = (float)c.cut2_pad * (1.0f + 1.0e-6f);
  // this block's j segment, in tiles (multiples of 8, so that every ring keeps its phase arithmetic)
  const int nt8 = (int)(a.n_pad / (8 * N2));
  const int t_lo = 8 * ((seg * nt8) / S), t_hi = 8 * (((seg + 1) * nt8) / S);   // seg * nt8 < 2^31: nt8 = n_pad / 1024
  const int nt = t_hi - t_lo;

  // --- A tile (this block's i-particles) and the cut value of this thread's row ---
  float cut = -1.0e30f;
  if (consumer) {
    FilterRec f; f.x = f.y = f.z = 0.f; f.w = 1.0e30f;
    if (ig < a.i_hi) f = filter_record_tc(a.pos_cur[ig], c, rcut);
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
  if (tid == 0) {
    for (int s = 0; s < NPAIR; ++s) { mbar_init(smem_u32(&b_full[s]), 1); mbar_init(smem_u32(&b_empty[s]), 2); }
    for (int b = 0; b < NBUF; ++b) { mbar_init(smem_u32(&d_full[b]), 1); mbar_init(smem_u32(&d_empty[b]), CW); }
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == CW) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base_s)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // A tile: generic-proxy writes -> tensor core
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base_s;

  FlushCtx<R> fx;                                  // the exact path's state of this thread (consumers)
  fx.pos = a.pos_cur; fx.c = &c; fx.ig = ig; fx.col0 = (int64_t)t_lo * N2 + half * 32;
  fx.f[0] = fx.f[1] = fx.f[2] = (R)0; fx.in_range = 0;
  if (warp == CW + NCTL) {
    if (lane == 0) {
      // ===== TMA producer: one 16 KB bulk copy per pair of tiles, as far ahead as the ring allows =====
      const int np = nt / 2;
      const unsigned bfull0 = smem_u32(&b_full[0]), bempty0 = smem_u32(&b_empty[0]), sB0 = smem_u32(sB);
      const unsigned char* src = bops + (size_t)t_lo * TILE_BYTES;
      for (int P = 0; P < np; ++P) {
        const int s = P % NPAIR;
        mbar_wait(bempty0 + 8u * s, (unsigned)(((P / NPAIR) & 1) ^ 1));   // both MMA warps done with the slot (fresh: passes)
        mbar_expect_tx(bfull0 + 8u * s, 2 * TILE_BYTES);
        bulk_copy_g2s(sB0 + (unsigned)(s * 2 * TILE_BYTES), src + (size_t)P * 2 * TILE_BYTES, 2 * TILE_BYTES, bfull0 + 8u * s);
      }
    }
  } else if (warp >= CW) {
    if (lane == 0) {
      // ===== MMA warp w: issues the tiles t = w, w + NCTL, ... into accumulator w =====
      const int w = warp - CW;
      const int ntw = (nt - w + NCTL - 1) / NCTL;
      const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N2 >> 3) << 17) | ((unsigned)(M >> 4) << 24);
      const unsigned long long da0 = make_desc(smem_u32(sA)), da1 = make_desc(smem_u32(sA) + M * 32);
      const unsigned acc = tm + (unsigned)(w * N2);
      const unsigned dfull = smem_u32(&d_full[w]), dempty = smem_u32(&d_empty[w]);
      const unsigned bfull0 = smem_u32(&b_full[0]), bempty0 = smem_u32(&b_empty[0]), sB0 = smem_u32(sB);
      for (int k = 0; k < ntw; ++k) {
        const int t = w + NCTL * k, P = t >> 1, s = P % NPAIR;
        mbar_wait(bfull0 + 8u * s, (unsigned)((P / NPAIR) & 1));                        // tile pair landed
        mbar_wait(dempty, (unsigned)((k & 1) ^ 1));                                     // accumulator drained
        asm volatile("tcgen05.fence::after_thread_sync;");
        const unsigned sb = sB0 + (unsigned)((2 * s + (t & 1)) * TILE_BYTES);
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const unsigned long long db = make_desc(sb + kb * N2 * 32);
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
              :: "r"(acc), "l"(kb ? da1 : da0), "l"(db), "r"(idesc), "r"((unsigned)kb),
                 "r"(0u), "r"(0u), "r"(0u), "r"(0u));
        }
        umma_commit(dfull);                       // both arrive when the two MMAs above have completed
        umma_commit(bempty0 + 8u * s);
      }
    }
  } else {
    // ===== consumers: TMEM -> registers -> minimum against cut -> candidate queue; exact path between calls =====
    const unsigned q = smem_u32(sB + RING_BYTES) + 2u * (unsigned)tid;     // this thread's queue, entry e at + e * 2 QSTRIDE bytes
    const unsigned lane_base = tm + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)(half * 32);
    const unsigned dbar0 = smem_u32(&d_bar[0]);
    int t = 0;
    while (t < nt) {
      const unsigned long long r = consume_tiles(cut, q, lane_base, dbar0, nt, t);
      flush_queue<R>(&fx, q, (int)(unsigned)r, t);
      t = (int)(r >> 32);
    }
    unsigned in_range = fx.in_range;
    for (int off = 16; off > 0; off >>= 1) in_range += __shfl_xor_sync(0xffffffffu, in_range, off);
    if (lane == 0 && in_range) atomicAdd(pair_count, (unsigned long long)in_range);
  }

  // --- the column groups of a row meet (fixed order) ---
  asm volatile("tcgen05.fence::before_thread_sync;");
  R* const f = fx.f;
  if (consumer && half > 0) { fpart[half][0][row] = f[0]; fpart[half][1][row] = f[1]; fpart[half][2][row] = f[2]; }
  __syncthreads();
  const bool owner = consumer && half == 0;        // the thread that finishes row `row`
  if (owner) {
#pragma unroll
    for (int g = 1; g < NQ; ++g) {
      f[0] = r_add(f[0], fpart[g][0][row]); f[1] = r_add(f[1], fpart[g][1][row]); f[2] = r_add(f[2], fpart[g][2][row]);
    }
  }
  if (warp == CW) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
  }
  // --- the j segments of an i-block meet in global scratch; the block that arrives last adds them in
  //     segment order and finishes the rows ---
  bool finish = true;
  if (S > 1) {
    if (owner) {
      R* p = js.part + (size_t)blockIdx.x * 3 * M;
      p[row] = f[0]; p[M + row] = f[1]; p[2 * M + row] = f[2];
      __threadfence();
    }
    __syncthreads();
    if (tid == 0) {
      const unsigned prev = atomicAdd(&js.arrived[ib], 1u);
      last_s = (prev == (unsigned)S - 1u) ? 1 : 0;
      if (last_s) js.arrived[ib] = 0u;             // every segment has arrived: clean for the next launch
    }
    __syncthreads();
    finish = last_s != 0;
    if (finish && owner) {
      __threadfence();
      const volatile R* p = js.part + (size_t)ib * S * 3 * M;
      f[0] = p[row]; f[1] = p[M + row]; f[2] = p[2 * M + row];
      for (int s = 1; s < S; ++s) {
        const volatile R* ps = p + (size_t)s * 3 * M;
        f[0] = r_add(f[0], ps[row]); f[1] = r_add(f[1], ps[M + row]); f[2] = r_add(f[2], ps[2 * M + row]);
      }
    }
  }
  R first = (R)0, second = (R)0;
  if (finish && owner && ig < a.i_hi) step_epilogue<R, MODE>(a, ig, f[0], f[1], f[2], first, second);
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, TW>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

}  // namespace tc
}  // namespace psim
