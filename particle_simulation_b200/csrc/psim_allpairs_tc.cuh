// psim_allpairs_tc.cuh — the all-pairs force pass with the distance filter as ONE kind::f16 tcgen05.mma per
// tile, FP16 accumulators in TMEM, packed TMEM loads and a sign test.
//
// Why FP16 (measured on B200, tools/microbench/tc_f16.cu, profiles/r2_microbench_tc_f16.log).  In round 2's first,
// TF32 form of this kernel (split-TF32 operands, FP32 accumulators; git history) a tile of 128 x 128 pair tests costs 128 cycles of tensor pipe (two K = 8 MMAs), and
// 128 cycles of ALU pipe per SM sub-partition (16 FMNMX3 per 32 values, one every 2 clocks).  With FP16
// operands the same K = 16 product is ONE instruction at twice the rate (64 cycles per tile); with an FP16
// accumulator tcgen05.ld .pack::16b delivers two entries per register, so the reduction handles two pairs per
// lane-operation (64 cycles per tile).  The tensor core forms the 16-term sum at (about) FP32 width and rounds
// once (measured: error 3.4e-5 on sums whose terms are ~350), so the cancellation |y_i|^2 + |y_j|^2 - 2 y_i.y_j
// survives; the FP16 rounding of the RESULT only has to keep its sign.
//
// Mathematics.  With y = position relative to the box centre, scaled by a power of two S so that |S y|^2 fits
// FP16, every float is split into two FP16 pieces (11 significant bits each), weights into three:
//     A row i: x1 x1 x2 | y1 y1 y2 | z1 z1 z2 | 1  1  1  | -C1 -C2 -C3 | 0
//     B row j: X1 X2 X1 | Y1 Y2 Y1 | Z1 Z2 Z1 | w1 w2 w3 |  1   1   1  | 0        X = -2 S x_j ...
//     acc(i, j) = w_j - 2 S^2 y_i.y_j - C_i ,   C_i = S^2 r_cut^2 (1 + d) - w_i ,   w = S^2 |y|^2 (1 - k) - a
// so acc < 0 is the filter's "may be in range".  k = 2^-16 covers the dropped cross terms
// (<= 3 * 2^-22 (|Sy_i|^2 + |Sy_j|^2)) and the accumulation error with a factor > 5 to spare; a covers the rounding of
// the coordinates to float (16 u r_cut (|y| + 1), as in the TF32 kernel) plus 2^-13 for FP16 underflow of the
// smallest pieces and of the result.  Inactive and padding records get a row that can never pass (+60000);
// a particle so far outside the box that S^2 |y|^2 would overflow FP16 (it left the box and the next rebuild
// will deactivate it) gets a row that ALWAYS passes (-60000): every pair that passes is decided and evaluated
// exactly, in the working precision, by the exact path (pair_term: pair_force's arithmetic, psim_common.cuh), so the
// forces are unchanged; the kernel counts the in-range pairs it evaluated for the tests' audit.
//
// Organisation (one block of 21 warps per SM, four 128-column FP16 accumulators in TMEM):
//   * tc_prepare_kernel writes every particle's B row (32 B) once per force pass in UMMA tile layout;
//   * the TMA producer streams 32 KB (eight tiles) per bulk copy through a 2-slot ring;
//   * MMA warp w issues the tiles t = w (mod 4) into accumulator w: one instruction and one commit per tile,
//     one b_full wait per two tiles (the B slot is handed back by a consumer lane, not by a second commit);
//   * consumer warps: four groups of four (one warp per TMEM lane quadrant); group g serves accumulator g alone,
//     i.e. the tiles t = g (mod 4).  A warp reads its 32 rows x 128 columns with two
//     tcgen05.ld.32x32b.x32.pack::16b, releases the accumulator at once and only then ORs the 64 packed
//     registers (32 LOP3): any sign bit set = a candidate; rare hits go to the 16-bit queue.  (Four warps per
//     tile instead of eight: an accumulator waits for four releases, and no warp alternates between two
//     accumulators — tools/microbench/tc_rtt.cu measures 140 against 205 cycles per tile for the two layouts.)
//   * j-split: when a launch has too few i-blocks to fill the SMs evenly (a shard of a multi-GPU run, a small N)
//     the j range is cut into S segments, one block each; the partial forces of an i-block meet in global scratch
//     and the block that arrives last adds them in segment order (fixed order: run-to-run identical results) and
//     runs the kick/drift epilogue.
//   * sharded runs: the epilogue stores x_{n+1} into this GPU's buffer and into every peer's (plain stores over
//     NVLink, PeerPush in psim_allpairs.cuh) — the position exchange rides on the kernel's own blocks;
//   * TMA bulk copies here read this GPU's own memory only (the B rows are rebuilt locally from the exchanged
//     positions).
#pragma once

#include <cuda_fp16.h>

#include "psim_allpairs.cuh"

namespace psim {
namespace tc {

#ifndef PSIM_TC_FLUSH_U
#define PSIM_TC_FLUSH_U 1       // candidates evaluated side by side in the exact path
#endif

constexpr int MAX_JSPLIT = 8;                   // j segments per i-block (launch-time choice)


constexpr int M = 128;                          // i-particles per block = accumulator rows = TMEM lanes
constexpr int K = 16;                           // FP16 inner dimension: one MMA
constexpr int N2 = 128;                         // j records per tile = accumulator columns per buffer
constexpr int NBUF = 4;                         // TMEM ring: four accumulators of 128 columns
constexpr int NCTL = NBUF;                      // MMA warps
constexpr int TILE_BYTES = N2 * K * 2;          // 4 KB
constexpr int TPS = 8;                          // tiles per ring slot (one 32 KB bulk copy); a multiple of NBUF
constexpr int NSLOT = 2;
constexpr int SLOT_BYTES = TPS * TILE_BYTES;
constexpr int RING_BYTES = NSLOT * SLOT_BYTES;  // 64 KB
constexpr int CW = 16;                          // consumer warps: (accumulator) x (4 lane quadrants)
constexpr int TW = 32 * (CW + NCTL + 1);
constexpr int QC = 64, QPHYS = QC + 64;         // queue: logical capacity + one chunk's unbounded first pass (64 columns)
constexpr int QMAX_TILE = 5;
constexpr int QSTRIDE = 32 * CW;
constexpr int SEG = 256;                        // own tiles per consume call: a queue entry holds a 9-bit chunk offset (two chunks per tile)
constexpr int SMEM_BYTES = RING_BYTES + QPHYS * QSTRIDE * 2;   // 64 KB + 128 KB
constexpr unsigned DENSE = 0x8000u;             // queue entry: [dense:1][chunk of 64 columns since k_begin:9][column:6]
constexpr float BIG = 60000.0f;                 // fits FP16; larger than any S^2 |y|^2 the scaling admits

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }


__device__ __forceinline__ unsigned long long make_desc(unsigned addr) {
  const unsigned long long lbo = 128 >> 4, sbo = 256 >> 4;
  return (unsigned long long)((addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);   // version 1 (sm_100)
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


// the queue is addressed in the shared window (one register instead of a generic pointer's two)
__device__ __forceinline__ void q_store(unsigned qaddr, int e, unsigned v) {
  asm volatile("st.shared.u16 [%0], %1;" :: "r"(qaddr + (unsigned)(e * (QSTRIDE * 2))), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ unsigned q_load(unsigned qaddr, int e) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(qaddr + (unsigned)(e * (QSTRIDE * 2))) : "memory");
  return v;
}


// Scratch of the j-split (S > 1): partial forces of every (i-block, segment) and one arrival counter per
// i-block (zero between launches: the combining block clears it).
template <class R>
struct JSplit {
  int S;                     // segments per i-block (1: no split, scratch unused)
  R* part;                   // [gridDim.x][3][M]
  unsigned* arrived;         // [i-blocks]
};


// rows x 16 halves (32 B per row): 8-row groups 256 B apart (SBO), the two 16-byte K-chunks of a row 128 B
// apart (LBO), rows of a group 16 B apart: byte-identical to one K = 8 block of the TF32 layout.
__device__ __forceinline__ unsigned chunk_off16(int r, int chunk) {   // chunk = k / 8, 0..1
  return (unsigned)((r >> 3) * 256 + chunk * 128 + (r & 7) * 16);
}

struct alignas(16) Half8 { __half2 a, b, c, d; };

__device__ __forceinline__ void hsplit2(float x, __half& a, __half& b) {
  a = __float2half_rn(x); b = __float2half_rn(x - __half2float(a));
}
__device__ __forceinline__ void hsplit3(float x, __half& a, __half& b, __half& c) {
  a = __float2half_rn(x); const float r = x - __half2float(a);
  b = __float2half_rn(r); c = __float2half_rn(r - __half2float(b));
}

// The scale: the largest power of two S with (1.4 x half diagonal of the box x S)^2 <= 60000, so that anything
// inside the box, and well outside it, has a finite FP16 weight.  Computed on the host.
struct Scale16 { float S, S2; };

// What one particle contributes to the operands.  kind: 0 = normal, 1 = never passes (inactive, padding),
// 2 = always passes (beyond the FP16 range).
struct Rec16 { float x, y, z, w; int kind; };

template <class R>
__device__ __forceinline__ Rec16 filter_record_tc(const Real4<R>& p, const Consts<R>& c, float rcut, Scale16 sc) {
  constexpr float u32 = 5.9604645e-8f;
  Rec16 f;
  const float fx = (float)r_sub(p.x, r_mul((R)0.5, c.area[0]));
  const float fy = (float)r_sub(p.y, r_mul((R)0.5, c.area[1]));
  const float fz = (float)r_sub(p.z, r_mul((R)0.5, c.area[2]));
  const float n2 = fx * fx + fy * fy + fz * fz;
  f.x = fx * sc.S; f.y = fy * sc.S; f.z = fz * sc.S;
  const float n2s = n2 * sc.S2;
  f.w = n2s * (1.0f - 1.52587890625e-5f) - sc.S2 * 16.0f * u32 * rcut * (sqrtf(n2) + 1.0f) - 1.220703125e-4f;
  f.kind = 0;
  if (!(n2s < 0.9f * BIG)) f.kind = 2;         // also catches NaN / inf coordinates
  if (p.w >= c.huge) f.kind = 1;
  return f;
}

// B operand rows of all particles, tile by tile (4 KB per 128 records), in the layout of chunk_off16(); also
// clears the pair counter of the force kernel that follows.
template <class R>
__global__ void __launch_bounds__(256)
tc_prepare_kernel(const Real4<R>* __restrict__ pos, unsigned char* __restrict__ bops, int64_t n_pad, Consts<R> c,
                    Scale16 sc, unsigned long long* __restrict__ pair_count) {
  const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (j == 0) *pair_count = 0ull;
  if (j >= n_pad) return;
  const Rec16 f = filter_record_tc(pos[j], c, (float)c.r_cut, sc);
  const __half zero = __float2half_rn(0.f), one = __float2half_rn(1.f);
  __half x1 = zero, x2 = zero, y1 = zero, y2 = zero, z1 = zero, z2 = zero, w1, w2 = zero, w3 = zero;
  if (f.kind == 0) {
    hsplit2(-2.0f * f.x, x1, x2); hsplit2(-2.0f * f.y, y1, y2); hsplit2(-2.0f * f.z, z1, z2);
    hsplit3(f.w, w1, w2, w3);
  } else {
    w1 = __float2half_rn(f.kind == 1 ? BIG : -BIG);
  }
  unsigned char* b = bops + (j / N2) * TILE_BYTES;
  const int r = (int)(j % N2);
  Half8 c0, c1;
  c0.a = __halves2half2(x1, x2); c0.b = __halves2half2(x1, y1); c0.c = __halves2half2(y2, y1); c0.d = __halves2half2(z1, z2);
  c1.a = __halves2half2(z1, w1); c1.b = __halves2half2(w2, w3); c1.c = __halves2half2(one, one); c1.d = __halves2half2(one, zero);
  *reinterpret_cast<Half8*>(b + chunk_off16(r, 0)) = c0;
  *reinterpret_cast<Half8*>(b + chunk_off16(r, 1)) = c1;
}

// A operand row of one i-particle (same record, the other role).
__device__ __forceinline__ void store_a_row(unsigned char* sA, int row, const Rec16& f, float cut2s) {
  const __half zero = __float2half_rn(0.f), one = __float2half_rn(1.f);
  __half x1 = zero, x2 = zero, y1 = zero, y2 = zero, z1 = zero, z2 = zero, o = zero, c1 = zero, c2 = zero, c3 = zero;
  if (f.kind == 0) {
    hsplit2(f.x, x1, x2); hsplit2(f.y, y1, y2); hsplit2(f.z, z1, z2);
    hsplit3(-(cut2s - f.w), c1, c2, c3);
    o = one;
  } else {
    c1 = __float2half_rn(f.kind == 1 ? BIG : -BIG);      // times the 1 in column 12 of every B row
  }
  Half8 a0, a1;
  a0.a = __halves2half2(x1, x1); a0.b = __halves2half2(x2, y1); a0.c = __halves2half2(y1, y2); a0.d = __halves2half2(z1, z1);
  a1.a = __halves2half2(z2, o); a1.b = __halves2half2(o, o); a1.c = __halves2half2(c1, c2); a1.d = __halves2half2(c3, zero);
  *reinterpret_cast<Half8*>(sA + chunk_off16(row, 0)) = a0;
  *reinterpret_cast<Half8*>(sA + chunk_off16(row, 1)) = a1;
}

#define PSIM_TMEM_LD32_PACK16(r, taddr)                                                                            \
  asm volatile(                                                                                                    \
      "tcgen05.ld.sync.aligned.32x32b.x32.pack::16b.b32 "                                                          \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                    \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"                  \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),            \
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),      \
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),    \
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])     \
      : "r"(taddr))

constexpr unsigned SIGNS = 0x80008000u;

// -DPSIM_TC_PROF: clock64 timeline of block 0 (tiles 0..511), printed at the end of the kernel.
// events: 0..3 per MMA warp: [4 w + 0] b_full/d_empty seen, [4 w + 1] issued + committed;
//         16 + 5 * consumer warp + {0: d_full seen, 1: load complete, 2: accumulator released, 3: slot bookkeeping done,
//                                   4: reduction done}
#ifdef PSIM_TC_PROF
__device__ long long g_trace[16 + 5 * CW][512];
#ifndef PSIM_TC_PROF_BLOCK
#define PSIM_TC_PROF_BLOCK 1000
#endif
__device__ long long g_flush[CW][10];   // per consumer warp of block PSIM_TC_PROF_BLOCK: start; per call: consume returned, flush done, entries
#ifdef PSIM_TC_PROF_LIGHT                  // only the per-block stamps (g_flush): the tile loop runs undisturbed
#define TR(ev, tile) {}
#else
#define TR(ev, tile) { if (blockIdx.x == 0 && (tile) < 512) g_trace[ev][tile] = clock64(); }
#endif
#else
#define TR(ev, tile) {}
#endif

// 64 columns of this thread's row as 32 packed registers (low half = even column): OR them (16 LOP3,
// groups of 9, 9, 9 and 5 registers); a sign bit in the result = some entry is negative = a candidate.
// The rare hit pushes the negative columns (tile-relative code `tcode`); more than QMAX_TILE of them become
// ONE dense entry, so a chunk adds at most QMAX_TILE logical entries — the first pass may write up to 64
// entries past cnt, hence QPHYS.
__device__ __forceinline__ void queue_chunk(const unsigned (&r)[32], unsigned tcode, unsigned q, int& cnt) {
  // grouped tree (9, 9, 9, 5 registers) and an unrolled descent into the groups
  unsigned mg[4];
#pragma unroll
  for (int g = 0; g < 3; ++g)
    mg[g] = (r[9 * g] | r[9 * g + 1] | r[9 * g + 2]) | (r[9 * g + 3] | r[9 * g + 4] | r[9 * g + 5]) |
            (r[9 * g + 6] | r[9 * g + 7] | r[9 * g + 8]);
  mg[3] = (r[27] | r[28] | r[29]) | r[30] | r[31];
  const unsigned m = (mg[0] | mg[1] | mg[2]) | mg[3];
  if (m & SIGNS) {
    const int c0 = cnt;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (mg[g] & SIGNS) {
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          const int u = 9 * g + k;                 // static after unrolling: r[] must stay in registers
          if (u < 32) {
            const unsigned v = r[u < 32 ? u : 0];
            if (v & 0x00008000u) { q_store(q, cnt, tcode + 2 * u); ++cnt; }
            if (v & 0x80000000u) { q_store(q, cnt, tcode + 2 * u + 1); ++cnt; }
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


// What the exact path needs of Consts, and the exact path's state: ONE 128-byte line of local memory per thread.
// (The consumer loop is a function of its own with 64 accumulator registers live, so its caller can keep next to
// nothing in registers across the call; behind a pointer to the kernel's parameter copy the exact path started
// every call with a chain of three local-memory misses, ~4,000 cycles during which the tensor pipeline stood.)
template <class R>
struct PairConsts { R r_cut, cut2_pad, huge, r_min, F_rcut, epsilon; };
template <class R>
struct __align__(128) FlushCtx {
  const Real4<R>* pos; int64_t ig;
  int64_t col0;              // global index of column 0 of this thread's FIRST own tile in the block's j segment
  R f[3]; unsigned in_range;
  PairConsts<R> c;
};

// k_begin: the own-tile index that queue codes are relative to; own tile k is tile 4 k + group of the segment,
// chunk c since k_begin = columns 64 (c & 1) ... of own tile k_begin + (c >> 1).
// Entries are evaluated U at a time and added in queue order (fixed summation order).  U = 1 measured best on
// B200 (f64: 3.01 ms per step at N = 262,144 against 3.11 with U = 2 and 3.17 with U = 3): the exact path is a
// chain of ~45 dependent FP64 operations per candidate (~1,000 cycles), and while it runs the warp's accumulator
// and, through the ring, the tensor pipeline wait: ~6 % of a block's time (tools/tc_timeline.py, -DPSIM_TC_PROF_LIGHT).
template <class R>
__device__ __noinline__ void flush_queue(FlushCtx<R>* x, unsigned q, int cnt, int k_begin) {
  if (cnt == 0) return;
  constexpr int U = PSIM_TC_FLUSH_U;
  const Real4<R>* __restrict__ pos = x->pos;
  const PairConsts<R> c = x->c;
  const int64_t ig = x->ig, col0 = x->col0;
  const Real4<R> pi = pos[ig];
  auto first_j = [&](unsigned v) {
    const int ch = (int)((v >> 6) & 511u);
    return (int64_t)(k_begin + (ch >> 1)) * (NBUF * N2) + col0 + (int64_t)((ch & 1) * 64) + (int64_t)((v & DENSE) ? 0u : (v & 63u));
  };
  R f0 = x->f[0], f1 = x->f[1], f2 = x->f[2];
  unsigned in_range = x->in_range;
  for (int e = 0; e < cnt; e += U) {
    unsigned v[U]; int64_t jg[U]; Real4<R> pj[U];
    unsigned dense = 0u;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const bool ok = e + u < cnt;
      v[u] = ok ? q_load(q, e + u) : 0u;
      jg[u] = ok ? first_j(v[u]) : ig;
      dense |= v[u] & DENSE;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) pj[u] = pos[jg[u]];
    if (dense) {                                   // rare: a chunk with more than QMAX_TILE candidates in one row
      R fd[3] = {f0, f1, f2};
#pragma unroll                                     // static indices: v, jg, pj must stay in registers
      for (int u = 0; u < U; ++u) {
        if (e + u >= cnt) continue;
        if (v[u] & DENSE) {                        // all 64 columns of the chunk
          Real4<R> pc = pj[u];
#pragma unroll 1
          for (int w = 0; w < 64; ++w) {
            Real4<R> pn = pc;
            if (w + 1 < 64) pn = pos[jg[u] + w + 1];
            R ax, ay, az;
            in_range += pair_term(c, ig, pi, jg[u] + w, pc, ax, ay, az) ? 1u : 0u;
            fd[0] = r_add(fd[0], ax); fd[1] = r_add(fd[1], ay); fd[2] = r_add(fd[2], az);
            pc = pn;
          }
        } else {
          R ax, ay, az;
          in_range += pair_term(c, ig, pi, jg[u], pj[u], ax, ay, az) ? 1u : 0u;
          fd[0] = r_add(fd[0], ax); fd[1] = r_add(fd[1], ay); fd[2] = r_add(fd[2], az);
        }
      }
      f0 = fd[0]; f1 = fd[1]; f2 = fd[2];
    } else {
      R ax[U], ay[U], az[U];
#pragma unroll
      for (int u = 0; u < U; ++u) in_range += pair_term(c, ig, pi, jg[u], pj[u], ax[u], ay[u], az[u]) ? 1u : 0u;
#pragma unroll
      for (int u = 0; u < U; ++u) { f0 = r_add(f0, ax[u]); f1 = r_add(f1, ay[u]); f2 = r_add(f2, az[u]); }
    }
  }
  x->f[0] = f0; x->f[1] = f1; x->f[2] = f2; x->in_range = in_range;
}

// The consumer tile loop of one warp (a function of its own, so that the exact path's double-precision
// register needs stay out of it; loop state in uniform registers).
// The warp's group g owns accumulator g: own tile k is tile 4 k + g, phase k & 1 of the accumulator's barriers.
// Per tile: wait for the accumulator, load the warp's 32 rows x 128 columns, HAND THE ACCUMULATOR BACK, then
// reduce from registers.  What bounds this kernel is the round trip of an accumulator (consumers' release ->
// MMA warp wakes -> issue -> tensor pipe -> commit -> consumers wake -> load; tools/microbench/tc_rtt.cu):
// 512 TMEM columns per round trip is all the tensor pipe gets, so the release comes as early as the data
// allows and the reduction of tile k overlaps the MMA side of the next round instead of delaying it.
// lane_base addresses the warp's rows of accumulator g.  dfull / dempty: shared addresses of the accumulator's
// two barriers.  releases_slots: this warp hands the ring slot of its tiles back to the TMA producer: lane 0
// arrives on b_empty[(t / TPS) % NSLOT] once per own tile, together with the accumulator's release (the MMA
// that read the B tile is complete then); bempty0 = shared address of b_empty[0].
// ntk (own tiles): multiple of 2.  Returns (next own tile << 32) | queue length; queue codes are relative to
// k_begin.  Returns at the end, every SEG own tiles, and when some lane's queue could overflow in the next pair.
__device__ __noinline__ unsigned long long consume_tiles(unsigned q, unsigned lane_base, unsigned dfull, unsigned dempty,
                                                           unsigned bempty0, int releases_slots, int ntk_v, int k_begin_v) {
  const int lane = threadIdx.x & 31;
  const int k_begin = (int)__reduce_max_sync(0xffffffffu, (unsigned)k_begin_v);
  const int ntk = (int)__reduce_max_sync(0xffffffffu, (unsigned)ntk_v);
  const bool release_slot = __reduce_max_sync(0xffffffffu, (unsigned)releases_slots) != 0u;   // uniform
  // own tile k = tile 4 k + g: its ring slot is ((4 k + g) / TPS) % NSLOT = (k / 2) % NSLOT for TPS = 8
#define PSIM_TC_TRT(kk) (4 * (kk) + (int)(threadIdx.x >> 7))
#define PSIM_TC_TILE(par_, kk, code)                                                           \
  {                                                                                              \
    mbar_wait_warp(dfull, (par_));                  /* all lanes leave together: .aligned load next */ \
    if (lane == 0) TR(16 + 5 * (int)(threadIdx.x >> 5), PSIM_TC_TRT(kk))                         \
    asm volatile("tcgen05.fence::after_thread_sync;");                                           \
    PSIM_TMEM_LD32_PACK16(ra, lane_base);                                                        \
    PSIM_TMEM_LD32_PACK16(rb, lane_base + 64u);                                                  \
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");                                 \
    if (lane == 0) TR(17 + 5 * (int)(threadIdx.x >> 5), PSIM_TC_TRT(kk))                         \
    asm volatile("tcgen05.fence::before_thread_sync;");   /* values are in registers */          \
    __syncwarp();                                                                                \
    /* A real branch: ptxas schedules within basic blocks; as predicated code the release ended up */ \
    /* behind half of the reduction, and the reduction's first half between the two loads. */     \
    if (lane == 0) {                                                                             \
      mbar_arrive(dempty);                                                                       \
      TR(18 + 5 * (int)(threadIdx.x >> 5), PSIM_TC_TRT(kk))                                      \
      if (release_slot) mbar_arrive(bempty0 + 8u * (unsigned)(((kk) / (TPS / NBUF)) % NSLOT));   \
      TR(19 + 5 * (int)(threadIdx.x >> 5), PSIM_TC_TRT(kk))                                      \
    }                                                                                            \
    __syncwarp();                                                                                \
    queue_chunk(ra, (code), q, cnt);                                                             \
    queue_chunk(rb, (code) + 64u, q, cnt);                                                       \
    if (lane == 0) TR(20 + 5 * (int)(threadIdx.x >> 5), PSIM_TC_TRT(kk))                         \
  }
  unsigned ra[32], rb[32];
  int cnt = 0, k0 = k_begin;
  for (;;) {
    const unsigned tc0 = (unsigned)(k0 - k_begin) << 7;
    PSIM_TC_TILE(0u, k0, tc0)
    PSIM_TC_TILE(1u, k0 + 1, tc0 + 128u)
    k0 += 2;
    // the next pair of tiles adds at most 4 QMAX_TILE logical entries
    if (k0 >= ntk || k0 - k_begin >= SEG || __any_sync(0xffffffffu, cnt > QC - 4 * QMAX_TILE)) break;
  }
#undef PSIM_TC_TILE
#undef PSIM_TC_TRT
  return ((unsigned long long)(unsigned)k0 << 32) | (unsigned)cnt;
}

template <class R, int MODE>
__global__ void __launch_bounds__(TW, 1)
allpairs_tc_kernel(const StepArgs<R> a, const unsigned char* __restrict__ bops, unsigned long long* __restrict__ pair_count,
                     const JSplit<R> js, const Scale16 sc) {
  extern __shared__ __align__(1024) unsigned char sB[];          // ring of B tiles, then the candidate queues
  __shared__ __align__(128) unsigned char sA[M * K * 2];          // 4 KB
  __shared__ __align__(8) unsigned long long b_full[NSLOT], b_empty[NSLOT], d_bar[2 * NBUF];   // d_bar: full[0..3], empty[0..3]
  unsigned long long* const d_full = d_bar;
  unsigned long long* const d_empty = d_bar + NBUF;
  __shared__ unsigned tmem_base_s;
  __shared__ int last_s;
  __shared__ R fpart[4][3][M];

  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool consumer = warp < CW;
  const int row = (warp & 3) * 32 + lane;          // accumulator row = TMEM lane (consumers)
  const int grp = (warp >> 2) & 3;                 // consumers: the accumulator this warp's group serves
  const int S = js.S;
  const int ib = (int)blockIdx.x / S, seg = (int)blockIdx.x - ib * S;
  const int64_t ig = a.i_lo + (int64_t)ib * M + row;
  const float rcut = (float)c.r_cut;
  const float cut2s = (float)c.cut2_pad * (1.0f + 1.0e-6f) * sc.S2;
  // this block's j segment, in tiles (multiples of 8: whole ring slots, whole phase pairs)
  const int nt8 = (int)(a.n_pad / (8 * N2));
  const int t_lo = 8 * ((seg * nt8) / S), t_hi = 8 * (((seg + 1) * nt8) / S);
  const int nt = t_hi - t_lo;

  // --- A tile: this block's i-particles ---
  if (consumer && grp == 0) {
    Rec16 f; f.x = f.y = f.z = f.w = 0.f; f.kind = 1;
    if (ig < a.i_hi) f = filter_record_tc(a.pos_cur[ig], c, rcut, sc);
    store_a_row(sA, row, f, cut2s);
  }
  if (tid == 0) {
    for (int s = 0; s < NSLOT; ++s) { mbar_init(smem_u32(&b_full[s]), 1); mbar_init(smem_u32(&b_empty[s]), TPS); }
    for (int b = 0; b < NBUF; ++b) { mbar_init(smem_u32(&d_full[b]), 1); mbar_init(smem_u32(&d_empty[b]), CW / NBUF); }
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

  FlushCtx<R> fx;
  fx.pos = a.pos_cur; fx.ig = ig; fx.col0 = (int64_t)(t_lo + grp) * N2;
  fx.f[0] = fx.f[1] = fx.f[2] = (R)0; fx.in_range = 0;
  fx.c.r_cut = c.r_cut; fx.c.cut2_pad = c.cut2_pad; fx.c.huge = c.huge; fx.c.r_min = c.r_min; fx.c.F_rcut = c.F_rcut;
  fx.c.epsilon = c.epsilon;
  if (warp == CW + NCTL) {
    if (lane == 0) {
      // ===== TMA producer: one bulk copy per ring slot (this GPU's memory), as far ahead as the ring allows =====
      const int nq = nt / TPS;
      const unsigned bfull0 = smem_u32(&b_full[0]), bempty0 = smem_u32(&b_empty[0]), sB0 = smem_u32(sB);
      const unsigned char* src = bops + (size_t)t_lo * TILE_BYTES;
      for (int Q = 0; Q < nq; ++Q) {
        const int s = Q % NSLOT;
        mbar_wait(bempty0 + 8u * s, (unsigned)(((Q / NSLOT) & 1) ^ 1));   // all tiles of the slot multiplied (fresh: passes)
        mbar_expect_tx(bfull0 + 8u * s, SLOT_BYTES);
        bulk_copy_g2s(sB0 + (unsigned)(s * SLOT_BYTES), src + (size_t)Q * SLOT_BYTES, SLOT_BYTES, bfull0 + 8u * s);
      }
    }
  } else if (warp >= CW) {
    if (lane == 0) {
      // ===== MMA warp w: issues the tiles t = w, w + 4, ... into accumulator w: one instruction, one commit =====
      const int w = warp - CW;
      const int ntw = nt / NCTL;                   // nt is a multiple of 8
      // D = F16 (c_format 0), A and B = F16 (0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
      const unsigned idesc = ((unsigned)(N2 >> 3) << 17) | ((unsigned)(M >> 4) << 24);
      const unsigned long long da = make_desc(smem_u32(sA));
      const unsigned acc = tm + (unsigned)(w * N2);
      const unsigned dfull = smem_u32(&d_full[w]), dempty = smem_u32(&d_empty[w]);
      const unsigned bfull0 = smem_u32(&b_full[0]), sB0 = smem_u32(sB);
      for (int k = 0; k < ntw; ++k) {
        constexpr int KPS = TPS / NCTL;            // own tiles per fill
        const int Q = k / KPS, s = Q % NSLOT;      // tile t = w + 4 k sits in ring fill Q = t / TPS, at position w + 4 (k % KPS)
        if (k % KPS == 0) mbar_wait(bfull0 + 8u * s, (unsigned)((Q / NSLOT) & 1));      // the slot's tiles landed
        TR(4 * w + 2, w + 4 * k)
        mbar_wait(dempty, (unsigned)((k & 1) ^ 1));                                     // accumulator drained
        TR(4 * w + 0, w + 4 * k)
        asm volatile("tcgen05.fence::after_thread_sync;");
        const unsigned long long db = make_desc(sB0 + (unsigned)(s * SLOT_BYTES + (w + 4 * (k % KPS)) * TILE_BYTES));
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}\n"
            :: "r"(acc), "l"(da), "l"(db), "r"(idesc), "r"(0u), "r"(0u), "r"(0u), "r"(0u), "r"(0u));
        umma_commit(dfull);
        TR(4 * w + 1, w + 4 * k)
      }
    }
  } else {
    // ===== consumers =====
    const unsigned q = smem_u32(sB + RING_BYTES) + 2u * (unsigned)tid;
    const unsigned lane_base = tm + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)(grp * N2);
    const unsigned dfull = smem_u32(&d_full[grp]), dempty = smem_u32(&d_empty[grp]);
    const unsigned bempty0 = smem_u32(&b_empty[0]);
    const int releases = (warp & 3) == 0 ? 1 : 0;     // one warp per group hands the ring slots back (TPS arrivals per slot)
    const int ntk = nt / NBUF;
    int k = 0;
#ifdef PSIM_TC_PROF
    int call = 0;
    const bool prof_block = blockIdx.x == PSIM_TC_PROF_BLOCK && lane == 0;
    if (prof_block) g_flush[warp][0] = clock64();
#endif
    while (k < ntk) {
      const unsigned long long r = consume_tiles(q, lane_base, dfull, dempty, bempty0, releases, ntk, k);
#ifdef PSIM_TC_PROF
      if (prof_block && call < 3) { g_flush[warp][1 + 3 * call] = clock64(); g_flush[warp][3 + 3 * call] = (long long)(unsigned)r; }
#endif
      flush_queue<R>(&fx, q, (int)(unsigned)r, k);
#ifdef PSIM_TC_PROF
      if (prof_block && call < 3) g_flush[warp][2 + 3 * call] = clock64();
      ++call;
#endif
      k = (int)(r >> 32);
    }
    unsigned in_range = fx.in_range;
    for (int off = 16; off > 0; off >>= 1) in_range += __shfl_xor_sync(0xffffffffu, in_range, off);
    if (lane == 0 && in_range) atomicAdd(pair_count, (unsigned long long)in_range);
  }

  // --- the four threads of a row meet (fixed order) ---
  asm volatile("tcgen05.fence::before_thread_sync;");
  R* const f = fx.f;
  if (consumer && grp > 0) { fpart[grp][0][row] = f[0]; fpart[grp][1][row] = f[1]; fpart[grp][2][row] = f[2]; }
  __syncthreads();
  const bool owner = consumer && grp == 0;
  if (owner) {
#pragma unroll
    for (int g = 1; g < 4; ++g) {
      f[0] = r_add(f[0], fpart[g][0][row]); f[1] = r_add(f[1], fpart[g][1][row]); f[2] = r_add(f[2], fpart[g][2][row]);
    }
  }
  if (warp == CW) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
  }
  // --- the j segments of an i-block meet in global scratch (psim_allpairs_tc.cuh) ---
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
      if (last_s) js.arrived[ib] = 0u;
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
#ifdef PSIM_TC_PROF
  __syncthreads();
  if (blockIdx.x == PSIM_TC_PROF_BLOCK && tid == 0) {
    const long long z = g_flush[0][0], e = clock64();
    for (int w = 0; w < CW; ++w)
      printf("FL w%d: start %lld | returned %lld flushed %lld entries %lld | returned %lld flushed %lld entries %lld | block end %lld\n", w,
             g_flush[w][0] - z, g_flush[w][1] - z, g_flush[w][2] - z, g_flush[w][3], g_flush[w][4] - z, g_flush[w][5] - z, g_flush[w][6], e - z);
  }
  if (blockIdx.x == 0 && tid == 0) {
    const long long z = g_trace[2][256];
    for (int t = 256; t < 268; ++t) {
      const int w = t & 3;
      printf("TR t=%d mma: b_full %lld d_empty %lld committed %lld |", t, g_trace[4 * w + 2][t] - z, g_trace[4 * w][t] - z, g_trace[4 * w + 1][t] - z);
      for (int cw = 4 * w; cw < 4 * w + 4; ++cw)
        printf(" w%d: %lld %lld %lld %lld %lld", cw, g_trace[16 + 5 * cw][t] - z, g_trace[17 + 5 * cw][t] - z, g_trace[18 + 5 * cw][t] - z,
               g_trace[19 + 5 * cw][t] - z, g_trace[20 + 5 * cw][t] - z);
      printf("\n");
    }
  }
#endif
  R first = (R)0, second = (R)0;
  if (finish && owner && ig < a.i_hi) step_epilogue<R, MODE>(a, ig, f[0], f[1], f[2], first, second);
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, TW>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

}  // namespace tc
}  // namespace psim
