// psim_allpairs_tc3.cuh — EXPERIMENTAL (PSIM_OPT_EXACT_DIFF_FILTER = 6): the tensor-core all-pairs pass of
// psim_allpairs_tc2.cuh with a software-pipelined consumer.  NOT the default and not exercised by the
// default tests or bench.py:
//
//   status (end of round 1): 5.4 ms/step at N = 262,144 f64 against 6.8 ms for psim_allpairs_tc2.cuh, parity
//   suite green (69 tests), filter_pairs == exact pair count at full size — but ONE "unspecified launch
//   failure" in ~16 bench processes (in the e2e loop, force-only launches) and a timed-out 8-GPU
//   N = 1,048,576 run.  Cause not found; the round's GPU budget ran out.  What is known:
//     * psim_allpairs_tc2.cuh never failed in the same runs (full suites, 1/2/4/8-GPU benches);
//     * a visible difference in SASS: tc2 has BSYNC + WARPSYNC.ALL between every mbarrier.try_wait spin loop
//       and the .sync.aligned tcgen05.ld; here ptxas keeps the loop state in uniform registers, treats the
//       try_wait result as warp-uniform and emits neither — an explicit __syncwarp() after the wait is
//       compiled to NOPs; the wait now loops on __all_sync (mbar_wait_warp), unverified.  If lanes CAN leave the spin an iteration apart, the .aligned load runs with a
//       partial warp; if the hardware guarantees a uniform result, this is not the cause.
//     * next steps: identify the failing launch with error checks after every launch in a loop of a few
//       hundred force calls; bisect dense entries / segment re-entry / the REDUX'ed start tile / the 3-slot
//       ring; compute-sanitizer is closed on this pool.
//
// Differences from psim_allpairs_tc2.cuh:
//   - the consumer tile loop is a function of its own (consume_tiles): ptxas gives it its own register
//     allocation (a few scalars + two sets of 32 accumulator values, no spills), so the tcgen05.ld of tile t + 1
//     is in flight while tile t is reduced; its loop state lives in uniform registers;
//   - 16-bit tile-relative candidate queues (96 logical entries per thread), at most 5 entries per tile
//     (more hits in a 32-column chunk become one "dense" entry that the exact path expands), flushed between
//     calls of consume_tiles, never while accumulators are held;
//   - a 3-slot TMA ring.
#pragma once

#include "psim_allpairs_tc.cuh"
#include "psim_allpairs_tc2.cuh"

namespace psim {
namespace tcp {

using namespace tc;                             // the mathematics, descriptors, mbar_wait, PSIM_TMEM_LD32: psim_allpairs_tc.cuh

constexpr int N2 = 128;                         // j records per tile of this kernel = accumulator columns per buffer
constexpr int NBUF = 512 / N2;                  // TMEM ring: four accumulators of 128 columns
constexpr int NCTL = NBUF;                      // MMA warps: warp w owns accumulator w, tiles t = w (mod NCTL)
constexpr int NPAIR = 3;                        // shared-memory ring: 3 slots of two tiles (one 16 KB bulk copy each; 2: slower, 4: same)
constexpr int STAGES = 2 * NPAIR;               // 6 x 8 KB
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

// tc_prepare_kernel (same 128-record B tiles), mbar_init / mbar_arrive / mbar_expect_tx / bulk_copy_g2s /
// umma_commit: psim_allpairs_tc2.cuh.

// Wait for an mbarrier phase with the whole warp and leave the spin loop in the SAME iteration in all lanes
// (the vote is a real warp-synchronous instruction; a bare __syncwarp() after mbar_wait is elided by ptxas here).
__device__ __forceinline__ void mbar_wait_warp(unsigned addr, unsigned parity) {
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  } while (!__all_sync(0xffffffffu, done));
}

// The exact path: every queued candidate j of row ig in the working precision, with the reference's own
// arithmetic (pair_force, psim_common.cuh).  A dense entry expands to its 32 columns.
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
  const Real4<R>* pos; const Consts<R>* c; int64_t ig; int col0;   // col0: first column of this thread's group
  R f[3]; unsigned in_range;                                        // force on row ig; ordered in-range pairs evaluated
};

// seg_begin: the tile that queue codes are relative to.
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
// divergence barriers and spills, 1.3x slower).  nt comes straight from the kernel parameters; the start
// tile comes back from the previous call, so it is passed through a warp reduction (REDUX writes a uniform
// register) — it is the same in all lanes anyway.
// Returns at a group boundary with no load in flight: at the end, every SEG tiles (a queue code holds a
// 10-bit tile offset), and when some lane's queue could overflow during the next two groups (one is already
// committed when the decision is taken).  The caller flushes the queue and calls again.  A flush in
// mid-stream stalls the accumulator ring, so the capacity is sized to make the last kind rare.  The flush
// is NOT called from here: values live across a call must sit above the callee chain's registers, and with
// pair_force<double> below that left too few for this loop.
// Returns (next tile << 32) | queue length.  Queue codes are relative to t_begin.
// dbar0: shared address of d_full[0]; d_empty[b] lives 8 * NBUF bytes after d_full[b].  nt: multiple of 8.
__device__ __noinline__ unsigned long long consume_tiles(float cut, unsigned q, unsigned lane_base, unsigned dbar0, int nt, int t_begin_v) {
  const int lane = threadIdx.x & 31;
  const int t_begin = (int)__reduce_max_sync(0xffffffffu, (unsigned)t_begin_v);
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

template <class R, int MODE>
__global__ void __launch_bounds__(TW, 1)
allpairs_tcws_kernel(const StepArgs<R> a, const unsigned char* __restrict__ bops, unsigned long long* __restrict__ pair_count) {
  extern __shared__ __align__(1024) unsigned char sB[];          // ring of B tiles, then the candidate queues
  __shared__ __align__(128) unsigned char sA[M * K * 4];          // 8 KB
  __shared__ __align__(8) unsigned long long b_full[NPAIR], b_empty[NPAIR], d_bar[2 * NBUF];   // d_bar: full[0..3], empty[0..3]
  unsigned long long* const d_full = d_bar;
  unsigned long long* const d_empty = d_bar + NBUF;
  __shared__ unsigned tmem_base_s;
  __shared__ R fpart[NQ][3][M];

  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool consumer = warp < CW;
  const int row = (warp & 3) * 32 + lane;          // accumulator row = TMEM lane (consumers)
  const int half = (warp >> 2) % NQ;               // which 32 columns of every tile (consumers)
  const int64_t ig = a.i_lo + (int64_t)blockIdx.x * M + row;
  const float rcut = (float)c.r_cut;
  const float cut2f = (float)c.cut2_pad * (1.0f + 1.0e-6f);
  const int nt = (int)(a.n_pad / N2);

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
  fx.pos = a.pos_cur; fx.c = &c; fx.ig = ig; fx.col0 = half * 32;
  fx.f[0] = fx.f[1] = fx.f[2] = (R)0; fx.in_range = 0;
  if (warp == CW + NCTL) {
    if (lane == 0) {
      // ===== TMA producer: one 16 KB bulk copy per pair of tiles, as far ahead as the ring allows =====
      const int np = nt / 2;
      const unsigned bfull0 = smem_u32(&b_full[0]), bempty0 = smem_u32(&b_empty[0]), sB0 = smem_u32(sB);
      for (int P = 0; P < np; ++P) {
        const int s = P % NPAIR;
        mbar_wait(bempty0 + 8u * s, (unsigned)(((P / NPAIR) & 1) ^ 1));   // both MMA warps done with the slot (fresh: passes)
        mbar_expect_tx(bfull0 + 8u * s, 2 * TILE_BYTES);
        bulk_copy_g2s(sB0 + (unsigned)(s * 2 * TILE_BYTES), bops + (size_t)P * 2 * TILE_BYTES, 2 * TILE_BYTES, bfull0 + 8u * s);
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

  // --- the column groups of a row meet (fixed order); epilogue by the thread of group 0 ---
  asm volatile("tcgen05.fence::before_thread_sync;");
  R* const f = fx.f;
  if (consumer && half > 0) { fpart[half][0][row] = f[0]; fpart[half][1][row] = f[1]; fpart[half][2][row] = f[2]; }
  __syncthreads();
  R first = (R)0, second = (R)0;
  if (consumer && half == 0 && ig < a.i_hi) {
#pragma unroll
    for (int g = 1; g < NQ; ++g) {
      f[0] = r_add(f[0], fpart[g][0][row]); f[1] = r_add(f[1], fpart[g][1][row]); f[2] = r_add(f[2], fpart[g][2][row]);
    }
    step_epilogue<R, MODE>(a, ig, f[0], f[1], f[2], first, second);
  }
  if (warp == CW) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
  }
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, TW>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

}  // namespace tcp
}  // namespace psim
