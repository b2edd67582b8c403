// psim_allpairs_tc2.cuh — the default all-pairs pass: distance filter on the tensor cores,
// warp-specialised (psim_allpairs_tc.cuh has the mathematics and the first, lock-step version;
// psim_allpairs_tc3.cuh an experimental successor with software-pipelined consumers).
//
// What the first version lost (ncu, profiles/): eight warps marched through
// stage -> barrier -> wait -> reduce together (IPC 0.28, tensor pipe 16 %), and
// each of the 2,048 blocks re-derived the same B operand rows from the positions.
// Here
//   * tc_prepare_kernel builds the B operand ONCE per force pass, in global
//     memory, already in the UMMA shared-memory tile layout (8 KB per tile of
//     128 j records);
//   * inside the force kernel a producer warp streams those tiles into a
//     shared-memory ring with TMA bulk copies (cp.async.bulk, completion on an
//     mbarrier, 16 KB = two tiles per copy) and four MMA warps (one elected lane
//     each, one TMEM accumulator each) issue the tcgen05.mma pairs — one thread
//     doing all of it sustains only ~1 tile per 950 cycles (measured,
//     tools/microbench/tc_gram.cu);
//   * sixteen consumer warps do nothing but tcgen05.ld + min-reduction and queue
//     the rare candidates for the exact path; rings of mbarriers (B full/empty,
//     TMEM full/empty) replace every block-wide barrier of the loop.
// PSIM_TC_PROF builds add clock64 instrumentation and an event timeline (block 0).
#pragma once

#include "psim_allpairs_tc.cuh"

namespace psim {
namespace tc {

constexpr int N2 = 128;                         // j records per tile of this kernel = accumulator columns per buffer
constexpr int NBUF = 512 / N2;                  // TMEM ring: four accumulators of 128 columns
constexpr int NCTL = NBUF;                      // MMA warps: warp w owns accumulator w, tiles t = w (mod NCTL)
constexpr int NPAIR = 4;                        // shared-memory ring: 4 slots of two tiles (one 16 KB bulk copy each)
constexpr int STAGES = 2 * NPAIR;               // 8 x 8 KB
constexpr int TILE_BYTES = N2 * K * 4;          // 8 KB
constexpr int NQ = 4;                           // column groups: warp w reads TMEM lanes 32 (w % 4).., columns 32 (w / 4)..
constexpr int CW = 4 * NQ;                      // consumer warps
constexpr int TW = 32 * (CW + NCTL + 1);        // + the MMA warps + the TMA producer warp
constexpr int QC = 48;                          // candidate queue entries per consumer thread (96 KB in all)
constexpr int RING_BYTES = STAGES * TILE_BYTES;
constexpr int SMEM_BYTES = RING_BYTES + QC * 32 * CW * 4;

// B operand rows of all particles, tile by tile, in the canonical layout of chunk_off().
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

__device__ __forceinline__ void mbar_init(unsigned addr, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(addr), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(addr) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned addr, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(addr), "r"(bytes) : "memory");
}
// TMA bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_copy_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned mbar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(mbar) : "memory");
}

// One 32-column chunk: running minima; candidates are only QUEUED here (their exact evaluation
// would stall this warp for hundreds of cycles, and every accumulator buffer waits for all warps).
// A shared-memory address the compiler may not rematerialise: left alone it recomputes the address of a
// __shared__ object before every use (S2UR SR_CgaCtaId -> ULEA -> LEA, ~100 cycles of uniform-datapath latency).
__device__ __forceinline__ unsigned pinned(unsigned v) {
  unsigned r;
  asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}

// Spin on the non-blocking test_wait (try_wait suspends the thread in hardware; measured below).
__device__ __forceinline__ void mbar_spin(unsigned addr, unsigned parity) {
  unsigned done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  }
}
#ifdef PSIM_TC_SPIN
#define PSIM_MBAR_WAIT mbar_spin
#else
#define PSIM_MBAR_WAIT mbar_wait
#endif

template <class R>
__device__ __noinline__ void flush_queue(const Real4<R>* __restrict__ pos, const Consts<R>& c, int64_t ig, const int* q, int cnt, R* f,
                                         unsigned* in_range) {
  if (cnt == 0) return;                            // rows past i_hi / inactive rows never queue anything
  const Real4<R> pi = pos[ig];
  for (int e = 0; e < cnt; ++e) {
    const int64_t jg = (int64_t)q[e * (32 * CW)];
    if (jg == ig) continue;
    const Real4<R> pj = pos[jg];
    if (pj.w >= c.huge) continue;
    const R dx = r_sub(pi.x, pj.x), dy = r_sub(pi.y, pj.y), dz = r_sub(pi.z, pj.z);
    const R r2 = r_add(r_add(r_mul(dx, dx), r_mul(dy, dy)), r_mul(dz, dz));
    if (r2 < c.cut2_pad) {
      pair_force(pi.x, pi.y, pi.z, pj.x, pj.y, pj.z, c, f[0], f[1], f[2]);
      if (r_sqrt(r2) < c.r_cut) ++*in_range;       // the predicate of count_pairs_kernel (psim_allpairs.cuh)
    }
  }
}

__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(a, fminf(b, c)); }   // one FMNMX3

// 32 columns of this thread's row: 16 three-input minima (groups of 9, 9, 9 and 5 columns), one compare;
// the rare hit descends to the groups and pushes the candidate columns.
template <class R>
__device__ __forceinline__ void queue_chunk(const unsigned (&r)[32], float cut, int j0, int* q, int& cnt) {
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
  if (m < cut) {                                   // the caller guarantees 32 free queue entries
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (mg[g] < cut) {
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          const int u = 9 * g + k;                 // static after unrolling: r[] must stay in registers
          if (u < 32 && __uint_as_float(r[u < 32 ? u : 0]) < cut) {
            q[cnt * (32 * CW)] = j0 + u;
            ++cnt;
          }
        }
      }
    }
  }
}

#ifdef PSIM_TC_PROF
__device__ long long g_trace[40][512];
#define TR(e, t) { if (blockIdx.x == 0 && (t) < 512) g_trace[e][t] = clock64(); }
#else
#define TR(e, t) {}
#endif

template <class R, int MODE>
__global__ void __launch_bounds__(TW, 1)
allpairs_tcws_kernel(const StepArgs<R> a, const unsigned char* __restrict__ bops, unsigned long long* __restrict__ pair_count) {
  extern __shared__ __align__(1024) unsigned char sB[];          // STAGES x 16 KB ring (+ padding, see launch)
  __shared__ __align__(128) unsigned char sA[M * K * 4];          // 8 KB
  __shared__ __align__(8) unsigned long long b_full[NPAIR], b_empty[NPAIR], d_full[NBUF], d_empty[NBUF];
  __shared__ unsigned tmem_base_s;
  __shared__ R fpart[NQ][3][M];

  const Consts<R>& c = a.c;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool consumer = warp < CW;
  const int row = (warp & 3) * 32 + lane;          // accumulator row = TMEM lane (consumers)
  const int half = (warp >> 2) % NQ;               // which N / NQ columns of every tile (consumers)
  const int64_t ig = a.i_lo + (int64_t)blockIdx.x * M + row;
  const float rcut = (float)c.r_cut;
  const float cut2f = (float)c.cut2_pad * (1.0f + 1.0e-6f);
  const int nt = (int)(a.n_pad / N2);

  float cut = -1.0e30f;
  if (consumer) {
    FilterRec f; f.x = f.y = f.z = 0.f; f.w = 1.0e30f;
    if (ig < a.i_hi) {
      const Real4<R> p = a.pos_cur[ig];
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

  R f[3] = {(R)0, (R)0, (R)0};
  unsigned in_range = 0;                           // ordered in-range pairs this thread's exact path evaluated
  if (warp == CW + NCTL) {
    if (lane == 0) {
      // ===== TMA producer: one 16 KB bulk copy per pair of tiles, as far ahead as the ring allows =====
      const int np = nt / 2;
      const unsigned bfull0 = pinned(smem_u32(&b_full[0])), bempty0 = pinned(smem_u32(&b_empty[0])), sB0 = pinned(smem_u32(sB));
      for (int P = 0; P < np; ++P) {
        const int s = P % NPAIR;
        PSIM_MBAR_WAIT(bempty0 + 8u * s, (unsigned)(((P / NPAIR) & 1) ^ 1));                 // both MMA warps done with the slot
        TR(7, 2 * P) TR(7, 2 * P + 1)
        mbar_expect_tx(bfull0 + 8u * s, 2 * TILE_BYTES);
        bulk_copy_g2s(sB0 + (unsigned)(s * 2 * TILE_BYTES), bops + (size_t)P * 2 * TILE_BYTES, 2 * TILE_BYTES, bfull0 + 8u * s);
      }
    }
  } else if (warp >= CW) {
    if (lane == 0) {
      // ===== MMA warp w: issues the tiles t = w, w + NCTL, ... into accumulator w =====
      // One thread sustains only ~1 tile per 950 cycles (every mbarrier wait / commit / MMA issue costs it
      // 70-190 cycles, measured: tools/microbench/tc_gram.cu), so the tiles are dealt to NCTL threads, each
      // with its own accumulator and barriers, and the TMA ring has its own warp.
      const int w = warp - CW;
      const int ntw = (nt - w + NCTL - 1) / NCTL;
      const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N2 >> 3) << 17) | ((unsigned)(M >> 4) << 24);
      const unsigned long long da0 = make_desc(smem_u32(sA)), da1 = make_desc(smem_u32(sA) + M * 32);
      const unsigned acc = tm + (unsigned)(w * N2);
      const unsigned dfull = pinned(smem_u32(&d_full[w])), dempty = pinned(smem_u32(&d_empty[w]));
      const unsigned bfull0 = pinned(smem_u32(&b_full[0])), bempty0 = pinned(smem_u32(&b_empty[0])), sB0 = pinned(smem_u32(sB));
#ifdef PSIM_TC_PROF
      long long p_bfull = 0, p_dempty = 0, p_issue = 0, p0 = clock64();
#define PT(acc) { long long p1 = clock64(); acc += p1 - p0; p0 = p1; }
#else
#define PT(acc) {}
#endif
      for (int k = 0; k < ntw; ++k) {
        const int t = w + NCTL * k, P = t >> 1, s = P % NPAIR;
        PSIM_MBAR_WAIT(bfull0 + 8u * s, (unsigned)((P / NPAIR) & 1));                        // tile pair landed
        PT(p_bfull) TR(0, t)
        PSIM_MBAR_WAIT(dempty, (unsigned)((k & 1) ^ 1));                                     // accumulator drained
        PT(p_dempty) TR(1, t)
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
        PT(p_issue) TR(2, t)
      }
#ifdef PSIM_TC_PROF
      if (blockIdx.x == 0) printf("mma warp %d per own tile: bfull %lld dempty %lld issue %lld\n", w, p_bfull / ntw, p_dempty / ntw, p_issue / ntw);
#endif
#undef PT
    }
  } else {
    // ===== consumers: TMEM -> registers -> running minimum (-> rare exact path) =====
    constexpr int COLS = N2 / NQ;                  // 32 columns = one load per tile
    int* q = reinterpret_cast<int*>(sB + RING_BYTES) + tid;      // this thread's queue: entry e at q[e * 32 CW]
    int cnt = 0;
    auto flush = [&]() { flush_queue<R>(a.pos_cur, c, ig, q, cnt, f, &in_range); cnt = 0; };
    const unsigned lane_base = tm + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)(half * COLS);
    const unsigned dfull0 = pinned(smem_u32(&d_full[0])), dempty0 = pinned(smem_u32(&d_empty[0]));
    const int jbase = half * COLS;
    // (Measured twice: a second register set to overlap the load of tile t + 1 with the reduction of tile t
    // makes ptxas spill inside the loop at the 96-register cap of 672 threads: 7.6 ms against 6.8 ms.  Overlap
    // comes from the 4 consumer warps per scheduler instead.)
#ifdef PSIM_TC_PROF
    long long c_wait = 0, c_ld = 0, c_q = 0, p0 = clock64();
#define PT(acc) { long long p1 = clock64(); acc += p1 - p0; p0 = p1; }
#else
#define PT(acc) {}
#endif
    for (int t0 = 0; t0 < nt; t0 += NBUF) {        // nt is a multiple of 8 (n_pad is a multiple of 1024)
      const unsigned par = (unsigned)((t0 / NBUF) & 1);
#pragma unroll
      for (int buf = 0; buf < NBUF; ++buf) {
        const int t = t0 + buf;
        PSIM_MBAR_WAIT(dfull0 + 8u * buf, par);    // (polling with one lane + __syncwarp: slower, measured)
        PT(c_wait) if (lane == 0) TR(8 + 2 * warp, t)
        asm volatile("tcgen05.fence::after_thread_sync;");
        unsigned r0[32];
        PSIM_TMEM_LD32(r0, lane_base + (unsigned)(buf * N2));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;");     // data is in registers: hand the buffer back
        __syncwarp();
        if (lane == 0) mbar_arrive(dempty0 + 8u * buf);
        PT(c_ld) if (lane == 0) TR(9 + 2 * warp, t)
        queue_chunk<R>(r0, cut, t * N2 + jbase, q, cnt);
        if (cnt > QC - 32) flush();                // rare; the 32 values are dead here, nothing to save around the call
        PT(c_q)
      }
    }
#ifdef PSIM_TC_PROF
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 15)) printf("consumer %d per tile: wait %lld ld %lld queue %lld\n", warp, c_wait / nt, c_ld / nt, c_q / nt);
#endif
#undef PT
    flush();                                       // the exact path, once the pipeline is done
    for (int off = 16; off > 0; off >>= 1) in_range += __shfl_xor_sync(0xffffffffu, in_range, off);
    if (lane == 0 && in_range) atomicAdd(pair_count, (unsigned long long)in_range);
  }

  // --- the column groups of a row meet (fixed order); epilogue by the thread of group 0 ---
  asm volatile("tcgen05.fence::before_thread_sync;");
  if (consumer && half > 0) { fpart[half][0][row] = f[0]; fpart[half][1][row] = f[1]; fpart[half][2][row] = f[2]; }
  __syncthreads();
  R first = (R)0, second = (R)0;
  if (consumer && half == 0 && ig < a.i_hi) {
#pragma unroll
    for (int q = 1; q < NQ; ++q) {
      f[0] = r_add(f[0], fpart[q][0][row]); f[1] = r_add(f[1], fpart[q][1][row]); f[2] = r_add(f[2], fpart[q][2][row]);
    }
    step_epilogue<R, MODE>(a, ig, f[0], f[1], f[2], first, second);
  }
  if (warp == CW) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
  }
#ifdef PSIM_TC_PROF
  __syncthreads();
  if (blockIdx.x == 0 && tid == 0) {
    const long long z = g_trace[7][256];
    for (int t = 256; t < 264; ++t) {
      printf("TR t=%d tma %lld mma: bfull %lld dempty %lld committed %lld\n", t, g_trace[7][t] - z, g_trace[0][t] - z, g_trace[1][t] - z, g_trace[2][t] - z);
      for (int w = 0; w < CW; ++w) printf("TR    warp %2d dfull %lld released %lld next-dfull %lld\n", w, g_trace[8 + 2 * w][t] - z, g_trace[9 + 2 * w][t] - z, g_trace[8 + 2 * w][t + 1] - z);
    }
  }
#endif
  if (MODE == MODE_FORCE_KICK_DRIFT) block_top2_store<R, TW>(first, second, a.top2_part, a.ticket, a.rank_pair);
}

}  // namespace tc
}  // namespace psim
