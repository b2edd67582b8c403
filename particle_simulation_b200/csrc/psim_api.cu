// psim_api.cu — the C ABI of include/psim.h over the kernels in this directory.
//
// One psim_sim owns one CUDA device context's worth of state: double-buffered
// position records, velocity / force / displacement records for the owned
// i-slice, a device-resident control block for the rebuild trigger, a frame
// ring of pinned host buffers, and (sharded runs) an NCCL communicator.
// Nothing here falls back to the CPU: without a device every compute entry
// point returns PSIM_ERR_NO_DEVICE.

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "psim_allpairs.cuh"
#include "psim_allpairs2.cuh"
#include "psim_allpairs2.cuh"
#include "psim_allpairs_tc.cuh"
#include "psim_cells.cuh"
#include "psim_integrate.cuh"

namespace psim {

static thread_local std::string g_last_error;

#define PSIM_CK(call)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      g_last_error = std::string(#call) + ": " + cudaGetErrorString(e_);               \
      return e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? PSIM_ERR_NO_DEVICE : PSIM_ERR_CUDA; \
    }                                                                                  \
  } while (0)

// ---- NCCL, bound lazily so that single-GPU users never load it -------------
struct NcclApi {
  void* so = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::once_flag once;
  bool ok = false;
  std::string why;
  // Hosts may drive several handles from several threads: bind exactly once.
  bool load() {
    std::call_once(once, [this] { ok = load_once(); if (!ok) why = g_last_error; });
    if (!ok) g_last_error = why;
    return ok;
  }
  bool load_once() {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      so = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (so) break;
    }
    if (!so) { g_last_error = std::string("dlopen libnccl: ") + dlerror(); return false; }
#define BIND(f) f = (decltype(f))dlsym(so, "nccl" #f); if (!f) { g_last_error = "dlsym nccl" #f; return false; }
    BIND(GetUniqueId) BIND(CommInitRank) BIND(CommDestroy) BIND(AllGather) BIND(GroupStart) BIND(GroupEnd)
    BIND(GetErrorString)
#undef BIND
    return true;
  }
};
static NcclApi g_nccl;

#define PSIM_NCCL(call)                                                                \
  do {                                                                                 \
    ncclResult_t r_ = (call);                                                          \
    if (r_ != ncclSuccess) {                                                           \
      g_last_error = std::string(#call) + ": " + g_nccl.GetErrorString(r_);            \
      return PSIM_ERR_COMM;                                                            \
    }                                                                                  \
  } while (0)

template <class R> struct NcclType;
template <> struct NcclType<float>  { static constexpr ncclDataType_t v = ncclFloat; };
template <> struct NcclType<double> { static constexpr ncclDataType_t v = ncclDouble; };

// stoermer:68-81 on the host in working precision R, plus the pre-filter margins.
template <class R>
static Consts<R> derive_consts(const psim_params& p) {
  Consts<R> c;
  const R sigma = (R)p.sigma, mass = (R)p.mass, rcf = (R)p.r_cut_factor;
  c.delta_t = (R)p.delta_t;
  c.epsilon = (R)p.epsilon;
  c.r_min = std::pow((R)2, (R)1 / (R)6) * sigma;
  c.r_cut = rcf * sigma;
  for (int d = 0; d < 3; ++d) {
    c.area[d] = (R)p.length_of_area[d];
    c.cell[d] = (R)p.length_of_cell[d];
    c.nc[d] = (int)std::lround(c.area[d] / c.cell[d]);     // nint, module_particle_management.f08:62
  }
  c.min_cell = std::min(c.cell[0], std::min(c.cell[1], c.cell[2]));
  c.dr_max_tolerable = c.min_cell - c.r_cut;
  // the host compiler targets baseline x86-64 (no FMA contraction), so these
  // round once per operation exactly like the oracle's -ffp-contract=off build
  const R q = c.r_min / c.r_cut;
  const R q2 = q * q;
  const R s = q2 * q2 * q2;
  c.F_rcut = (s / (c.r_cut * c.r_cut)) * ((R)1 - s);
  c.tmf = (R)0.5 * c.delta_t / mass;
  const R eps = RealTraits<R>::unit_roundoff;
  c.one_minus_k = (R)1 - (R)32 * eps;
  c.cut2_pad = c.r_cut * c.r_cut * ((R)1 + (R)32 * eps);
  c.huge = RealTraits<R>::huge;
  c.boff = (R)0;
  c.bslack = (R)0;
  if (sizeof(R) == 8) {
    // |x|^2 of anything that can be active: inside the box plus one cell of excursion
    R b = 0;
    for (int d = 0; d < 3; ++d) b += (c.area[d] + c.min_cell) * (c.area[d] + c.min_cell);
    c.boff = b;
    c.bslack = (R)16 * eps * b;
  }
  c.slices = std::max<int>(1, p.nb_of_threads);
  return c;
}

}  // namespace psim

using namespace psim;

// ---------------------------------------------------------------------------
struct psim_sim {
  virtual ~psim_sim() {}
  virtual psim_status advance(int64_t nsteps) = 0;
  virtual psim_status advance_timed(int64_t nsteps, double* ms) = 0;
  virtual psim_status synchronize() = 0;
  virtual psim_status get_frame(int32_t* id, void* pos, void* vel, uint8_t* active) = 0;
  virtual psim_status get_aux(void* disp, void* f_old, void* f_new) = 0;
  virtual psim_status get_counters(psim_counters* out) = 0;
  virtual psim_status set_option(int opt, int64_t value) = 0;
  virtual psim_status set_positions(const void* pos, const uint8_t* active) = 0;
  virtual psim_status snapshot_async(int slot) = 0;
  virtual psim_status snapshot_wait(int slot, const int32_t** id, const void** pos, const void** vel,
                                    const uint8_t** active) = 0;
  virtual psim_status count_pairs(int64_t* pairs) = 0;
  virtual psim_status energy(double* k, double* u, double* mom) = 0;
  virtual psim_status get_shard(int64_t* lo, int64_t* hi) = 0;
};

namespace psim {

template <class R>
class SimT final : public psim_sim {
 public:
  static constexpr int T = 128;                                   // threads per force block = j-tile
  static constexpr int RI = 2;                                    // i-particles per thread
  static constexpr int TS = 256;                                  // threads of the streaming kernels
  static constexpr int MAX_CLUSTER = 8;                           // portable cluster size limit
  int variant = PSIM_ALLPAIRS_TENSOR;   // psim_allpairs_kernel
  int jsplit_forced = 0;                // PSIM_JSPLIT: j segments per i-block of the tensor-core kernel (0 = choose)
  int ri_sel = 0;       // force i-particles per lane of the production all-pairs kernel (6, 7, 8; 0 = choose)
  int cluster_forced = 0;
  int warps_forced = 0;
  int sm_count = 148;

  int device = 0;
  int64_t n = 0, n_pad = 0, shard = 0, i_lo = 0, i_hi = 0;
  int rank = 0, world = 1;
  Consts<R> c;
  double mass = 1.0;
  cudaStream_t stream = nullptr, side = nullptr;
  Real4<R>* pos[2] = {nullptr, nullptr};
  int cur = 0;
  Real4<R>*vel = nullptr, *f_old = nullptr, *f_new = nullptr, *disp = nullptr;
  R* top2_part = nullptr;
  int64_t top2_cap = 0;
  R* top2_rank = nullptr;       // [2 * world]
  Control* ctrl = nullptr;
  int parity = 0;
  bool pending_drift = false;   // the previous fused kernel already produced x_{n+1}
  std::vector<int32_t> ids;
  R* tmp3 = nullptr;            // [3 * n_pad] scratch for pack/unpack
  unsigned char* tmp_active = nullptr;
  unsigned char* tc_bops = nullptr;     // tensor-core variant: B operand tiles of all particles
  unsigned long long* tc_pairs = nullptr;  // ... and the in-range pair count of its last launch (psim_counters.filter_pairs)
  R* tc_part = nullptr;                 // ... j-split: partial forces [grid][3][128]
  unsigned* tc_arrived = nullptr;       // ... j-split: arrival counter per i-block (zero between launches)
  int64_t tc_part_cap = 0;
  double* red_scratch = nullptr;        // 8 doubles: reductions of count_pairs / energy (no allocation per call)
  bool attr_tc[3] = {false, false, false};
  tc::Scale16 scale16 = {1.0f, 1.0f};
  CellPath<R> cells;
  // options
  int algo = PSIM_FORCE_ALL_PAIRS;
  bool fused = true, timing = false;
  // counters
  int64_t steps = 0, launches = 0, force_launches = 0, force_timed = 0;
  double force_ms = 0.0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pool;
  size_t ev_used = 0;
  // frame ring
  struct Slot { R* d_pos = nullptr; R* d_vel = nullptr; unsigned char* d_act = nullptr;
                R* h_pos = nullptr; R* h_vel = nullptr; unsigned char* h_act = nullptr;
                cudaEvent_t ready = nullptr, done = nullptr; bool used = false; };
  Slot slots[PSIM_FRAME_SLOTS];
  ncclComm_t comm = nullptr;
  // Peer stores (PeerPush, psim_allpairs.cuh): every rank's position buffers, trigger table and flag array, mapped
  // through CUDA IPC at create; the fused force kernel then stores x_{n+1} of its particles into all of them and
  // the per-step all-gather shrinks to exchange_kernel.  Off (NCCL all-gather every step) when any
  // rank could not map its peers, with PSIM_GATHER=nccl, and for the unfused drift.
  bool peer_push = false;
  Real4<R>* peer_pos[2][MAX_PEERS] = {};
  R* peer_top2[MAX_PEERS] = {};
  unsigned long long* peer_flags[MAX_PEERS] = {};
  unsigned long long* flags = nullptr;   // [world] this rank's flags, written by the peers
  unsigned long long epoch = 0;          // steps exchanged by peer stores so far (the same on every rank)
  static constexpr unsigned long long kAwaitTimeoutNs = 30000000000ull;   // a peer that silent has died
  int64_t peer_push_steps = 0;
  std::vector<void*> ipc_opened;

  ~SimT() override { destroy(); }

  void destroy() {
    if (device >= 0) cudaSetDevice(device);
    if (stream) cudaStreamSynchronize(stream);
    if (side) cudaStreamSynchronize(side);
    for (void* p : ipc_opened) cudaIpcCloseMemHandle(p);
    ipc_opened.clear(); peer_push = false;
    if (comm) { g_nccl.CommDestroy(comm); comm = nullptr; }
    cudaFree(flags); flags = nullptr;
    cells.destroy();
    for (auto& s : slots) {
      cudaFree(s.d_pos); cudaFree(s.d_vel); cudaFree(s.d_act);
      cudaFreeHost(s.h_pos); cudaFreeHost(s.h_vel); cudaFreeHost(s.h_act);
      if (s.ready) cudaEventDestroy(s.ready);
      if (s.done) cudaEventDestroy(s.done);
      s = Slot();
    }
    for (auto& e : ev_pool) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    ev_pool.clear();
    cudaFree(pos[0]); cudaFree(pos[1]); cudaFree(vel); cudaFree(f_old); cudaFree(f_new); cudaFree(disp);
    cudaFree(top2_part); cudaFree(top2_rank); cudaFree(ctrl); cudaFree(tmp3); cudaFree(tmp_active); cudaFree(tc_bops); cudaFree(tc_pairs); cudaFree(tc_part); cudaFree(tc_arrived); cudaFree(red_scratch);
    tc_bops = nullptr; tc_pairs = nullptr; tc_part = nullptr; tc_arrived = nullptr; red_scratch = nullptr; tc_part_cap = 0;
    pos[0] = pos[1] = vel = f_old = f_new = disp = nullptr;
    top2_part = top2_rank = nullptr; ctrl = nullptr; tmp3 = nullptr; tmp_active = nullptr;
    if (stream) cudaStreamDestroy(stream);
    if (side) cudaStreamDestroy(side);
    stream = side = nullptr;
  }

  static int64_t blocks(int64_t count, int per) { return (count + per - 1) / per; }

  psim_status init(const psim_params& p, int64_t n_, const int32_t* id, const void* pos_h, const void* vel_h,
                   const uint8_t* active_h, int device_, int rank_, int world_, const void* uid) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
      g_last_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
      cudaGetLastError();
      return PSIM_ERR_NO_DEVICE;
    }
    if (device_ < 0) PSIM_CK(cudaGetDevice(&device_));
    device = device_;
    PSIM_CK(cudaSetDevice(device));
    n = n_; rank = rank_; world = world_;
    PSIM_CK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    if (const char* e = std::getenv("PSIM_RI")) { const int v = std::atoi(e); if (v >= 6 && v <= 8) ri_sel = v; }   // tuning knobs
    if (const char* e = std::getenv("PSIM_WARPS")) { const int v = std::atoi(e); if (v == 4 || v == 8 || v == 16) warps_forced = v; }
    if (const char* e = std::getenv("PSIM_ALLPAIRS_KERNEL")) {   // testing: run everything through one kernel
      const int v = std::atoi(e);
      if (v == PSIM_ALLPAIRS_TENSOR || v == PSIM_ALLPAIRS_DIFFERENCE_FORM || v == PSIM_ALLPAIRS_CUDA_CORE) variant = v;
    }
    if (const char* e = std::getenv("PSIM_JSPLIT")) { const int v = std::atoi(e); if (v >= 1 && v <= tc::MAX_JSPLIT) jsplit_forced = v; }
    if (const char* e = std::getenv("PSIM_CLUSTER")) {
      const int v = std::atoi(e);
      if (v >= 1 && v <= MAX_CLUSTER) cluster_forced = v;
    }
    c = derive_consts<R>(p);
    mass = p.mass;
    {   // FP16 operand scale of the tensor-core filter: largest power of two with (1.4 x half diagonal x S)^2 <= 60000
      double d2 = 0;
      for (int d = 0; d < 3; ++d) d2 += p.length_of_area[d] * p.length_of_area[d];
      const double lim = std::sqrt((double)tc::BIG) / (1.4 * 0.5 * std::sqrt(d2));
      const float S = (float)std::ldexp(1.0, (int)std::floor(std::log2(lim)));
      scale16.S = S; scale16.S2 = S * S;
    }
    if (!(c.min_cell >= c.r_cut)) {
      // the reference's lists would then drop pairs inside the cutoff
      // (module_particle_management.f08:384 keeps r <= min_length_of_cell only)
      g_last_error = "LENGTH_OF_CELL smaller than R_CUT*SIGMA is not supported";
      return PSIM_ERR_UNSUPPORTED;
    }
    const int64_t gran = (int64_t)T * MAX_CLUSTER * world;   // every rank's j range splits into 1..8 equal tile runs
    n_pad = (n + gran - 1) / gran * gran;
    shard = n_pad / world;
    i_lo = std::min<int64_t>(n, shard * rank);
    i_hi = std::min<int64_t>(n, shard * (rank + 1));
    PSIM_CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    PSIM_CK(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
    const size_t rec = sizeof(Real4<R>) * (size_t)n_pad;
    PSIM_CK(cudaMalloc(&pos[0], rec)); PSIM_CK(cudaMalloc(&pos[1], rec));
    PSIM_CK(cudaMalloc(&vel, rec)); PSIM_CK(cudaMalloc(&f_old, rec));
    PSIM_CK(cudaMalloc(&f_new, rec)); PSIM_CK(cudaMalloc(&disp, rec));
    // most blocks any drift-producing kernel launches (tensor-core kernel: i-blocks x j segments)
    top2_cap = std::max<int64_t>(std::max<int64_t>(blocks(n_pad, 64), blocks(n_pad, 32 * 6) * MAX_CLUSTER),
                                 blocks(shard, tc::M) * tc::MAX_JSPLIT) + 8;
    PSIM_CK(cudaMalloc(&top2_part, sizeof(R) * 2 * top2_cap));
    PSIM_CK(cudaMalloc(&top2_rank, sizeof(R) * 2 * c.slices * world));   // [world][slices][2]
    PSIM_CK(cudaMemsetAsync(top2_rank, 0, sizeof(R) * 2 * c.slices * world, stream));
    PSIM_CK(cudaMalloc(&red_scratch, 8 * sizeof(double)));
    PSIM_CK(cudaMalloc(&ctrl, sizeof(Control)));
    PSIM_CK(cudaMemsetAsync(ctrl, 0, sizeof(Control), stream));
    PSIM_CK(cudaMalloc(&tmp3, sizeof(R) * 3 * n_pad));
    PSIM_CK(cudaMalloc(&tmp_active, (size_t)n_pad));
    ids.assign(id, id + n);
    if (world > 1) {
      if (!g_nccl.load()) return PSIM_ERR_COMM;
      ncclUniqueId nid;
      std::memcpy(&nid, uid, sizeof nid);
      PSIM_NCCL(g_nccl.CommInitRank(&comm, world, nid, rank));
      psim_status st = map_peers();
      if (st != PSIM_OK) return st;
    }
    // velocities, then positions + initial activation + initial forces
    if (vel_h) PSIM_CK(cudaMemcpyAsync(tmp3, vel_h, sizeof(R) * 3 * n, cudaMemcpyHostToDevice, stream));
    unpack3_kernel<R, TS><<<blocks(n_pad, TS), TS, 0, stream>>>(vel, vel_h ? tmp3 : nullptr, n, n_pad); ++launches;
    PSIM_CK(cudaStreamSynchronize(stream));   // tmp3 is reused below
    psim_status st = set_positions(pos_h, active_h);
    if (st != PSIM_OK) return st;
    return PSIM_OK;
  }

  // Upload positions, redo the start-up activation, zero displacement and
  // forces, and evaluate the initial forces (program_...:77).
  psim_status set_positions(const void* pos_h, const uint8_t* active_h) override {
    PSIM_CK(cudaSetDevice(device));
    if (!pos_h) return PSIM_ERR_INVALID_ARGUMENT;
    PSIM_CK(cudaMemcpyAsync(tmp3, pos_h, sizeof(R) * 3 * n, cudaMemcpyHostToDevice, stream));
    if (active_h) PSIM_CK(cudaMemcpyAsync(tmp_active, active_h, (size_t)n, cudaMemcpyHostToDevice, stream));
    init_positions_kernel<R, TS><<<blocks(n_pad, TS), TS, 0, stream>>>(pos[0], pos[1], tmp3,
                                                                     active_h ? tmp_active : nullptr, n, n_pad, c);
    unpack3_kernel<R, TS><<<blocks(n_pad, TS), TS, 0, stream>>>(disp, nullptr, n, n_pad);
    unpack3_kernel<R, TS><<<blocks(n_pad, TS), TS, 0, stream>>>(f_old, nullptr, n, n_pad);
    unpack3_kernel<R, TS><<<blocks(n_pad, TS), TS, 0, stream>>>(f_new, nullptr, n, n_pad);
    launches += 4;
    cur = 0; pending_drift = false;
    cells.invalidate();
    psim_status st = launch_force(MODE_FORCE);
    if (st != PSIM_OK) return st;
    PSIM_CK(cudaGetLastError());
    return PSIM_OK;
  }

  // Exchange CUDA IPC handles of the shared arrays over the communicator and map every peer's.  Collective.  Any
  // rank that cannot (same process, no peer access, PSIM_GATHER=nccl) makes ALL ranks keep the NCCL all-gather.
  psim_status map_peers() {
    struct Handles { cudaIpcMemHandle_t pos0, pos1, top2, flags; int ok; int pad[15]; };
    static_assert(sizeof(Handles) == 4 * 64 + 64, "handle record");
    PSIM_CK(cudaMalloc(&flags, sizeof(unsigned long long) * MAX_PEERS));
    PSIM_CK(cudaMemsetAsync(flags, 0, sizeof(unsigned long long) * MAX_PEERS, stream));
    Handles mine;
    std::memset(&mine, 0, sizeof mine);
    const char* env = std::getenv("PSIM_GATHER");
    mine.ok = world <= MAX_PEERS && !(env && std::strcmp(env, "nccl") == 0);
    if (mine.ok) {
      mine.ok = cudaIpcGetMemHandle(&mine.pos0, pos[0]) == cudaSuccess && cudaIpcGetMemHandle(&mine.pos1, pos[1]) == cudaSuccess &&
                cudaIpcGetMemHandle(&mine.top2, top2_rank) == cudaSuccess && cudaIpcGetMemHandle(&mine.flags, flags) == cudaSuccess;
      cudaGetLastError();
    }
    unsigned char* d_all = nullptr;
    PSIM_CK(cudaMalloc(&d_all, sizeof(Handles) * world));
    std::vector<Handles> all(world);
    auto exchange = [&]() -> psim_status {
      PSIM_CK(cudaMemcpyAsync(d_all + sizeof(Handles) * rank, &mine, sizeof mine, cudaMemcpyHostToDevice, stream));
      PSIM_NCCL(g_nccl.AllGather(d_all + sizeof(Handles) * rank, d_all, sizeof(Handles), ncclUint8, comm, stream));
      PSIM_CK(cudaMemcpyAsync(all.data(), d_all, sizeof(Handles) * world, cudaMemcpyDeviceToHost, stream));
      PSIM_CK(cudaStreamSynchronize(stream));
      return PSIM_OK;
    };
    psim_status st = exchange();
    if (st != PSIM_OK) { cudaFree(d_all); return st; }
    bool ok = true;
    for (int r = 0; r < world; ++r) ok = ok && all[r].ok;
    if (ok) {
      for (int r = 0; r < world && ok; ++r) {
        if (r == rank) { peer_pos[0][r] = pos[0]; peer_pos[1][r] = pos[1]; peer_top2[r] = top2_rank; peer_flags[r] = flags; continue; }
        void* q[4] = {nullptr, nullptr, nullptr, nullptr};
        const cudaIpcMemHandle_t* hs[4] = {&all[r].pos0, &all[r].pos1, &all[r].top2, &all[r].flags};
        for (int k = 0; k < 4 && ok; ++k) {
          ok = cudaIpcOpenMemHandle(&q[k], *hs[k], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
          if (ok) ipc_opened.push_back(q[k]);
        }
        cudaGetLastError();
        peer_pos[0][r] = (Real4<R>*)q[0]; peer_pos[1][r] = (Real4<R>*)q[1];
        peer_top2[r] = (R*)q[2]; peer_flags[r] = (unsigned long long*)q[3];
      }
    }
    // second round: did every rank map every peer?
    mine.ok = ok ? 1 : 0;
    st = exchange();
    cudaFree(d_all);
    if (st != PSIM_OK) return st;
    for (int r = 0; r < world; ++r) ok = ok && all[r].ok;
    if (!ok) {
      for (void* p : ipc_opened) cudaIpcCloseMemHandle(p);
      ipc_opened.clear();
      cudaGetLastError();
    }
    peer_push = ok;
    return PSIM_OK;
  }

  StepArgs<R> step_args() const {
    StepArgs<R> a;
    a.pos_cur = pos[cur]; a.pos_nxt = pos[cur ^ 1];
    a.vel = vel; a.f_old = f_old; a.f_new = f_new; a.disp = disp;
    a.top2_part = top2_part;
    a.rank_pair = top2_rank + 2 * c.slices * rank;   // slices > 1: overwritten per slice by slice_top2_kernel
    a.ticket = &ctrl->ticket;
    a.n_pad = n_pad; a.i_lo = i_lo; a.i_hi = i_hi;
    a.c = c;
    a.peers.world = 0; a.peers.rank = rank;
    if (peer_push && fused) {            // only MODE_FORCE_KICK_DRIFT kernels store x_{n+1}
      a.peers.world = world;
      for (int r = 0; r < world; ++r) a.peers.pos_nxt[r] = peer_pos[cur ^ 1][r];
    }
    return a;
  }

  // j segments per i-block of the tensor-core kernel: a block is 128 rows (the MMA's M), one block per SM,
  // so a launch with few i-blocks (a shard, a small N) leaves SMs idle in its last wave; cutting j into S
  // segments multiplies the blocks.  The smallest S that fills the waves wins (each block pays a fixed
  // set-up, and the segments meet through global scratch).
  int pick_jsplit(int64_t iblocks) const {
    const int nt8 = (int)(n_pad / (8 * tc::N2));
    const int smax = std::max(1, std::min(tc::MAX_JSPLIT, nt8));
    if (jsplit_forced) return std::min(jsplit_forced, smax);
    int best = 1;
    double best_eff = -1.0;
    for (int S = 1; S <= smax; ++S) {
      const int64_t g = iblocks * S, waves = (g + sm_count - 1) / sm_count;
      const double eff = (double)g / (double)(waves * sm_count) * (1.0 - 0.004 * (S - 1));
      if (eff > best_eff + 1e-9) { best_eff = eff; best = S; }
    }
    return best;
  }

  // Tiling of the production all-pairs kernel: RI i-particles per lane (i-tile = 32 RI), W warps
  // per block and S blocks per cluster (the j range is split W x S ways).  Registers allow 16 warps
  // per SM, i.e. 16/W blocks.  Small i-counts (one GPU at N = 65,536, or any sharded run) need the
  // j split to fill the machine; the split costs little inside a block (shared-memory reduction)
  // and more across a cluster (measured: clusters schedule less densely), so W grows first.
  struct Tiling { int ri, w, s; };
  Tiling pick_tiling() const {
    Tiling best = {8, 4, 1};
    if (!ri_sel && !cluster_forced && !warps_forced && blocks(i_hi - i_lo, 32 * 8) >= (int64_t)sm_count * 4)
      return best;       // more than a wave of full-size tiles: the block scheduler backfills well enough
    double best_eff = -1.0;
    for (int ri = 8; ri >= 6; --ri) {
      if (ri_sel && ri != ri_sel) continue;
      const int64_t itiles = blocks(i_hi - i_lo, 32 * ri);
      for (int w = 4; w <= 16; w *= 2) {
        if (warps_forced && w != warps_forced) continue;
        const int64_t slots = (int64_t)sm_count * (16 / w);
        for (int S = 1; S <= MAX_CLUSTER; ++S) {
          if (cluster_forced && S != cluster_forced) continue;
          const int64_t ctas = itiles * S;
          const int64_t waves = (ctas + slots - 1) / slots;
          double eff = (double)ctas / (double)(waves * slots);
          eff *= 1.0 - 0.015 * (8 - ri);                     // fewer i per lane: a little less efficient
          if (w == 8) eff *= 1.03;                           // measured: two 8-warp blocks per SM run a little denser
          if (S > 1) eff *= 0.95 - 0.03 * (S - 1);           // measured cost of cluster scheduling
          if (eff > best_eff) { best_eff = eff; best = {ri, w, S}; }
        }
      }
    }
    return best;
  }

  template <int MODE, int RIv, int Wv>
  cudaError_t launch_cluster(const StepArgs<R>& a, int S, int64_t* grid_out) {
    const int64_t itiles = blocks(i_hi - i_lo, 32 * RIv);
    auto kern = allpairs_warp_kernel<R, RIv, Wv, MODE>;
    constexpr size_t smem = allpairs_warp_smem<R, RIv, Wv>();
    {   // static tiles + dynamic accumulators can exceed the 48 KB default (a host-side call, no shared state)
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(itiles * S));
    cfg.blockDim = dim3(32 * Wv);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = S; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
    cfg.attrs = &attr; cfg.numAttrs = 1;
    *grid_out = itiles * S;
    return cudaLaunchKernelEx(&cfg, kern, a, S);
  }

  template <int MODE, int RIv>
  cudaError_t launch_cluster_w(const StepArgs<R>& a, int w, int S, int64_t* grid_out) {
    if (w == 16) return launch_cluster<MODE, RIv, 16>(a, S, grid_out);
    if (w == 8) return launch_cluster<MODE, RIv, 8>(a, S, grid_out);
    return launch_cluster<MODE, RIv, 4>(a, S, grid_out);
  }

  template <int MODE>
  psim_status launch_allpairs(const StepArgs<R>& a, int64_t* grid_out) {
    if (variant == PSIM_ALLPAIRS_TENSOR) {   // distance filter on the tensor cores (tcgen05 + TMEM): the default
      const int64_t iblocks = blocks(i_hi - i_lo, tc::M);
      tc::JSplit<R> js;
      js.S = pick_jsplit(iblocks); js.part = nullptr; js.arrived = nullptr;
      *grid_out = iblocks * js.S;
      if (!tc_bops) PSIM_CK(cudaMalloc(&tc_bops, (size_t)n_pad * tc::K * 2));   // 32 B per particle
      if (!tc_pairs) PSIM_CK(cudaMalloc(&tc_pairs, sizeof *tc_pairs));
      if (js.S > 1) {
        if (tc_part_cap < *grid_out) {
          cudaFree(tc_part); tc_part = nullptr;
          PSIM_CK(cudaMalloc(&tc_part, sizeof(R) * 3 * tc::M * (size_t)*grid_out));
          tc_part_cap = *grid_out;
        }
        if (!tc_arrived) {
          const size_t cnt = (size_t)blocks(shard, tc::M) + 1;
          PSIM_CK(cudaMalloc(&tc_arrived, sizeof(unsigned) * cnt));
          PSIM_CK(cudaMemsetAsync(tc_arrived, 0, sizeof(unsigned) * cnt, stream));
        }
        js.part = tc_part; js.arrived = tc_arrived;
      }
      tc::tc_prepare_kernel<R><<<(unsigned)blocks(n_pad, 256), 256, 0, stream>>>(a.pos_cur, tc_bops, n_pad, c, scale16, tc_pairs);
      ++launches;
      constexpr int kSmem = tc::SMEM_BYTES;
      if (!attr_tc[MODE]) {
        PSIM_CK(cudaFuncSetAttribute(tc::allpairs_tc_kernel<R, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        attr_tc[MODE] = true;
      }
      tc::allpairs_tc_kernel<R, MODE><<<(unsigned)*grid_out, tc::TW, kSmem, stream>>>(a, tc_bops, tc_pairs, js, scale16);
      return PSIM_OK;
    }
    if (variant == PSIM_ALLPAIRS_DIFFERENCE_FORM) {   // the literal reading of stoermer:117-121 in the working precision
      *grid_out = blocks(i_hi - i_lo, T * RI);
      allpairs_step_kernel<R, T, RI, MODE, true><<<*grid_out, T, 0, stream>>>(a);
      return PSIM_OK;
    }
    const Tiling t = pick_tiling();
    if (t.ri == 6) PSIM_CK((launch_cluster_w<MODE, 6>(a, t.w, t.s, grid_out)));
    else if (t.ri == 7) PSIM_CK((launch_cluster_w<MODE, 7>(a, t.w, t.s, grid_out)));
    else PSIM_CK((launch_cluster_w<MODE, 8>(a, t.w, t.s, grid_out)));
    return PSIM_OK;
  }

  psim_status launch_force(int mode) {
    if (i_hi <= i_lo) return PSIM_OK;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (timing) {
      if (ev_used == ev_pool.size()) {
        if (ev_pool.size() >= 8192) { psim_status st = drain_events(); if (st != PSIM_OK) return st; }
        else {
          cudaEvent_t a, b;
          PSIM_CK(cudaEventCreate(&a)); PSIM_CK(cudaEventCreate(&b));
          ev_pool.emplace_back(a, b);
        }
      }
      e0 = ev_pool[ev_used].first; e1 = ev_pool[ev_used].second; ++ev_used;
      PSIM_CK(cudaEventRecord(e0, stream));
    }
    if (algo == PSIM_FORCE_CELL_BINNED) {
      psim_status st = cells.force(*this, mode);
      if (st != PSIM_OK) return st;
    } else {
      const StepArgs<R> a = step_args();
      int64_t grid = 0;
      psim_status st;
      if (mode == MODE_FORCE) st = launch_allpairs<MODE_FORCE>(a, &grid);
      else if (mode == MODE_FORCE_KICK) st = launch_allpairs<MODE_FORCE_KICK>(a, &grid);
      else st = launch_allpairs<MODE_FORCE_KICK_DRIFT>(a, &grid);
      if (st != PSIM_OK) return st;
      ++launches;
    }
    ++force_launches;
    if (timing) PSIM_CK(cudaEventRecord(e1, stream));
    return PSIM_OK;
  }

  psim_status drain_events() {
    PSIM_CK(cudaStreamSynchronize(stream));
    for (size_t k = 0; k < ev_used; ++k) {
      float ms = 0;
      PSIM_CK(cudaEventElapsedTime(&ms, ev_pool[k].first, ev_pool[k].second));
      force_ms += ms; ++force_timed;
    }
    ev_used = 0;
    return PSIM_OK;
  }

  // One cooperative launch: trigger, rebuild side effects, (cell path) re-binning and the
  // cell-ordered position copy.  pos_new is the buffer that is current from here on.
  int house_grid_max = 0;
  psim_status launch_housekeeping(Real4<R>* pos_new, Real4<R>* pos_other, bool with_trigger, bool force_rebin) {
    if (!house_grid_max) {
      int per_sm = 0;
      PSIM_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, housekeeping_kernel<R>, 256, 0));
      house_grid_max = std::max(1, per_sm) * sm_count;
    }
    HouseArgs<R> h = {};
    h.pos_new = pos_new; h.pos_other = pos_other; h.disp = disp;
    h.rank_pairs = with_trigger ? top2_rank : nullptr;
    h.world = world;
    h.ctrl = ctrl; h.parity = parity;
    h.n = n; h.i_lo = i_lo; h.i_hi = i_hi; h.c = c;
    h.cells = (algo == PSIM_FORCE_CELL_BINNED && cells.allocated()) ? 1 : 0;
    h.force_rebin = force_rebin ? 1 : 0;
    // a rank without particles never builds the cell tables, yet must keep its copies of the inactive records
    h.carry_all = (algo == PSIM_FORCE_CELL_BINNED && !h.cells && peer_push) ? 1 : 0;
    if (h.cells) cells.fill_args(h);
    const int grid = (int)std::min<int64_t>(house_grid_max, std::max<int64_t>(2, blocks(n, 256)));
    void* args[] = {&h};
    PSIM_CK(cudaLaunchCooperativeKernel((const void*)housekeeping_kernel<R>, dim3(grid), dim3(256), args, 0, stream));
    ++launches;
    if (h.cells && (force_rebin || cells.binned)) cells.binned = true;
    return PSIM_OK;
  }

  // After x_{n+1} of the owned slice is in pos[cur^1]: exchange (sharded), then housekeeping, swap.
  // pushed: the drift came from a fused force kernel that stored x_{n+1} into the peers' buffers as well
  psim_status finish_drift(bool pushed = false) {
    Real4<R>* nxt = pos[cur ^ 1];
    const int ns = c.slices;
    R* mine = top2_rank + 2 * ns * rank;
    if (i_hi <= i_lo) PSIM_CK(cudaMemsetAsync(mine, 0, 2 * ns * sizeof(R), stream));   // nothing owned
    else if (ns > 1) {
      // NB_OF_THREADS > 1: the reference's per-slice trigger (program:71,86,92) over the part of every
      // slice this handle owns; sharded handles exchange the per-slice pairs below.
      slice_top2_kernel<R><<<ns, 256, 0, stream>>>(nxt, disp, n, i_lo, i_hi, mine, c);
      ++launches;
    }
    if (world > 1 && pushed) {
      ++epoch; ++peer_push_steps;
      PeerMail<R> m;
      m.world = world; m.rank = rank; m.nvals = 2 * ns; m.mine = mine;
      for (int r = 0; r < world; ++r) { m.top2[r] = peer_top2[r]; m.flags[r] = peer_flags[r]; }
      exchange_kernel<R><<<1, 64, 0, stream>>>(m, epoch, flags, ctrl, kAwaitTimeoutNs);
      ++launches;
    } else if (world > 1) {
      PSIM_NCCL(g_nccl.GroupStart());
      PSIM_NCCL(g_nccl.AllGather(nxt + shard * rank, nxt, (size_t)shard * 4, NcclType<R>::v, comm, stream));
      PSIM_NCCL(g_nccl.AllGather(mine, top2_rank, (size_t)2 * ns, NcclType<R>::v, comm, stream));
      PSIM_NCCL(g_nccl.GroupEnd());
    }
    psim_status st = launch_housekeeping(nxt, pos[cur], true, false);
    if (st != PSIM_OK) return st;
    parity ^= 1;
    cur ^= 1;
    return PSIM_OK;
  }

  psim_status advance(int64_t nsteps) override {
    PSIM_CK(cudaSetDevice(device));
    if (nsteps < 0) return PSIM_ERR_INVALID_ARGUMENT;
    const bool can_fuse = fused;
    for (int64_t s = 0; s < nsteps; ++s) {
      if (!pending_drift) {
        if (i_hi > i_lo) {
          const int64_t grid = blocks(i_hi - i_lo, TS);
          drift_kernel<R, TS><<<grid, TS, 0, stream>>>(pos[cur], pos[cur ^ 1], vel, f_old, f_new, disp, top2_part,
                                                       &ctrl->ticket, top2_rank + 2 * c.slices * rank, i_lo, i_hi, c);
          ++launches;
        }
      }
      // the previous trip of THIS call ran a fused kernel (whether or not this rank owns particles): every rank
      // takes the same branch
      const bool pushed = peer_push && can_fuse && s > 0;
      pending_drift = false;
      psim_status st = finish_drift(pushed);
      if (st != PSIM_OK) return st;
      const bool last = (s == nsteps - 1);
      if (can_fuse) {
        st = launch_force(last ? MODE_FORCE_KICK : MODE_FORCE_KICK_DRIFT);
        if (st != PSIM_OK) return st;
        pending_drift = !last && (i_hi > i_lo);
      } else {
        st = launch_force(MODE_FORCE);
        if (st != PSIM_OK) return st;
        if (i_hi > i_lo) {
          kick_kernel<R, TS><<<blocks(i_hi - i_lo, TS), TS, 0, stream>>>(pos[cur], vel, f_old, f_new, i_lo, i_hi, c);
          ++launches;
        }
      }
      ++steps;
    }
    PSIM_CK(cudaGetLastError());
    return PSIM_OK;
  }

  psim_status advance_timed(int64_t nsteps, double* ms) override {
    PSIM_CK(cudaSetDevice(device));
    struct Events {                                 // released on every return path
      cudaEvent_t a = nullptr, b = nullptr;
      ~Events() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    } ev;
    PSIM_CK(cudaEventCreate(&ev.a)); PSIM_CK(cudaEventCreate(&ev.b));
    PSIM_CK(cudaStreamSynchronize(stream));
    PSIM_CK(cudaEventRecord(ev.a, stream));
    psim_status st = advance(nsteps);
    if (st != PSIM_OK) return st;
    PSIM_CK(cudaEventRecord(ev.b, stream));
    PSIM_CK(cudaEventSynchronize(ev.b));
    float f = 0;
    PSIM_CK(cudaEventElapsedTime(&f, ev.a, ev.b));
    if (ms) *ms = f;
    return PSIM_OK;
  }

  psim_status synchronize() override {
    PSIM_CK(cudaSetDevice(device));
    PSIM_CK(cudaStreamSynchronize(stream));
    PSIM_CK(cudaStreamSynchronize(side));
    if (peer_push && peer_push_steps > 0) {          // did every await see its peers?
      unsigned t = 0;
      PSIM_CK(cudaMemcpy(&t, &ctrl->comm_timeout, sizeof t, cudaMemcpyDeviceToHost));
      if (t) { g_last_error = "a peer's positions did not arrive within 30 s (peer-store exchange)"; return PSIM_ERR_COMM; }
    }
    return PSIM_OK;
  }

  psim_status fetch3(void* host, const Real4<R>* src, int64_t lo, int64_t hi) {
    if (!host || hi <= lo) return PSIM_OK;
    pack3_kernel<R, TS><<<blocks(hi - lo, TS), TS, 0, stream>>>(tmp3, src, lo, hi); ++launches;
    PSIM_CK(cudaMemcpyAsync((R*)host + 3 * lo, tmp3 + 3 * lo, sizeof(R) * 3 * (hi - lo), cudaMemcpyDeviceToHost, stream));
    PSIM_CK(cudaStreamSynchronize(stream));
    return PSIM_OK;
  }

  psim_status get_frame(int32_t* id, void* pos_h, void* vel_h, uint8_t* active_h) override {
    PSIM_CK(cudaSetDevice(device));
    if (id) std::memcpy(id, ids.data(), sizeof(int32_t) * n);
    psim_status st = fetch3(pos_h, pos[cur], 0, n);
    if (st != PSIM_OK) return st;
    st = fetch3(vel_h, vel, i_lo, i_hi);
    if (st != PSIM_OK) return st;
    if (active_h) {
      pack_active_kernel<R, TS><<<blocks(n, TS), TS, 0, stream>>>(tmp_active, pos[cur], n, c, ctrl); ++launches;
      PSIM_CK(cudaMemcpyAsync(active_h, tmp_active, (size_t)n, cudaMemcpyDeviceToHost, stream));
      PSIM_CK(cudaStreamSynchronize(stream));
    }
    return PSIM_OK;
  }

  psim_status get_aux(void* d, void* fo, void* fn) override {
    PSIM_CK(cudaSetDevice(device));
    psim_status st;
    if ((st = fetch3(d, disp, i_lo, i_hi)) != PSIM_OK) return st;
    if ((st = fetch3(fo, f_old, i_lo, i_hi)) != PSIM_OK) return st;
    if ((st = fetch3(fn, f_new, i_lo, i_hi)) != PSIM_OK) return st;
    return PSIM_OK;
  }

  psim_status get_counters(psim_counters* out) override {
    PSIM_CK(cudaSetDevice(device));
    if (!out) return PSIM_ERR_INVALID_ARGUMENT;
    psim_status st = drain_events();
    if (st != PSIM_OK) return st;
    PSIM_CK(cudaMemsetAsync(&ctrl->n_active, 0, sizeof(unsigned long long), stream));
    pack_active_kernel<R, TS><<<blocks(n, TS), TS, 0, stream>>>(nullptr, pos[cur], n, c, ctrl); ++launches;
    Control h;
    PSIM_CK(cudaMemcpyAsync(&h, ctrl, sizeof h, cudaMemcpyDeviceToHost, stream));
    unsigned long long fp = 0;
    const bool tracked = algo == PSIM_FORCE_ALL_PAIRS && variant == PSIM_ALLPAIRS_TENSOR && tc_pairs;
    if (tracked) PSIM_CK(cudaMemcpyAsync(&fp, tc_pairs, sizeof fp, cudaMemcpyDeviceToHost, stream));
    PSIM_CK(cudaStreamSynchronize(stream));
    out->filter_pairs = tracked ? (int64_t)fp : -1;
    out->filter_tests = -1;
    if (algo == PSIM_FORCE_ALL_PAIRS) out->filter_tests = (i_hi - i_lo) * (n - 1);
    else if (cells.tested && timing) {
      unsigned long long ft = 0;
      PSIM_CK(cudaMemcpyAsync(&ft, cells.tested, sizeof ft, cudaMemcpyDeviceToHost, stream));
      PSIM_CK(cudaStreamSynchronize(stream));
      out->filter_tests = (int64_t)ft;
    }
    if (h.comm_timeout) { g_last_error = "a peer's positions did not arrive within 30 s (peer-store exchange)"; return PSIM_ERR_COMM; }
    out->steps = steps;
    out->rebuilds = (int64_t)h.rebuilds;
    out->kernel_launches = launches + cells.launches;
    out->force_launches = force_launches;
    out->force_ms = force_ms;
    out->force_timed = force_timed;
    out->peer_push_steps = peer_push_steps;
    out->n_active = (int64_t)h.n_active;
    return PSIM_OK;
  }

  psim_status set_option(int opt, int64_t value) override {
    switch (opt) {
      case PSIM_OPT_FORCE_ALGO:
        if (value != PSIM_FORCE_ALL_PAIRS && value != PSIM_FORCE_CELL_BINNED) return PSIM_ERR_INVALID_ARGUMENT;
        algo = (int)value; cells.invalidate(); return PSIM_OK;
      case PSIM_OPT_FUSED_STEP: fused = value != 0; return PSIM_OK;
      case PSIM_OPT_KERNEL_TIMING: timing = value != 0; return PSIM_OK;
      case PSIM_OPT_ALLPAIRS_KERNEL:
        // 3 and 6 selected the round-1 tensor-core kernels: both are today's default
        if (value == 3 || value == 6) value = PSIM_ALLPAIRS_TENSOR;
        if (value != PSIM_ALLPAIRS_TENSOR && value != PSIM_ALLPAIRS_DIFFERENCE_FORM && value != PSIM_ALLPAIRS_CUDA_CORE)
          return PSIM_ERR_INVALID_ARGUMENT;
        variant = (int)value; return PSIM_OK;
      default: return PSIM_ERR_INVALID_ARGUMENT;
    }
  }

  psim_status ensure_slot(Slot& s) {
    if (s.d_pos) return PSIM_OK;
    PSIM_CK(cudaMalloc(&s.d_pos, sizeof(R) * 3 * n)); PSIM_CK(cudaMalloc(&s.d_vel, sizeof(R) * 3 * n));
    PSIM_CK(cudaMalloc(&s.d_act, (size_t)n));
    PSIM_CK(cudaMallocHost(&s.h_pos, sizeof(R) * 3 * n)); PSIM_CK(cudaMallocHost(&s.h_vel, sizeof(R) * 3 * n));
    PSIM_CK(cudaMallocHost(&s.h_act, (size_t)n));
    PSIM_CK(cudaEventCreateWithFlags(&s.ready, cudaEventDisableTiming));
    PSIM_CK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    return PSIM_OK;
  }

  // Pack the post-kick state into the slot's device staging on the stepping
  // stream (cheap, O(N)), then drain it to pinned memory on the side stream so
  // the time loop never waits for PCIe.
  psim_status snapshot_async(int slot) override {
    PSIM_CK(cudaSetDevice(device));
    if (slot < 0 || slot >= PSIM_FRAME_SLOTS) return PSIM_ERR_INVALID_ARGUMENT;
    Slot& s = slots[slot];
    psim_status st = ensure_slot(s);
    if (st != PSIM_OK) return st;
    if (s.used) PSIM_CK(cudaStreamWaitEvent(stream, s.done, 0));   // previous drain of this slot
    // while a fused kernel has already drifted, pos[cur] still holds x_n
    pack3_kernel<R, TS><<<blocks(n, TS), TS, 0, stream>>>(s.d_pos, pos[cur], 0, n);
    if (i_hi > i_lo) pack3_kernel<R, TS><<<blocks(i_hi - i_lo, TS), TS, 0, stream>>>(s.d_vel, vel, i_lo, i_hi);
    pack_active_kernel<R, TS><<<blocks(n, TS), TS, 0, stream>>>(s.d_act, pos[cur], n, c, ctrl);
    launches += 3;
    PSIM_CK(cudaEventRecord(s.ready, stream));
    PSIM_CK(cudaStreamWaitEvent(side, s.ready, 0));
    PSIM_CK(cudaMemcpyAsync(s.h_pos, s.d_pos, sizeof(R) * 3 * n, cudaMemcpyDeviceToHost, side));
    if (i_hi > i_lo)
      PSIM_CK(cudaMemcpyAsync(s.h_vel + 3 * i_lo, s.d_vel + 3 * i_lo, sizeof(R) * 3 * (i_hi - i_lo),
                              cudaMemcpyDeviceToHost, side));
    PSIM_CK(cudaMemcpyAsync(s.h_act, s.d_act, (size_t)n, cudaMemcpyDeviceToHost, side));
    PSIM_CK(cudaEventRecord(s.done, side));
    s.used = true;
    return PSIM_OK;
  }

  psim_status snapshot_wait(int slot, const int32_t** id, const void** pos_h, const void** vel_h,
                            const uint8_t** active_h) override {
    PSIM_CK(cudaSetDevice(device));
    if (slot < 0 || slot >= PSIM_FRAME_SLOTS || !slots[slot].used) return PSIM_ERR_INVALID_ARGUMENT;
    Slot& s = slots[slot];
    PSIM_CK(cudaEventSynchronize(s.done));
    if (id) *id = ids.data();
    if (pos_h) *pos_h = s.h_pos;
    if (vel_h) *vel_h = s.h_vel;
    if (active_h) *active_h = s.h_act;
    return PSIM_OK;
  }

  psim_status count_pairs(int64_t* pairs) override {
    PSIM_CK(cudaSetDevice(device));
    unsigned long long* d = reinterpret_cast<unsigned long long*>(red_scratch);
    PSIM_CK(cudaMemsetAsync(d, 0, sizeof *d, stream));
    if (i_hi > i_lo) {
      count_pairs_kernel<R, 128><<<blocks(i_hi - i_lo, 128), 128, 0, stream>>>(pos[cur], n_pad, i_lo, i_hi, c, d);
      ++launches;
    }
    unsigned long long h = 0;
    PSIM_CK(cudaMemcpyAsync(&h, d, sizeof h, cudaMemcpyDeviceToHost, stream));
    PSIM_CK(cudaStreamSynchronize(stream));
    if (pairs) *pairs = (int64_t)h;
    return PSIM_OK;
  }

  psim_status energy(double* k, double* u, double* mom) override {
    PSIM_CK(cudaSetDevice(device));
    double* d = red_scratch;
    PSIM_CK(cudaMemsetAsync(d, 0, 5 * sizeof(double), stream));
    if (i_hi > i_lo) {
      energy_kernel<R, 128><<<blocks(i_hi - i_lo, 128), 128, 0, stream>>>(pos[cur], vel, n_pad, i_lo, i_hi, c, mass, d);
      ++launches;
    }
    double h[5];
    PSIM_CK(cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, stream));
    PSIM_CK(cudaStreamSynchronize(stream));
    if (k) *k = h[0];
    if (u) *u = h[1];
    if (mom) { mom[0] = h[2]; mom[1] = h[3]; mom[2] = h[4]; }
    return PSIM_OK;
  }

  psim_status get_shard(int64_t* lo, int64_t* hi) override {
    if (lo) *lo = i_lo;
    if (hi) *hi = i_hi;
    return PSIM_OK;
  }
};

template <class R>
static psim_status create_t(const psim_params* p, int64_t n, const int32_t* id, const void* pos, const void* vel,
                            const uint8_t* active, int device, int rank, int world, const void* uid,
                            psim_sim** out) {
  SimT<R>* s = new SimT<R>();
  psim_status st = s->init(*p, n, id, pos, vel, active, device, rank, world, uid);
  if (st != PSIM_OK) { delete s; return st; }
  *out = s;
  return PSIM_OK;
}

static psim_status create_any(const psim_params* p, int precision, int64_t n, const int32_t* id, const void* pos,
                              const void* vel, const uint8_t* active, int device, int rank, int world,
                              const void* uid, psim_sim** out) {
  if (!p || !out || !pos) return PSIM_ERR_INVALID_ARGUMENT;
  if (precision != 4 && precision != 8) return PSIM_ERR_INVALID_ARGUMENT;
  if (n <= 0) return PSIM_ERR_NO_PARTICLES;
  if (world < 1 || rank < 0 || rank >= world) return PSIM_ERR_INVALID_ARGUMENT;
  if (world > 1 && !uid) return PSIM_ERR_INVALID_ARGUMENT;
  *out = nullptr;
  std::vector<int32_t> seq;
  if (!id) { seq.resize(n); for (int64_t i = 0; i < n; ++i) seq[i] = (int32_t)(i + 1); id = seq.data(); }
  if (precision == 4) return create_t<float>(p, n, id, pos, vel, active, device, rank, world, uid, out);
  return create_t<double>(p, n, id, pos, vel, active, device, rank, world, uid, out);
}

template <class R>
static psim_status fma_peak_t(int device, double* tflops, double* sm_mhz) {
  PSIM_CK(cudaSetDevice(device));
  cudaDeviceProp prop;
  PSIM_CK(cudaGetDeviceProperties(&prop, device));
  constexpr int ILP = 8;
  const int grid = prop.multiProcessorCount * 8, block = 256, iters = 4096;
  R* out = nullptr; long long* cyc = nullptr;
  PSIM_CK(cudaMalloc(&out, sizeof(R) * (size_t)grid * block));
  PSIM_CK(cudaMalloc(&cyc, sizeof(long long) * grid));
  cudaEvent_t a, b;
  PSIM_CK(cudaEventCreate(&a)); PSIM_CK(cudaEventCreate(&b));
  double best = 0, best_mhz = 0;
  for (int rep = 0; rep < 8; ++rep) {
    PSIM_CK(cudaEventRecord(a, 0));
    fma_peak_kernel<R, ILP><<<grid, block>>>(out, iters, (R)1.0000001, (R)1e-7, cyc);
    PSIM_CK(cudaEventRecord(b, 0));
    PSIM_CK(cudaEventSynchronize(b));
    float ms = 0;
    PSIM_CK(cudaEventElapsedTime(&ms, a, b));
    const double flops = 2.0 * ILP * (double)iters * grid * block;
    const double tf = flops / (ms * 1e-3) / 1e12;
    if (rep >= 2 && tf > best) {
      best = tf;
      // each SM runs 8 blocks of this grid concurrently -> per-block cycle count ~ kernel duration
      long long h = 0;
      PSIM_CK(cudaMemcpy(&h, cyc, sizeof h, cudaMemcpyDeviceToHost));
      best_mhz = (double)h / (ms * 1e-3) / 1e6;
    }
  }
  cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(out); cudaFree(cyc);
  PSIM_CK(cudaGetLastError());
  if (tflops) *tflops = best;
  if (sm_mhz) *sm_mhz = best_mhz;
  return PSIM_OK;
}

// Dense tcgen05.mma rate: one thread per SM issues `iters` M128 N256 instructions back to back into two
// alternating TMEM accumulators (operands: zeros in the canonical no-swizzle K-major layout, 32 B rows) and
// commits once at the end.  KIND 0: kind::tf32, K = 8; 1: kind::f16, K = 16.
template <int KIND>
__global__ void __launch_bounds__(32) tensor_peak_kernel(long long* cycles, int iters) {
  constexpr int MM = 128, NN = 256;
  __shared__ __align__(128) unsigned char sA[MM * 32];
  __shared__ __align__(128) unsigned char sB[NN * 32];
  __shared__ __align__(8) unsigned long long mbar;
  __shared__ unsigned tmem_base;
  const int tid = threadIdx.x;
  for (int e = tid; e < MM * 32 / 4; e += 32) reinterpret_cast<unsigned*>(sA)[e] = 0u;
  for (int e = tid; e < NN * 32 / 4; e += 32) reinterpret_cast<unsigned*>(sB)[e] = 0u;
  if (tid == 0) {
    tc::mbar_init(tc::smem_u32(&mbar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(tc::smem_u32(&tmem_base)));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const unsigned tm = tmem_base;
  if (tid == 0) {
    const unsigned fmt = KIND == 0 ? ((2u << 7) | (2u << 10)) : 0u;     // a/b format: TF32 = 2, F16 = 0; D = F32
    const unsigned idesc = (1u << 4) | fmt | ((unsigned)(NN >> 3) << 17) | ((unsigned)(MM >> 4) << 24);
    const unsigned long long da = tc::make_desc(tc::smem_u32(sA)), db = tc::make_desc(tc::smem_u32(sB));
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
    tc::umma_commit(tc::smem_u32(&mbar));
    tc::mbar_wait(tc::smem_u32(&mbar), 0u);
    cycles[blockIdx.x] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tm));
}

static psim_status tensor_peak(int kind, int device, double* tflops, double* sm_mhz) {
  PSIM_CK(cudaSetDevice(device));
  int nsm = 0;
  PSIM_CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device));
  struct Res {                                      // released on every return path
    long long* cyc = nullptr; cudaEvent_t a = nullptr, b = nullptr;
    ~Res() { cudaFree(cyc); if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
  } r;
  PSIM_CK(cudaMalloc(&r.cyc, sizeof(long long) * nsm));
  PSIM_CK(cudaEventCreate(&r.a)); PSIM_CK(cudaEventCreate(&r.b));
  const int iters = 20000;
  double best = 0, best_mhz = 0;
  for (int rep = 0; rep < 4; ++rep) {
    PSIM_CK(cudaEventRecord(r.a, 0));
    if (kind == 0) tensor_peak_kernel<0><<<nsm, 32>>>(r.cyc, iters);
    else tensor_peak_kernel<1><<<nsm, 32>>>(r.cyc, iters);
    PSIM_CK(cudaEventRecord(r.b, 0));
    PSIM_CK(cudaEventSynchronize(r.b));
    float ms = 0;
    PSIM_CK(cudaEventElapsedTime(&ms, r.a, r.b));
    const double flops = 2.0 * 128 * 256 * (kind == 0 ? 8 : 16) * (double)iters * nsm;
    const double tf = flops / (ms * 1e-3) / 1e12;
    if (rep >= 1 && tf > best) {
      best = tf;
      long long h = 0;
      PSIM_CK(cudaMemcpy(&h, r.cyc, sizeof h, cudaMemcpyDeviceToHost));
      best_mhz = (double)h / (ms * 1e-3) / 1e6;
    }
  }
  PSIM_CK(cudaGetLastError());
  if (tflops) *tflops = best;
  if (sm_mhz) *sm_mhz = best_mhz;
  return PSIM_OK;
}

}  // namespace psim

// ------------------------------ extern "C" --------------------------------
extern "C" {

psim_status psim_create(const psim_params* params, int precision, int64_t n, const int32_t* id, const void* pos,
                        const void* vel, int device, psim_sim** out) {
  return create_any(params, precision, n, id, pos, vel, nullptr, device, 0, 1, nullptr, out);
}

psim_status psim_create_sharded(const psim_params* params, int precision, int64_t n, const int32_t* id,
                                const void* pos, const void* vel, int device, int rank, int world,
                                const void* nccl_unique_id, psim_sim** out) {
  return create_any(params, precision, n, id, pos, vel, nullptr, device, rank, world, nccl_unique_id, out);
}

psim_status psim_comm_unique_id(void* out128) {
  if (!out128) return PSIM_ERR_INVALID_ARGUMENT;
  if (!g_nccl.load()) return PSIM_ERR_COMM;
  ncclUniqueId id;
  PSIM_NCCL(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  std::memcpy(out128, &id, 128);
  return PSIM_OK;
}

psim_status psim_destroy(psim_sim* sim) { delete sim; return PSIM_OK; }

#define PSIM_FWD(sim, expr) ((sim) ? (sim)->expr : PSIM_ERR_INVALID_ARGUMENT)

psim_status psim_set_option(psim_sim* sim, int option, int64_t value) { return PSIM_FWD(sim, set_option(option, value)); }
psim_status psim_advance(psim_sim* sim, int64_t nsteps) { return PSIM_FWD(sim, advance(nsteps)); }
psim_status psim_synchronize(psim_sim* sim) { return PSIM_FWD(sim, synchronize()); }
psim_status psim_advance_timed(psim_sim* sim, int64_t nsteps, double* ms) { return PSIM_FWD(sim, advance_timed(nsteps, ms)); }
psim_status psim_get_frame(psim_sim* sim, int32_t* id, void* pos, void* vel, uint8_t* active) {
  return PSIM_FWD(sim, get_frame(id, pos, vel, active));
}
psim_status psim_get_aux(psim_sim* sim, void* d, void* fo, void* fn) { return PSIM_FWD(sim, get_aux(d, fo, fn)); }
psim_status psim_get_counters(psim_sim* sim, psim_counters* out) { return PSIM_FWD(sim, get_counters(out)); }
psim_status psim_get_shard(psim_sim* sim, int64_t* lo, int64_t* hi) { return PSIM_FWD(sim, get_shard(lo, hi)); }
psim_status psim_set_positions(psim_sim* sim, const void* pos, const uint8_t* active) {
  return PSIM_FWD(sim, set_positions(pos, active));
}
psim_status psim_snapshot_async(psim_sim* sim, int slot) { return PSIM_FWD(sim, snapshot_async(slot)); }
psim_status psim_snapshot_wait(psim_sim* sim, int slot, const int32_t** id, const void** pos, const void** vel,
                               const uint8_t** active) {
  return PSIM_FWD(sim, snapshot_wait(slot, id, pos, vel, active));
}
psim_status psim_count_pairs_in_range(psim_sim* sim, int64_t* pairs) { return PSIM_FWD(sim, count_pairs(pairs)); }
psim_status psim_energy(psim_sim* sim, double* k, double* u, double mom[3]) { return PSIM_FWD(sim, energy(k, u, mom)); }

psim_status psim_compute_forces_host(const psim_params* params, int precision, int64_t n, const void* pos,
                                     const uint8_t* active, int algo, int device, void* force,
                                     int64_t* pairs_in_range) {
  if (!force) return PSIM_ERR_INVALID_ARGUMENT;
  psim_sim* s = nullptr;
  std::vector<uint8_t> all;
  if (!active) { all.assign((size_t)std::max<int64_t>(n, 0), 1); active = all.data(); }
  psim_status st = create_any(params, precision, n, nullptr, pos, nullptr, active, device, 0, 1, nullptr, &s);
  if (st != PSIM_OK) return st;
  if (algo != PSIM_FORCE_ALL_PAIRS) {
    st = s->set_option(PSIM_OPT_FORCE_ALGO, algo);
    if (st == PSIM_OK) st = s->set_positions(pos, active);
  }
  if (st == PSIM_OK) st = s->get_aux(nullptr, nullptr, force);
  if (st == PSIM_OK && pairs_in_range) st = s->count_pairs(pairs_in_range);
  delete s;
  return st;
}

psim_status psim_measure_fma_peak(int precision, int device, double* tflops, double* sm_mhz) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return PSIM_ERR_NO_DEVICE; }
  if (device < 0) device = 0;
  if (precision == 4) return fma_peak_t<float>(device, tflops, sm_mhz);
  if (precision == 8) return fma_peak_t<double>(device, tflops, sm_mhz);
  return PSIM_ERR_INVALID_ARGUMENT;
}

psim_status psim_measure_tensor_peak(int kind, int device, double* tflops, double* sm_mhz) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return PSIM_ERR_NO_DEVICE; }
  if (device < 0) device = 0;
  if (kind != 0 && kind != 1) return PSIM_ERR_INVALID_ARGUMENT;
  return tensor_peak(kind, device, tflops, sm_mhz);
}

const char* psim_status_string(int status) {
  switch (status) {
    case PSIM_OK: return "ok";
    case PSIM_ERR_INVALID_ARGUMENT: return "invalid argument";
    case PSIM_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
    case PSIM_ERR_CUDA: return "CUDA error";
    case PSIM_ERR_NO_PARTICLES: return "No particles in particle file";
    case PSIM_ERR_UNSUPPORTED: return "unsupported configuration";
    case PSIM_ERR_COMM: return "communicator (NCCL) error";
    case PSIM_ERR_IO: return "I/O error";
    default: return "unknown status";
  }
}

const char* psim_last_error(void) { return g_last_error.c_str(); }
const char* psim_version(void) { return "particle_simulation_b200 0.1 (sm_100a)"; }

}  // extern "C"
