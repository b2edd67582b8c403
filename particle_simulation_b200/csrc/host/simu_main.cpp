// simu_main.cpp — the reference's command line over the B200 library:
//
//   simu_b200 <parameter-file> <particle-file> <output-file> [--precision 4|8]
//             [--algo cells|allpairs] [--device N] [--quiet] [--sync-io]
//
// Follows program_particle_simulation_openmp.f08: exactly three positional
// arguments (:146-154), namelist and particle file as read there (:180-286),
// records 1..3 = length_of_area (:54-56), then the time loop (:80-117) with a
// frame every WRITE_TIME_STEP steps (:109-115).  Errors the reference reports
// with stop_program (message, then a bare `stop`, i.e. exit status 0,
// module_auxiliary_routines.f08:24-49) behave the same way here; failures of
// the GPU library itself exit with status 2.
//
// Frames leave through a three-stage pipeline so that write_data (:327-365) never
// holds up the stepping: the GPU keeps stepping, a side stream drains snapshot k
// into one of PSIM_FRAME_SLOTS pinned slots, and a writer thread filters inactive
// particles and appends frame k-1, k-2, ... to the file (--sync-io: no thread).
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "host_io.hpp"
#include "psim.h"

using namespace psim_host;

static bool g_quiet = false;
static void info(const std::string& s) { if (!g_quiet) std::printf(" [INFO] %s\n", s.c_str()); }

[[noreturn]] static void stop_program(const std::string& msg) {   // aux:24-49
  std::fprintf(stderr, " %s\n", msg.c_str());
  std::exit(0);
}

static bool file_exists(const std::string& p) { std::ifstream f(p); return (bool)f; }

static void gpu_check(psim_status st, const char* where) {
  if (st == PSIM_OK) return;
  std::fprintf(stderr, " [ERROR] %s: %s (%s)\n", where, psim_status_string(st), psim_last_error());
  std::exit(2);
}

// Writer thread of the frame pipeline.  The stepping thread hands over (slot, step) after
// psim_snapshot_async; the writer blocks in psim_snapshot_wait (the only library call it makes: it
// reads that slot's event and pinned buffers, nothing the stepping thread touches), writes the frame
// and frees the slot.  Frames are written in hand-over order; a slot is reused only after the frame
// that occupied it is on its way to the file (acquire()).
class FramePipe {
 public:
  FramePipe(psim_sim* sim, FrameWriter* out, int64_t n, bool threaded) : sim_(sim), out_(out), n_(n), threaded_(threaded) {
    if (threaded_) th_ = std::thread([this] { run(); });
  }
  void acquire() {                                   // before psim_snapshot_async on the next slot
    if (!threaded_) return;
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [this] { return in_flight_ < PSIM_FRAME_SLOTS; });
  }
  void submit(int slot, int step) {
    if (!threaded_) { write(slot, step); return; }
    { std::lock_guard<std::mutex> lk(m_); q_.push_back({slot, step}); ++in_flight_; }
    cv_.notify_all();
  }
  void finish() {
    if (!threaded_) return;
    { std::lock_guard<std::mutex> lk(m_); stop_ = true; }
    cv_.notify_all();
    if (th_.joinable()) th_.join();
  }

 private:
  struct Job { int slot, step; };
  void write(int slot, int step) {
    const int32_t* id; const void *p, *v; const uint8_t* act;
    gpu_check(psim_snapshot_wait(sim_, slot, &id, &p, &v, &act), "psim_snapshot_wait");
    info("write for time step: " + std::to_string(step));
    out_->write_frame(n_, id, p, v, act);
  }
  void run() {
    for (;;) {
      Job j;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
        if (q_.empty()) return;
        j = q_.front(); q_.pop_front();
      }
      write(j.slot, j.step);
      { std::lock_guard<std::mutex> lk(m_); --in_flight_; }
      cv_.notify_all();
    }
  }
  psim_sim* sim_; FrameWriter* out_; int64_t n_; bool threaded_;
  std::thread th_; std::mutex m_; std::condition_variable cv_; std::deque<Job> q_;
  int in_flight_ = 0; bool stop_ = false;
};

int main(int argc, char** argv) {
  std::vector<std::string> positional;
  int precision = 4;          // the reference computes in real32 (module_stoermer_verlet_method.f08:5)
  int device = -1;
  std::string algo = "cells";
  bool sync_io = false;       // write frames from the stepping thread (no writer thread)
  bool io_selftest = false;   // parse both inputs, write the INITIAL state as one frame, no GPU (host-logic tests)
  for (int a = 1; a < argc; ++a) {
    std::string s = argv[a];
    auto next = [&]() -> std::string { if (a + 1 >= argc) stop_program("missing value after " + s); return argv[++a]; };
    if (s == "--precision") precision = std::atoi(next().c_str());
    else if (s == "--device") device = std::atoi(next().c_str());
    else if (s == "--algo") algo = next();
    else if (s == "--quiet") g_quiet = true;
    else if (s == "--io-selftest") io_selftest = true;
    else if (s == "--sync-io") sync_io = true;
    else positional.push_back(s);
  }
  info("get command arguments");
  if (positional.size() != 3) {
    std::fprintf(stderr, " %s\n", positional.size() < 3 ? "Not enough arguments!" : "To many arguments!");
    stop_program(std::string("Usage is: ") + argv[0] + " <parameter-file> <particle-file> <output-file>");
  }
  if (precision != 4 && precision != 8) stop_program("--precision must be 4 or 8");
  for (int k = 0; k < 2; ++k)
    if (!file_exists(positional[k])) stop_program("File '" + positional[k] + "' does not exist.");

  info("get and set simulation parameters");
  Namelist nl;
  try { nl = read_namelist(positional[0]); }
  catch (const std::exception& e) {
    stop_program("READ failed for parameter-file '" + positional[0] + "' with message: " + e.what());
  }

  info("initialze particle array");
  Particles P;
  try { P = read_particles(positional[1]); }
  catch (const std::exception& e) { stop_program(e.what()); }
  if (P.size() == 0) stop_program("No particles in particle file");
  info("Number of particles in file: " + std::to_string(P.size()));

  info("write size of simulation arae to output file");
  FrameWriter* out = nullptr;
  try { out = new FrameWriter(positional[2], nl.params.length_of_area, precision); }
  catch (const std::exception& e) { stop_program(e.what()); }

  const int64_t n = (int64_t)P.size();
  std::vector<float> pos32, vel32;
  const void *pos = P.pos.data(), *vel = P.vel.data();
  if (precision == 4) {
    pos32.assign(P.pos.begin(), P.pos.end());
    vel32.assign(P.vel.begin(), P.vel.end());
    pos = pos32.data(); vel = vel32.data();
  }
  if (io_selftest) {
    std::vector<uint8_t> all(n, 1);
    out->write_frame(n, P.id.data(), pos, vel, all.data());
    out->close();
    std::printf("%d %d %d %.17g %.17g\n", (int)nl.params.nb_of_threads, nl.write_time_step, nl.max_time_step,
                nl.params.delta_t, nl.params.r_cut_factor);
    return 0;
  }
  psim_sim* sim = nullptr;
  gpu_check(psim_create(&nl.params, precision, n, P.id.data(), pos, vel, device, &sim), "psim_create");
  gpu_check(psim_set_option(sim, PSIM_OPT_FORCE_ALGO, algo == "allpairs" ? PSIM_FORCE_ALL_PAIRS : PSIM_FORCE_CELL_BINNED),
            "psim_set_option");

  info("start time integration");
  const auto t0 = std::chrono::steady_clock::now();
  const int wts = nl.write_time_step > 0 ? nl.write_time_step : nl.max_time_step + 1;
  FramePipe pipe(sim, out, n, !sync_io);
  int pending_slot = -1, pending_step = 0, slot = 0;   // --sync-io: write frame k-1 after enqueueing chunk k+1's snapshot
  int step = 0;
  while (step < nl.max_time_step) {
    const int chunk = std::min(wts - step % wts, nl.max_time_step - step);
    gpu_check(psim_advance(sim, chunk), "psim_advance");
    step += chunk;
    if (step % wts == 0) {
      pipe.acquire();                                                     // the frame that used this slot is written
      gpu_check(psim_snapshot_async(sim, slot), "psim_snapshot_async");   // drains on a side stream
      if (sync_io) {
        if (pending_slot >= 0) pipe.submit(pending_slot, pending_step);   // write the previous frame meanwhile
        pending_slot = slot; pending_step = step;
      } else {
        pipe.submit(slot, step);
      }
      slot = (slot + 1) % PSIM_FRAME_SLOTS;
    }
  }
  if (sync_io && pending_slot >= 0) pipe.submit(pending_slot, pending_step);
  pipe.finish();
  gpu_check(psim_synchronize(sim), "psim_synchronize");
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (!g_quiet) std::printf(" elapsed time in seconds: %.3f\n", secs);
  psim_counters c;
  gpu_check(psim_get_counters(sim, &c), "psim_get_counters");
  if (!g_quiet)
    std::printf(" [INFO] %lld steps, %lld rebuilds, %lld kernel launches, %lld active particles, %d frames\n",
                (long long)c.steps, (long long)c.rebuilds, (long long)c.kernel_launches, (long long)c.n_active,
                out->frames());
  info("clean up");
  psim_destroy(sim);
  out->close();
  delete out;
  return 0;
}
