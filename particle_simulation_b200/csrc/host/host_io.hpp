// host_io.hpp — the reference's three file formats for the C++ host.
//   parameter file : namelist ENVIRONMENTPARAMS (program_particle_simulation_openmp.f08:198-209)
//   particle file  : list-directed `id, location(3), velocity(3)` records (:252-267)
//   output file    : direct-access records of 4 (or 8) bytes (:54-56, 167-170, 327-365)
#pragma once

#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "psim.h"

namespace psim_host {

struct Namelist {
  psim_params params{};
  int write_time_step = 0;
  int max_time_step = 0;
  bool has[10] = {false};   // which of the ten keys were present
};

// Throws std::runtime_error with a gfortran-like message on malformed input.
Namelist read_namelist(const std::string& path);
Namelist parse_namelist(const std::string& text);

struct Particles {
  std::vector<int32_t> id;
  std::vector<double> pos, vel;   // packed xyz
  size_t size() const { return id.size(); }
};
Particles read_particles(const std::string& path);

// The `.output` stream.  nbytes = record length = sizeof(real(rt)).
class FrameWriter {
 public:
  FrameWriter(const std::string& path, const double area[3], int nbytes);
  ~FrameWriter();
  // `pos`/`vel` hold nbytes-wide reals, packed xyz; inactive particles are skipped (:342).
  void write_frame(int64_t n, const int32_t* id, const void* pos, const void* vel, const uint8_t* active);
  int frames() const { return frames_; }
  void close();

 private:
  void put_int(int32_t v);
  FILE* f_ = nullptr;
  int nbytes_;
  int frames_ = 0;
  std::vector<unsigned char> buf_;
};

}  // namespace psim_host
