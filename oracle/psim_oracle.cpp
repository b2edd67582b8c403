// psim_oracle.cpp — CPU restatement of the reference's hot path and of the
// neighbour management that feeds it.  See psim_oracle.h for the rules of use
// (test infrastructure only; parity unpinned) and the citation shorthand.
//
// Build: see oracle/Makefile.  The deterministic oracle is built with
// -ffp-contract=off so that the float instantiation has one well-defined
// rounding per operation; the timed baseline is the same source at
// -O3 -march=native -fopenmp (optionally -ffast-math, mirroring Makefile:5 of
// the reference, `-Ofast`).

#include "psim_oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

template <class R>
struct Constants {
  R delta_t, epsilon, r_min, r_cut, min_cell, dr_max_tolerable, F_rcut, tmf;
  R area[3], cell[3];
};

// stoermer:68-81, same expression order, all in R.
template <class R>
Constants<R> derive(const psim_params& p) {
  Constants<R> c;
  const R sigma = (R)p.sigma, mass = (R)p.mass, rcf = (R)p.r_cut_factor;
  c.delta_t = (R)p.delta_t;
  c.epsilon = (R)p.epsilon;
  c.r_min = std::pow((R)2, (R)1 / (R)6) * sigma;        // :70
  c.r_cut = rcf * sigma;                                // :71
  for (int d = 0; d < 3; ++d) {
    c.area[d] = (R)p.length_of_area[d];
    c.cell[d] = (R)p.length_of_cell[d];
  }
  c.min_cell = std::min(c.cell[0], std::min(c.cell[1], c.cell[2]));  // :74
  c.dr_max_tolerable = c.min_cell - c.r_cut;            // :75
  R q = c.r_min / c.r_cut;
  R q2 = q * q;
  R s = q2 * q2 * q2;                                   // :78  (..)**6
  c.F_rcut = (s / (c.r_cut * c.r_cut)) * ((R)1 - s);    // :79
  c.tmf = (R)0.5 * c.delta_t / mass;                    // :81
  return c;
}

template <class R>
inline R norm3(R a, R b, R c) {          // norm2() of a 3-vector (stoermer:119,165; pm:382)
  return std::sqrt(a * a + b * b + c * c);
}

// The scalar multiplying r_ij in `force` of stoermer:126, or 0 outside the cutoff.
template <class R>
inline bool pair_scalar(const Constants<R>& c, R r, R& g) {
  if (!(r < c.r_cut)) return false;                     // :121 strict <
  R q = c.r_min / r;
  R q2 = q * q;
  R s = q2 * q2 * q2;                                   // :123
  g = (s / (r * r)) * ((R)1 - s);                       // :124  F_rij
  return true;
}

struct SimBase {
  virtual ~SimBase() {}
  virtual void advance(int64_t nsteps) = 0;
  virtual void get_frame(int32_t*, void*, void*, uint8_t*) = 0;
  virtual void get_aux(void*, void*, void*) = 0;
  virtual void energy(double*, double*, double*) = 0;
  int64_t rebuilds = 0, steps = 0;
  double loop_seconds = 0.0;
  virtual int64_t n_active() = 0;
  virtual int64_t list_entries() = 0;
};

template <class R>
struct Sim : SimBase {
  Constants<R> c;
  int64_t n;
  int threads, slices;
  std::vector<int32_t> id;
  std::vector<R> x, disp, v, f_old, f_new;      // packed xyz
  std::vector<uint8_t> active;
  // particle_type%neighbor_list / nb_of_neighbors (stoermer:25-26)
  std::vector<std::vector<int32_t>> nbr;
  std::vector<int32_t> nb;
  // cell_space (pm:14-23): singly linked lists, newest entry first.
  int nc[3];
  std::vector<int32_t> head;   // first_number per cell, -1 = null
  std::vector<int32_t> next;   // next_number per particle
  std::vector<uint8_t> in_cell;

  Sim(const psim_params& p, int64_t n_, const int32_t* id_, const R* pos, const R* vel, int threads_)
      : c(derive<R>(p)), n(n_), threads(std::max(1, threads_)),
        slices(std::max<int>(1, p.nb_of_threads)) {
    id.assign(id_, id_ + n);
    x.assign(pos, pos + 3 * n);
    v.assign(vel, vel + 3 * n);
    disp.assign(3 * n, (R)0);
    f_old.assign(3 * n, (R)0);
    f_new.assign(3 * n, (R)0);
    active.assign(n, 1);
    nbr.resize(n);
    nb.assign(n, 0);
    next.assign(n, -1);
    in_cell.assign(n, 0);
    init_cell_space();        // pm:38
    update_neighbor_list();   // pm:43
    update_active();          // pm:46
    compute_forces();         // program:77
  }

  inline int cell_index(int cx, int cy, int cz) const {   // 1-based in, linear out
    return (cx - 1) + nc[0] * ((cy - 1) + nc[1] * (cz - 1));
  }

  // ceiling(location / length_of_cell) (pm:95, 440-441); returns false if the
  // index leaves 1..nc (pm:503).
  inline bool target_cell(int64_t i, int cc[3]) const {
    bool inside = true;
    for (int d = 0; d < 3; ++d) {
      cc[d] = (int)std::ceil(x[3 * i + d] / c.cell[d]);
      if (cc[d] < 1 || cc[d] > nc[d]) inside = false;
    }
    return inside;
  }

  // pm:52-108.  A particle outside [0, area] is never linked (pm:82-92).  The
  // reference then indexes the grid unchecked; a coordinate of exactly 0, or a
  // grid smaller than the box (SURVEY.md 3.3 edge cases) would index out of
  // bounds there — here such a particle is simply left unlinked (inactive).
  void init_cell_space() {
    for (int d = 0; d < 3; ++d) nc[d] = (int)std::lround(c.area[d] / c.cell[d]);   // nint, pm:62
    head.assign((size_t)nc[0] * nc[1] * nc[2], -1);
    for (int64_t i = 0; i < n; ++i) {
      bool outside = false;
      for (int d = 0; d < 3; ++d)
        if (x[3 * i + d] > c.area[d] || x[3 * i + d] < (R)0) outside = true;
      if (outside) continue;
      int cc[3];
      if (!target_cell(i, cc)) continue;
      int ci = cell_index(cc[0], cc[1], cc[2]);
      next[i] = head[ci];       // pm:100-104: new entry becomes the list head
      head[ci] = (int32_t)i;
      in_cell[i] = 1;
    }
  }

  // pm:374-397
  inline void add_neighbor(int32_t p1, int32_t p2) {
    R r = norm3(x[3 * p1] - x[3 * p2], x[3 * p1 + 1] - x[3 * p2 + 1], x[3 * p1 + 2] - x[3 * p2 + 2]);
    if (r <= c.min_cell) {
      int32_t k = nb[p1]++;
      if ((size_t)k >= nbr[p1].size()) nbr[p1].resize(nbr[p1].size() + 10);   // pm:401-412
      nbr[p1][k] = p2;
    }
  }

  // pm:343-370
  inline void add_from_other_cell(int a, int b) {
    if (head[b] < 0) return;
    for (int32_t p1 = head[a]; p1 >= 0; p1 = next[p1])
      for (int32_t p2 = head[b]; p2 >= 0; p2 = next[p2]) add_neighbor(p1, p2);
  }

  // pm:128-137 = within-cell pass (pm:141-175) then the 13 half-stencil cell
  // pairs in the reference's loop order (pm:237-337); that order fixes the
  // order of every neighbour list and hence the summation order of the forces.
  void update_neighbor_list() {
    const int X = nc[0], Y = nc[1], Z = nc[2];
    const size_t ncell = head.size();
    for (size_t ci = 0; ci < ncell; ++ci) {
      for (int32_t p1 = head[ci]; p1 >= 0; p1 = next[p1]) {
        disp[3 * p1] = disp[3 * p1 + 1] = disp[3 * p1 + 2] = (R)0;   // pm:158
        nb[p1] = 0;                                                   // pm:159
        for (int32_t p2 = next[p1]; p2 >= 0; p2 = next[p2]) add_neighbor(p1, p2);
      }
    }
    auto C = [&](int cx, int cy, int cz) { return cell_index(cx, cy, cz); };
    for (int cz = 1; cz <= Z; ++cz) {                      // pm:238-271
      for (int cy = 1; cy <= Y - 1; ++cy)
        for (int cx = 1; cx <= X; ++cx) add_from_other_cell(C(cx, cy, cz), C(cx, cy + 1, cz));
      for (int cy = 1; cy <= Y; ++cy)
        for (int cx = 1; cx <= X - 1; ++cx) add_from_other_cell(C(cx, cy, cz), C(cx + 1, cy, cz));
      for (int cy = 1; cy <= Y - 1; ++cy)
        for (int cx = 1; cx <= X - 1; ++cx) {
          add_from_other_cell(C(cx, cy, cz), C(cx + 1, cy + 1, cz));
          add_from_other_cell(C(cx + 1, cy, cz), C(cx, cy + 1, cz));
        }
    }
    for (int cz = 1; cz <= Z - 1; ++cz) {                  // pm:275-336
      for (int cy = 1; cy <= Y; ++cy)
        for (int cx = 1; cx <= X; ++cx) add_from_other_cell(C(cx, cy, cz), C(cx, cy, cz + 1));
      for (int cy = 1; cy <= Y - 1; ++cy)
        for (int cx = 1; cx <= X; ++cx) {
          add_from_other_cell(C(cx, cy, cz), C(cx, cy + 1, cz + 1));
          add_from_other_cell(C(cx, cy + 1, cz), C(cx, cy, cz + 1));
        }
      for (int cy = 1; cy <= Y; ++cy)
        for (int cx = 1; cx <= X - 1; ++cx) {
          add_from_other_cell(C(cx, cy, cz), C(cx + 1, cy, cz + 1));
          add_from_other_cell(C(cx + 1, cy, cz), C(cx, cy, cz + 1));
        }
      for (int cy = 1; cy <= Y - 1; ++cy)
        for (int cx = 1; cx <= X - 1; ++cx) {
          add_from_other_cell(C(cx, cy, cz), C(cx + 1, cy + 1, cz + 1));
          add_from_other_cell(C(cx + 1, cy + 1, cz), C(cx, cy, cz + 1));
          add_from_other_cell(C(cx + 1, cy, cz), C(cx, cy + 1, cz + 1));
          add_from_other_cell(C(cx, cy + 1, cz), C(cx + 1, cy, cz + 1));
        }
    }
  }

  // pm:416-513: walk every cell's list (z outermost, x innermost), unhook the
  // particles whose cell changed and push them on the head of their new cell,
  // or drop them when the new index is outside the grid (pm:503-504).
  void update_cell_space() {
    for (int cz = 1; cz <= nc[2]; ++cz)
      for (int cy = 1; cy <= nc[1]; ++cy)
        for (int cx = 1; cx <= nc[0]; ++cx) {
          const int ci = cell_index(cx, cy, cz);
          int32_t prev = -1, p = head[ci];
          while (p >= 0) {
            int cc[3];
            bool inside = target_cell(p, cc);
            if (cc[0] == cx && cc[1] == cy && cc[2] == cz) {
              prev = p;
              p = next[p];
              continue;
            }
            int32_t follower = next[p];
            if (prev >= 0) next[prev] = follower; else head[ci] = follower;
            if (inside) {
              int cj = cell_index(cc[0], cc[1], cc[2]);
              next[p] = head[cj];
              head[cj] = p;
            } else {
              next[p] = -1;
              in_cell[p] = 0;       // delete_particle_number, pm:517-530
            }
            p = follower;
          }
        }
  }

  void update_active() {            // pm:534-552
    for (int64_t i = 0; i < n; ++i) active[i] = in_cell[i];
  }

  void rebuild() {                  // pm:112-124
    update_cell_space();
    update_neighbor_list();
    update_active();
    ++rebuilds;
  }

  // stoermer:87-132.  Serial version: the reference's single-thread order.
  void compute_forces() {
    const R k12 = (R)12;
    if (threads == 1) {
      for (int64_t i = 0; i < n; ++i) {
        if (!active[i] || nb[i] <= 0) continue;           // :94
        const int32_t* L = nbr[i].data();
        for (int32_t t = 0; t < nb[i]; ++t) {
          const int32_t j = L[t];
          R rx = x[3 * i] - x[3 * j], ry = x[3 * i + 1] - x[3 * j + 1], rz = x[3 * i + 2] - x[3 * j + 2];
          R r = norm3(rx, ry, rz), F;
          if (!pair_scalar(c, r, F)) continue;
          R g = F - c.F_rcut;
          R fx = rx * k12 * c.epsilon * g;                // :126, left to right
          R fy = ry * k12 * c.epsilon * g;
          R fz = rz * k12 * c.epsilon * g;
          f_new[3 * i] -= fx; f_new[3 * i + 1] -= fy; f_new[3 * i + 2] -= fz;   // :128
          f_new[3 * j] += fx; f_new[3 * j + 1] += fy; f_new[3 * j + 2] += fz;   // :129
        }
      }
      return;
    }
    // Timed multi-thread baseline: contiguous i-slices per thread as
    // program:305-323; the reference's unsynchronised write to particle_j
    // (stoermer:129) is repaired with atomic updates.
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int64_t i = 0; i < n; ++i) {
      if (!active[i] || nb[i] <= 0) continue;
      const int32_t* L = nbr[i].data();
      R ax = 0, ay = 0, az = 0;
      for (int32_t t = 0; t < nb[i]; ++t) {
        const int32_t j = L[t];
        R rx = x[3 * i] - x[3 * j], ry = x[3 * i + 1] - x[3 * j + 1], rz = x[3 * i + 2] - x[3 * j + 2];
        R r = norm3(rx, ry, rz), F;
        if (!pair_scalar(c, r, F)) continue;
        R g = F - c.F_rcut;
        R fx = rx * k12 * c.epsilon * g, fy = ry * k12 * c.epsilon * g, fz = rz * k12 * c.epsilon * g;
        ax -= fx; ay -= fy; az -= fz;
#pragma omp atomic
        f_new[3 * j] += fx;
#pragma omp atomic
        f_new[3 * j + 1] += fy;
#pragma omp atomic
        f_new[3 * j + 2] += fz;
      }
#pragma omp atomic
      f_new[3 * i] += ax;
#pragma omp atomic
      f_new[3 * i + 1] += ay;
#pragma omp atomic
      f_new[3 * i + 2] += az;
    }
  }

  // stoermer:136-187 over one slice [lo, hi); returns the slice's flag.
  bool compute_locations(int64_t lo, int64_t hi) {
    R first = 0, second = 0;
    auto body = [&](int64_t i, R& a, R& b) {
      if (!active[i]) return;                                          // :151
      for (int d = 0; d < 3; ++d) {
        f_old[3 * i + d] = f_new[3 * i + d];                           // :155
        f_new[3 * i + d] = (R)0;                                       // :157
        R dx = c.delta_t * (v[3 * i + d] + c.tmf * f_old[3 * i + d]);  // :160
        x[3 * i + d] = x[3 * i + d] + dx;                              // :161
        disp[3 * i + d] = disp[3 * i + d] + dx;                        // :162
      }
      R dr = norm3(disp[3 * i], disp[3 * i + 1], disp[3 * i + 2]);     // :165
      if (dr > a) { b = a; a = dr; } else if (dr > b) { b = dr; }      // :168-173
    };
    if (threads == 1) {
      for (int64_t i = lo; i < hi; ++i) body(i, first, second);
    } else {
#pragma omp parallel num_threads(threads)
      {
        R a = 0, b = 0;
#pragma omp for schedule(static) nowait
        for (int64_t i = lo; i < hi; ++i) body(i, a, b);
#pragma omp critical
        {
          // merging per-thread top-2 pairs yields the slice's top-2 values
          R vals[2] = {a, b};
          for (R dr : vals) {
            if (dr > first) { second = first; first = dr; } else if (dr > second) { second = dr; }
          }
        }
      }
    }
    return (first + second) > c.dr_max_tolerable;                      // :177-185
  }

  void compute_velocities() {       // stoermer:191-206
#pragma omp parallel for schedule(static) num_threads(threads) if (threads > 1)
    for (int64_t i = 0; i < n; ++i) {
      if (!active[i]) continue;
      for (int d = 0; d < 3; ++d)
        v[3 * i + d] = v[3 * i + d] + c.tmf * (f_new[3 * i + d] + f_old[3 * i + d]);
    }
  }

  // program:305-323 (1-based, inclusive) -> 0-based half-open
  void slice_bounds(int t, int64_t& lo, int64_t& hi) const {
    int64_t per = n / slices;
    lo = per * t;
    hi = (t == slices - 1) ? n : per * (t + 1);
  }

  void advance(int64_t nsteps) override {   // program:80-117
    auto t0 = std::chrono::steady_clock::now();
    for (int64_t s = 0; s < nsteps; ++s) {
      bool any = false;
      for (int t = 0; t < slices; ++t) {
        int64_t lo, hi;
        slice_bounds(t, lo, hi);
        if (compute_locations(lo, hi)) any = true;     // :86
      }
      if (any) rebuild();                              // :92-93
      compute_forces();                                // :103
      compute_velocities();                            // :106
      ++steps;
    }
    loop_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  }

  void get_frame(int32_t* id_, void* pos, void* vel, uint8_t* act) override {
    if (id_) std::memcpy(id_, id.data(), n * sizeof(int32_t));
    if (pos) std::memcpy(pos, x.data(), 3 * n * sizeof(R));
    if (vel) std::memcpy(vel, v.data(), 3 * n * sizeof(R));
    if (act) std::memcpy(act, active.data(), n);
  }
  void get_aux(void* d, void* fo, void* fn) override {
    if (d) std::memcpy(d, disp.data(), 3 * n * sizeof(R));
    if (fo) std::memcpy(fo, f_old.data(), 3 * n * sizeof(R));
    if (fn) std::memcpy(fn, f_new.data(), 3 * n * sizeof(R));
  }
  int64_t n_active() override {
    int64_t k = 0;
    for (int64_t i = 0; i < n; ++i) k += active[i];
    return k;
  }
  int64_t list_entries() override {
    int64_t k = 0;
    for (int64_t i = 0; i < n; ++i) if (active[i]) k += nb[i];
    return k;
  }

  // Energy consistent with the force of stoermer:121-129:
  // U(r) = eps (s^2 - 2 s) - 6 eps F_rcut r^2 + C for r < r_cut, C from U(r_cut) = 0.
  void energy(double* kin, double* pot, double* mom) override {
    const double eps = (double)c.epsilon, rmin = (double)c.r_min, rcut = (double)c.r_cut;
    const double Frc = (double)c.F_rcut, mass = 0.5 * (double)c.delta_t / (double)c.tmf;
    const double sc = std::pow(rmin / rcut, 6.0);
    const double Cc = -(eps * (sc * sc - 2.0 * sc) - 6.0 * eps * Frc * rcut * rcut);
    double K = 0, U = 0, P[3] = {0, 0, 0};
    for (int64_t i = 0; i < n; ++i) {
      if (!active[i]) continue;
      for (int d = 0; d < 3; ++d) {
        double vd = (double)v[3 * i + d];
        K += 0.5 * mass * vd * vd;
        P[d] += mass * vd;
      }
    }
    // Pair sum over the CURRENT positions (the lists may be stale between
    // rebuilds): sweep along x with a cutoff window.
    std::vector<int64_t> order;
    for (int64_t i = 0; i < n; ++i) if (active[i]) order.push_back(i);
    std::sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return x[3 * a] < x[3 * b]; });
    for (size_t a = 0; a < order.size(); ++a) {
      int64_t i = order[a];
      for (size_t b = a + 1; b < order.size(); ++b) {
        int64_t j = order[b];
        double dx = (double)x[3 * j] - (double)x[3 * i];
        if (dx >= rcut) break;
        double dy = (double)x[3 * j + 1] - (double)x[3 * i + 1], dz = (double)x[3 * j + 2] - (double)x[3 * i + 2];
        double r2 = dx * dx + dy * dy + dz * dz;
        if (r2 < rcut * rcut) {
          double s = std::pow(rmin * rmin / r2, 3.0);
          U += eps * (s * s - 2.0 * s) - 6.0 * eps * Frc * r2 + Cc;
        }
      }
    }
    *kin = K; *pot = U;
    mom[0] = P[0]; mom[1] = P[1]; mom[2] = P[2];
  }
};

struct Kahan {
  double s = 0, c = 0;
  inline void add(double v) { double y = v - c, t = s + y; c = (t - s) - y; s = t; }
};

template <class R>
int64_t bruteforce(const psim_params& p, int64_t n, const R* pos, const uint8_t* active, double* out) {
  const double sigma = p.sigma, eps = p.epsilon;
  const double rmin = std::pow(2.0, 1.0 / 6.0) * sigma, rcut = p.r_cut_factor * sigma;
  const double sc = std::pow(rmin / rcut, 6.0);
  const double Frc = (sc / (rcut * rcut)) * (1.0 - sc);
  // sort by x to prune; accumulate per particle with compensation
  std::vector<int64_t> order;
  for (int64_t i = 0; i < n; ++i) if (!active || active[i]) order.push_back(i);
  std::sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return pos[3 * a] < pos[3 * b]; });
  std::vector<Kahan> acc(3 * n);
  int64_t pairs = 0;
  for (size_t a = 0; a < order.size(); ++a) {
    int64_t i = order[a];
    for (size_t b = a + 1; b < order.size(); ++b) {
      int64_t j = order[b];
      double dx = (double)pos[3 * i] - (double)pos[3 * j];
      if (-dx >= rcut) break;
      double dy = (double)pos[3 * i + 1] - (double)pos[3 * j + 1];
      double dz = (double)pos[3 * i + 2] - (double)pos[3 * j + 2];
      double r = std::sqrt(dx * dx + dy * dy + dz * dz);
      if (r < rcut) {
        double s = std::pow(rmin / r, 6.0);
        double g = 12.0 * eps * ((s / (r * r)) * (1.0 - s) - Frc);
        acc[3 * i].add(-dx * g); acc[3 * i + 1].add(-dy * g); acc[3 * i + 2].add(-dz * g);
        acc[3 * j].add(dx * g);  acc[3 * j + 1].add(dy * g);  acc[3 * j + 2].add(dz * g);
        pairs += 2;
      }
    }
  }
  for (int64_t k = 0; k < 3 * n; ++k) out[k] = acc[k].s;
  return pairs;
}

}  // namespace

struct psim_oracle { SimBase* impl; };

extern "C" {

psim_oracle* psim_oracle_create(const psim_params* params, int precision, int64_t n,
                                const int32_t* id, const void* pos, const void* vel, int threads) {
  if (!params || n <= 0 || !id || !pos || !vel) return nullptr;
  psim_oracle* o = new psim_oracle;
  if (precision == 4) o->impl = new Sim<float>(*params, n, id, (const float*)pos, (const float*)vel, threads);
  else if (precision == 8) o->impl = new Sim<double>(*params, n, id, (const double*)pos, (const double*)vel, threads);
  else { delete o; return nullptr; }
  return o;
}
void psim_oracle_destroy(psim_oracle* o) { if (o) { delete o->impl; delete o; } }
void psim_oracle_advance(psim_oracle* o, int64_t nsteps) { o->impl->advance(nsteps); }
void psim_oracle_get_frame(psim_oracle* o, int32_t* id, void* pos, void* vel, uint8_t* active) {
  o->impl->get_frame(id, pos, vel, active);
}
void psim_oracle_get_aux(psim_oracle* o, void* d, void* fo, void* fn) { o->impl->get_aux(d, fo, fn); }
int64_t psim_oracle_rebuilds(psim_oracle* o) { return o->impl->rebuilds; }
int64_t psim_oracle_steps(psim_oracle* o) { return o->impl->steps; }
int64_t psim_oracle_n_active(psim_oracle* o) { return o->impl->n_active(); }
int64_t psim_oracle_list_entries(psim_oracle* o) { return o->impl->list_entries(); }
double psim_oracle_loop_seconds(psim_oracle* o) { return o->impl->loop_seconds; }

void psim_oracle_constants(const psim_params* params, int precision, double out[7]) {
  if (precision == 4) {
    Constants<float> c = derive<float>(*params);
    double v[7] = {c.r_min, c.r_cut, c.min_cell, c.dr_max_tolerable, c.F_rcut, c.tmf, c.delta_t};
    std::memcpy(out, v, sizeof v);
  } else {
    Constants<double> c = derive<double>(*params);
    double v[7] = {c.r_min, c.r_cut, c.min_cell, c.dr_max_tolerable, c.F_rcut, c.tmf, c.delta_t};
    std::memcpy(out, v, sizeof v);
  }
}

double psim_oracle_pair_force_x(const psim_params* params, int precision, double r) {
  // particle i at (r,0,0), j at the origin: returns f_i.x = -(force.x) of stoermer:126-128
  if (precision == 4) {
    Constants<float> c = derive<float>(*params);
    float F;
    if (!pair_scalar(c, (float)r, F)) return 0.0;
    float fx = (float)r * 12.0f * c.epsilon * (F - c.F_rcut);
    return (double)(0.0f - fx);
  }
  Constants<double> c = derive<double>(*params);
  double F;
  if (!pair_scalar(c, r, F)) return 0.0;
  double fx = r * 12.0 * c.epsilon * (F - c.F_rcut);
  return 0.0 - fx;
}

int64_t psim_oracle_bruteforce(const psim_params* params, int precision, int64_t n, const void* pos,
                               const uint8_t* active, double* force_out) {
  if (precision == 4) return bruteforce<float>(*params, n, (const float*)pos, active, force_out);
  return bruteforce<double>(*params, n, (const double*)pos, active, force_out);
}

void psim_oracle_energy(psim_oracle* o, double* kinetic, double* potential, double momentum[3]) {
  o->impl->energy(kinetic, potential, momentum);
}

}  // extern "C"
