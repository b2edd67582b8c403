#include "host_io.hpp"

#include <algorithm>
#include <cctype>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace psim_host {

namespace {

std::string lower(std::string s) {
  std::transform(s.begin(), s.end(), s.begin(), [](unsigned char c) { return (char)std::tolower(c); });
  return s;
}

std::string slurp(const std::string& path) {
  std::ifstream in(path, std::ios::binary);
  if (!in) throw std::runtime_error("Cannot open file '" + path + "': " + std::strerror(errno));
  std::ostringstream ss;
  ss << in.rdbuf();
  return ss.str();
}

// Fortran real literal: 1.0, 1., .5, 1e-3, 1d-3
bool parse_real(std::string t, double& out) {
  for (char& ch : t) if (ch == 'd' || ch == 'D') ch = 'e';
  if (t.empty()) return false;
  char* end = nullptr;
  errno = 0;
  out = std::strtod(t.c_str(), &end);
  return end && *end == '\0' && errno != ERANGE;
}

bool parse_int(const std::string& t, long& out) {
  if (t.empty()) return false;
  char* end = nullptr;
  errno = 0;
  out = std::strtol(t.c_str(), &end, 10);
  return end && *end == '\0' && errno != ERANGE;
}

const char* kKeys[10] = {"nb_of_threads", "sigma", "epsilon", "mass", "r_cut",
                         "length_of_area", "length_of_cell", "delta_t", "write_time_step", "max_time_step"};

}  // namespace

Namelist parse_namelist(const std::string& text_in) {
  // strip `!` comments
  std::string text;
  {
    std::istringstream in(text_in);
    std::string line;
    while (std::getline(in, line)) {
      size_t bang = line.find('!');
      if (bang != std::string::npos) line.erase(bang);
      text += line;
      text += '\n';
    }
  }
  const std::string low = lower(text);
  size_t g = low.find("&environmentparams");
  if (g == std::string::npos) throw std::runtime_error("namelist group &ENVIRONMENTPARAMS not found");
  size_t pos = g + std::strlen("&environmentparams");
  Namelist nl;
  nl.params.nb_of_threads = 1;
  // tokenise: names, '=', values, '/' terminator
  auto skip_sep = [&](size_t& p) { while (p < text.size() && (std::isspace((unsigned char)text[p]) || text[p] == ',')) ++p; };
  int current = -1;
  int filled = 0;
  while (true) {
    skip_sep(pos);
    if (pos >= text.size()) break;                 // no terminator: end of file, which the reference ignores (:210)
    if (text[pos] == '/') break;
    size_t start = pos;
    while (pos < text.size() && !std::isspace((unsigned char)text[pos]) && text[pos] != ',' && text[pos] != '=' && text[pos] != '/') ++pos;
    std::string tok = text.substr(start, pos - start);
    size_t look = pos;
    while (look < text.size() && std::isspace((unsigned char)text[look])) ++look;
    if (look < text.size() && text[look] == '=') {  // a name
      pos = look + 1;
      std::string name = lower(tok);
      size_t paren = name.find('(');
      if (paren != std::string::npos) name.erase(paren);
      current = -1;
      for (int k = 0; k < 10; ++k) if (name == kKeys[k]) current = k;
      if (current < 0) throw std::runtime_error("Cannot match namelist object name " + tok);
      nl.has[current] = true;
      filled = 0;
      continue;
    }
    if (current < 0) throw std::runtime_error("namelist value without a name: " + tok);
    // value, possibly r*c
    int repeat = 1;
    size_t star = tok.find('*');
    if (star != std::string::npos) {
      long r;
      if (!parse_int(tok.substr(0, star), r) || r < 1) throw std::runtime_error("bad repeat count in namelist: " + tok);
      repeat = (int)r;
      tok = tok.substr(star + 1);
    }
    for (int r = 0; r < repeat; ++r) {
      const bool is_int = current == 0 || current == 8 || current == 9;
      const int cap = (current == 5 || current == 6) ? 3 : 1;
      if (filled >= cap) throw std::runtime_error(std::string("too many values for namelist object ") + kKeys[current]);
      if (is_int) {
        long v;
        if (!parse_int(tok, v)) throw std::runtime_error(std::string("Bad integer for item in namelist object ") + kKeys[current]);
        if (current == 0) nl.params.nb_of_threads = (int32_t)v;
        if (current == 8) nl.write_time_step = (int)v;
        if (current == 9) nl.max_time_step = (int)v;
      } else {
        double v;
        if (!parse_real(tok, v)) throw std::runtime_error(std::string("Bad real number in item of namelist object ") + kKeys[current]);
        switch (current) {
          case 1: nl.params.sigma = v; break;
          case 2: nl.params.epsilon = v; break;
          case 3: nl.params.mass = v; break;
          case 4: nl.params.r_cut_factor = v; break;
          case 5: nl.params.length_of_area[filled] = v; break;
          case 6: nl.params.length_of_cell[filled] = v; break;
          case 7: nl.params.delta_t = v; break;
        }
      }
      ++filled;
    }
  }
  return nl;
}

Namelist read_namelist(const std::string& path) { return parse_namelist(slurp(path)); }

// One list-directed READ of 7 items (program_...:252): values are taken from the
// current line and, if it runs short, from the following lines; what is left
// of the last line used is skipped.
Particles read_particles(const std::string& path) {
  std::ifstream in(path);
  if (!in) throw std::runtime_error("Could not open particle file '" + path + "'");
  Particles p;
  std::string line;
  std::vector<std::string> item;
  while (true) {
    item.clear();
    bool eof = false;
    while (item.size() < 7) {
      if (!std::getline(in, line)) { eof = true; break; }
      for (char& ch : line) if (ch == ',') ch = ' ';
      std::istringstream ls(line);
      std::string tok;
      while (item.size() < 7 && (ls >> tok)) item.push_back(tok);
    }
    if (eof) break;      // end of file, also in the middle of a record (istat < 0, :253-254)
    long idv;
    if (!parse_int(item[0], idv)) throw std::runtime_error("Could not read from particle file '" + path + "': Bad integer for item 1 in list input");
    double v[6];
    for (int k = 0; k < 6; ++k)
      if (!parse_real(item[1 + k], v[k])) throw std::runtime_error("Could not read from particle file '" + path + "': Bad real number in item " + std::to_string(k + 2) + " of list input");
    p.id.push_back((int32_t)idv);
    p.pos.insert(p.pos.end(), v, v + 3);
    p.vel.insert(p.vel.end(), v + 3, v + 6);
  }
  return p;
}

FrameWriter::FrameWriter(const std::string& path, const double area[3], int nbytes) : nbytes_(nbytes) {
  f_ = std::fopen(path.c_str(), "wb");                       // status = "replace" (:169)
  if (!f_) throw std::runtime_error("Could not open output file.");
  for (int d = 0; d < 3; ++d) {                              // records 1..3 (:54-56)
    if (nbytes_ == 4) { float v = (float)area[d]; std::fwrite(&v, 4, 1, f_); }
    else { double v = area[d]; std::fwrite(&v, 8, 1, f_); }
  }
  // record 4 (frames so far) is NOT written here: the reference writes it inside write_data only
  // (:336-337), so a run with MAX_TIME_STEP < WRITE_TIME_STEP leaves a file of three records
}

FrameWriter::~FrameWriter() { close(); }

void FrameWriter::close() {
  if (f_) { std::fclose(f_); f_ = nullptr; }
}

void FrameWriter::put_int(int32_t v) {
  unsigned char rec[8] = {0};
  std::memcpy(rec, &v, 4);                                   // low 4 bytes of the record
  std::fwrite(rec, (size_t)nbytes_, 1, f_);
}

void FrameWriter::write_frame(int64_t n, const int32_t* id, const void* pos, const void* vel, const uint8_t* active) {
  const size_t rb = (size_t)nbytes_;
  int64_t n_act = 0;
  for (int64_t i = 0; i < n; ++i) n_act += (!active || active[i]) ? 1 : 0;
  ++frames_;
  std::fseek(f_, (long)(3 * rb), SEEK_SET);                  // rec=4 rewritten every frame (:336-337)
  put_int(frames_);
  std::fseek(f_, 0, SEEK_END);
  buf_.assign((size_t)(7 * n_act + 1) * rb, 0);
  unsigned char* w = buf_.data();
  const int32_t count = (int32_t)(7 * n_act);                // :361
  std::memcpy(w, &count, 4);
  w += rb;
  const unsigned char* pp = (const unsigned char*)pos;
  const unsigned char* vv = (const unsigned char*)vel;
  for (int64_t i = 0; i < n; ++i) {
    if (active && !active[i]) continue;                      // :342
    std::memcpy(w, &id[i], 4); w += rb;
    std::memcpy(w, pp + 3 * rb * i, 3 * rb); w += 3 * rb;
    std::memcpy(w, vv + 3 * rb * i, 3 * rb); w += 3 * rb;
  }
  std::fwrite(buf_.data(), 1, buf_.size(), f_);
}

}  // namespace psim_host
