// ply.cpp -- PCL-free PLY reader/writer for the lidar cloud (host side of a1).
// Replaces pcl::io::loadPLYFile<LidarPoint> (src/lidar/ply.cc:14): vertex fields x y z and
// normal_x normal_y normal_z (also accepted: nx ny nz), ASCII or binary little-endian.
// Missing normal fields read as 0, like PCL leaves unmatched fields at their default.
#include <stdint.h>
#include <cmath>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <sstream>
#include <fstream>
#include <atomic>
#include <charconv>
#include <thread>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include "../../include/pcd_b200.h"

namespace pcd { void set_global_error(const std::string& s); }

namespace {

struct Prop { std::string name; int size; char kind; };  // kind: 'f' float, 'd' double, 'i' signed, 'u' unsigned

bool parse_type(const std::string& t, Prop& p) {
  if (t == "float" || t == "float32") { p.size = 4; p.kind = 'f'; return true; }
  if (t == "double" || t == "float64") { p.size = 8; p.kind = 'd'; return true; }
  if (t == "char" || t == "int8") { p.size = 1; p.kind = 'i'; return true; }
  if (t == "uchar" || t == "uint8") { p.size = 1; p.kind = 'u'; return true; }
  if (t == "short" || t == "int16") { p.size = 2; p.kind = 'i'; return true; }
  if (t == "ushort" || t == "uint16") { p.size = 2; p.kind = 'u'; return true; }
  if (t == "int" || t == "int32") { p.size = 4; p.kind = 'i'; return true; }
  if (t == "uint" || t == "uint32") { p.size = 4; p.kind = 'u'; return true; }
  return false;
}

float read_bin(const unsigned char* b, const Prop& p) {
  switch (p.kind) {
    case 'f': { float v; memcpy(&v, b, 4); return v; }
    case 'd': { double v; memcpy(&v, b, 8); return (float)v; }
    case 'i':
      if (p.size == 1) { int8_t v; memcpy(&v, b, 1); return (float)v; }
      if (p.size == 2) { int16_t v; memcpy(&v, b, 2); return (float)v; }
      { int32_t v; memcpy(&v, b, 4); return (float)v; }
    default:
      if (p.size == 1) { uint8_t v; memcpy(&v, b, 1); return (float)v; }
      if (p.size == 2) { uint16_t v; memcpy(&v, b, 2); return (float)v; }
      { uint32_t v; memcpy(&v, b, 4); return (float)v; }
  }
}

int slot_of(const std::string& n) {
  if (n == "x") return 0;
  if (n == "y") return 1;
  if (n == "z") return 2;
  if (n == "normal_x" || n == "nx") return 4;
  if (n == "normal_y" || n == "ny") return 5;
  if (n == "normal_z" || n == "nz") return 6;
  return -1;
}

int fail(int code, const std::string& msg) { pcd::set_global_error(msg); return code; }

// Splits [0, n) over the host threads (at most 16, at least `grain` items each) and runs fn(t, lo, hi).
template <class F>
void parallel_ranges(int64_t n, int64_t grain, F fn) {
  unsigned hw = std::thread::hardware_concurrency();
  int64_t T = hw ? (hw > 16 ? 16 : hw) : 1;
  if (T > (n + grain - 1) / grain) T = (n + grain - 1) / grain;
  if (T <= 1) { fn(0, (int64_t)0, n); return; }
  std::vector<std::thread> th;
  for (int64_t t = 0; t < T; t++) th.emplace_back(fn, (int)t, n * t / T, n * (t + 1) / T);
  for (auto& x : th) x.join();
}

// Read-only mapping of a whole file.
struct FileMap {
  const char* p = nullptr; size_t len = 0;
  bool open(const char* path) {
    int fd = ::open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0) { ::close(fd); return false; }
    len = (size_t)st.st_size;
    if (len > 0) {
      void* m = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
      if (m == MAP_FAILED) { ::close(fd); return false; }
      p = (const char*)m;
    }
    ::close(fd);
    return true;
  }
  ~FileMap() { if (p) munmap((void*)p, len); }
};

// One decimal token of an ASCII vertex line -> double, with strtod's value (the caller narrows to float).
// Returns the position after the token, or nullptr when there is none.
const char* parse_number(const char* s, const char* end, double& v) {
  while (s < end && (*s == ' ' || *s == '\t' || *s == '\r')) s++;
  if (s >= end) return nullptr;
  const char* t = (*s == '+') ? s + 1 : s;
  auto r = std::from_chars(t, end, v);
  if (r.ec == std::errc() && r.ptr > t) return r.ptr;
  // out-of-range values, hexadecimal floats ...: let strtod decide on a bounded copy of the token
  char tok[128]; size_t k = 0;
  while (s + k < end && k + 1 < sizeof tok && s[k] != ' ' && s[k] != '\t' && s[k] != '\r') { tok[k] = s[k]; k++; }
  tok[k] = 0;
  char* e = nullptr;
  v = strtod(tok, &e);
  return e == tok ? nullptr : s + (e - tok);
}

}  // namespace

extern "C" int pcd_ply_read(const char* path, float** aos8, int64_t* n_out) {
  if (!path || !aos8 || !n_out) return fail(PCD_E_INVALID, "pcd_ply_read: NULL argument");
  *aos8 = nullptr; *n_out = 0;
  std::ifstream f(path, std::ios::binary);
  if (!f) return fail(PCD_E_IO, std::string("Couldn't read file ") + path);
  std::string line;
  if (!std::getline(f, line) || line.compare(0, 3, "ply") != 0) return fail(PCD_E_IO, std::string(path) + ": not a PLY file");
  bool ascii = false, le = false, in_vertex = false, vertex_first = true, seen_vertex = false, header_done = false;
  int64_t n = 0;
  std::vector<Prop> props;
  int n_elements_before = 0;
  while (std::getline(f, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    std::istringstream ss(line);
    std::string tok; ss >> tok;
    if (tok == "format") {
      std::string fmt; ss >> fmt;
      if (fmt == "ascii") ascii = true;
      else if (fmt == "binary_little_endian") le = true;
      else return fail(PCD_E_IO, std::string(path) + ": unsupported PLY format " + fmt);
    } else if (tok == "element") {
      std::string name; int64_t cnt = 0; ss >> name >> cnt;
      in_vertex = (name == "vertex");
      if (in_vertex) { n = cnt; seen_vertex = true; vertex_first = (n_elements_before == 0); }
      n_elements_before++;
    } else if (tok == "property") {
      if (!in_vertex) continue;
      std::string type; ss >> type;
      if (type == "list") return fail(PCD_E_IO, std::string(path) + ": list property in vertex element");
      Prop p; if (!parse_type(type, p)) return fail(PCD_E_IO, std::string(path) + ": unknown property type " + type);
      ss >> p.name; props.push_back(p);
    } else if (tok == "end_header") { header_done = true; break; }
  }
  if (!header_done || !seen_vertex || (!ascii && !le)) return fail(PCD_E_IO, std::string(path) + ": malformed PLY header");
  if (!vertex_first) return fail(PCD_E_IO, std::string(path) + ": vertex must be the first element");
  if (n < 0) return fail(PCD_E_IO, std::string(path) + ": negative vertex count");
  std::vector<int> slot(props.size());
  size_t stride = 0;
  bool have_xyz[3] = {false, false, false};
  for (size_t i = 0; i < props.size(); i++) {
    slot[i] = slot_of(props[i].name);
    if (slot[i] >= 0 && slot[i] < 3) have_xyz[slot[i]] = true;
    stride += props[i].size;
  }
  if (!(have_xyz[0] && have_xyz[1] && have_xyz[2])) return fail(PCD_E_IO, std::string(path) + ": vertex has no x/y/z");
  const std::streamoff body_off = (std::streamoff)f.tellg();
  f.close();
  FileMap fm;
  if (!fm.open(path) || body_off < 0 || (size_t)body_off > fm.len) return fail(PCD_E_IO, std::string("Couldn't read file ") + path);
  const char* body = fm.p + body_off;
  const size_t body_len = fm.len - (size_t)body_off;
  if (!ascii && (stride == 0 || (size_t)n > body_len / stride)) return fail(PCD_E_IO, std::string(path) + ": truncated vertex data");
  float* out = (float*)calloc((size_t)(n > 0 ? n : 1) * 8, sizeof(float));
  if (!out) return fail(PCD_E_NOMEM, "pcd_ply_read: out of memory");
  const size_t np = props.size();
  if (ascii) {
    // one vertex per line.  The body is cut into byte ranges at line ends; a first pass counts the lines of every
    // range, so that the second pass knows which vertex each range starts with.
    unsigned hw = std::thread::hardware_concurrency();
    int T = hw ? (hw > 16 ? 16 : (int)hw) : 1;
    if ((size_t)T > body_len / (1 << 20) + 1) T = (int)(body_len / (1 << 20) + 1);
    std::vector<size_t> cut(T + 1);
    cut[0] = 0; cut[T] = body_len;
    for (int t = 1; t < T; t++) {
      size_t c = body_len / T * t;
      if (c < cut[t - 1]) c = cut[t - 1];
      const void* nl = c < body_len ? memchr(body + c, '\n', body_len - c) : nullptr;
      cut[t] = nl ? (size_t)((const char*)nl - body) + 1 : body_len;
    }
    std::vector<int64_t> first(T + 1, 0);
    {
      std::vector<int64_t> cnt(T, 0);
      std::vector<std::thread> th;
      for (int t = 0; t < T; t++)
        th.emplace_back([&, t] {
          int64_t c = 0;
          const char* q = body + cut[t]; const char* e = body + cut[t + 1];
          while (q < e) { const void* nl = memchr(q, '\n', e - q); if (!nl) { c++; break; } c++; q = (const char*)nl + 1; }
          cnt[t] = c;
        });
      for (auto& x : th) x.join();
      for (int t = 0; t < T; t++) first[t + 1] = first[t] + cnt[t];
    }
    if (first[T] < n) { free(out); return fail(PCD_E_IO, std::string(path) + ": truncated vertex data"); }
    std::atomic<int> bad(0);
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++)
      th.emplace_back([&, t] {
        const char* q = body + cut[t]; const char* e = body + cut[t + 1];
        for (int64_t i = first[t]; i < first[t + 1] && i < n && q < e; i++) {
          const void* nl = memchr(q, '\n', e - q);
          const char* le_ = nl ? (const char*)nl : e;
          const char* sp = q;
          for (size_t k = 0; k < np; k++) {
            double v;
            sp = parse_number(sp, le_, v);
            if (!sp) { bad.store(1); return; }
            if (slot[k] >= 0) out[8 * i + slot[k]] = (float)v;
          }
          q = le_ + 1;
        }
      });
    for (auto& x : th) x.join();
    if (bad.load()) { free(out); return fail(PCD_E_IO, std::string(path) + ": bad vertex line"); }
  } else {
    bool all_f32 = true;
    std::vector<size_t> offs(np);
    { size_t off = 0; for (size_t k = 0; k < np; k++) { offs[k] = off; off += props[k].size; all_f32 = all_f32 && props[k].kind == 'f'; } }
    const unsigned char* ub = (const unsigned char*)body;
    parallel_ranges(n, 1 << 18, [&](int, int64_t lo, int64_t hi) {
      for (int64_t i = lo; i < hi; i++) {
        const unsigned char* b = ub + stride * (size_t)i;
        float* o = out + 8 * i;
        if (all_f32) {
          for (size_t k = 0; k < np; k++) if (slot[k] >= 0) memcpy(o + slot[k], b + offs[k], 4);
        } else {
          for (size_t k = 0; k < np; k++) if (slot[k] >= 0) o[slot[k]] = read_bin(b + offs[k], props[k]);
        }
      }
    });
  }
  *aos8 = out; *n_out = n;
  return PCD_OK;
}

extern "C" int pcd_ply_write(const char* path, const float* aos8, int64_t n, int binary) {
  if (!path || (!aos8 && n > 0)) return fail(PCD_E_INVALID, "pcd_ply_write: NULL argument");
  FILE* f = fopen(path, "wb");
  if (!f) return fail(PCD_E_IO, std::string("cannot open ") + path);
  fprintf(f, "ply\nformat %s 1.0\nelement vertex %lld\n", binary ? "binary_little_endian" : "ascii", (long long)n);
  fprintf(f, "property float x\nproperty float y\nproperty float z\n");
  fprintf(f, "property float normal_x\nproperty float normal_y\nproperty float normal_z\nend_header\n");
  bool ok = true;
  if (binary) {
    const int64_t chunk = 1 << 16;
    std::vector<float> rows(6 * (size_t)chunk);
    for (int64_t base = 0; base < n && ok; base += chunk) {
      const int64_t m = n - base < chunk ? n - base : chunk;
      for (int64_t i = 0; i < m; i++) {
        const float* p = aos8 + 8 * (base + i);
        float* r = rows.data() + 6 * i;
        r[0] = p[0]; r[1] = p[1]; r[2] = p[2]; r[3] = p[4]; r[4] = p[5]; r[5] = p[6];
      }
      ok = fwrite(rows.data(), 6 * sizeof(float), (size_t)m, f) == (size_t)m;
    }
  } else {
    for (int64_t i = 0; i < n && ok; i++) {
      const float* p = aos8 + 8 * i;
      ok = fprintf(f, "%.9g %.9g %.9g %.9g %.9g %.9g\n", p[0], p[1], p[2], p[4], p[5], p[6]) > 0;
    }
  }
  ok = (fclose(f) == 0) && ok;
  if (!ok) return fail(PCD_E_IO, std::string("write error on ") + path);
  return PCD_OK;
}

// ------------------------------------------------------------------------------------------------------------
// N4  pose prior file "pose.ply" (src/controllers/incremental_mapper.cc:922-997, IncrementalMapperController::LoadPose):
// ASCII, one line per image after `end_header`: x y z roll pitch yaw in the LIDAR frame (radians).  image_id counts
// lines from 1; a line containing the token "nan" is skipped but still consumes its id.  Conversion to the camera
// frame: t_wc = (-y, -z, x); R_wc = AngleAxis(-yaw, Y) * AngleAxis(-pitch, X) * AngleAxis(roll, Z);
// R_cw = R_wc^T, t_cw = -R_cw t_wc, q_cw = Quaterniond(R_cw).  Output row: image_id, then tx ty tz qw qx qy qz (the
// order of the reference's trans_pose vector).  Host-side format code: a few hundred rows per reconstruction.
// ------------------------------------------------------------------------------------------------------------
static void rot_axis(int axis, double a, double R[9]) {           // Eigen::AngleAxisd(a, UnitX/Y/Z).toRotationMatrix()
  const double c = cos(a), s = sin(a);
  const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  memcpy(R, I, sizeof I);
  const int i = (axis + 1) % 3, j = (axis + 2) % 3;
  R[3 * i + i] = c; R[3 * i + j] = -s; R[3 * j + i] = s; R[3 * j + j] = c;
}
static void mat3_mul(const double A[9], const double B[9], double C[9]) {
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) C[3 * r + c] = A[3 * r] * B[c] + A[3 * r + 1] * B[3 + c] + A[3 * r + 2] * B[6 + c];
}
static void mat3_to_quat(const double m[9], double q[4]) {        // Eigen::Quaternion(Matrix3) (Shoemake), w x y z
  double t = m[0] + m[4] + m[8];
  if (t > 0) {
    t = sqrt(t + 1.0); q[0] = 0.5 * t; t = 0.5 / t;
    q[1] = (m[7] - m[5]) * t; q[2] = (m[2] - m[6]) * t; q[3] = (m[3] - m[1]) * t;
  } else {
    int i = 0;
    if (m[4] > m[0]) i = 1;
    if (m[8] > m[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrt(m[4 * i] - m[4 * j] - m[4 * k] + 1.0);
    q[1 + i] = 0.5 * t; t = 0.5 / t;
    q[0] = (m[3 * k + j] - m[3 * j + k]) * t;
    q[1 + j] = (m[3 * j + i] + m[3 * i + j]) * t;
    q[1 + k] = (m[3 * k + i] + m[3 * i + k]) * t;
  }
}

extern "C" int pcd_pose_ply_read(const char* path, int64_t* image_ids, double* tq7, int64_t capacity, int64_t* n_out) {
  if (!path || !n_out || (capacity > 0 && (!image_ids || !tq7))) return fail(PCD_E_INVALID, "pcd_pose_ply_read: NULL argument");
  *n_out = 0;
  std::ifstream f(path);
  if (!f) return fail(PCD_E_IO, std::string("failed to open the pose file ") + path);
  std::string line;
  bool header_done = false;
  int64_t image_id = 0, n = 0;
  while (std::getline(f, line)) {
    if (!line.empty() && line.back() == '\r') line.pop_back();
    if (!header_done) { if (line == "end_header") header_done = true; continue; }
    image_id++;
    std::istringstream tok(line);
    std::string s; bool has_nan = false;
    while (tok >> s) if (s == "nan") { has_nan = true; break; }
    if (has_nan) continue;
    std::istringstream ss(line);
    double v[6]; int k = 0; double d;
    while (k < 6 && (ss >> d)) v[k++] = d;
    if (k < 6) return fail(PCD_E_IO, std::string(path) + ": pose line with fewer than 6 numbers");
    const double twc[3] = {-v[1], -v[2], v[0]};
    double Rr[9], Rp[9], Ry[9], T[9], Rwc[9];
    rot_axis(2, v[3], Rr); rot_axis(0, -v[4], Rp); rot_axis(1, -v[5], Ry);
    mat3_mul(Ry, Rp, T); mat3_mul(T, Rr, Rwc);
    double Rcw[9];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Rcw[3 * r + c] = Rwc[3 * c + r];
    double q[4]; mat3_to_quat(Rcw, q);
    if (n < capacity) {
      image_ids[n] = image_id;
      double* o = tq7 + 7 * n;
      for (int r = 0; r < 3; r++) o[r] = -(Rcw[3 * r] * twc[0] + Rcw[3 * r + 1] * twc[1] + Rcw[3 * r + 2] * twc[2]);
      o[3] = q[0]; o[4] = q[1]; o[5] = q[2]; o[6] = q[3];
    }
    n++;
  }
  *n_out = n;            // rows available; call again with a larger buffer if n > capacity
  return PCD_OK;
}

// Writer of the same file (the reference writes it from GUI code, src/ui/main_window.cc:1096-1160): x y z roll pitch yaw.
extern "C" int pcd_pose_ply_write(const char* path, const double* xyzrpy, int64_t n) {
  if (!path || (!xyzrpy && n > 0)) return fail(PCD_E_INVALID, "pcd_pose_ply_write: NULL argument");
  FILE* f = fopen(path, "w");
  if (!f) return fail(PCD_E_IO, std::string("cannot open ") + path);
  fprintf(f, "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n"
             "property float roll\nproperty float pitch\nproperty float yaw\nend_header\n", (long long)n);
  for (int64_t i = 0; i < n; i++) {
    const double* p = xyzrpy + 6 * i;
    bool nan = false;
    for (int k = 0; k < 6; k++) nan |= std::isnan(p[k]);
    if (nan) fprintf(f, "nan nan nan nan nan nan\n");
    else fprintf(f, "%.17g %.17g %.17g %.17g %.17g %.17g\n", p[0], p[1], p[2], p[3], p[4], p[5]);
  }
  fclose(f);
  return PCD_OK;
}
