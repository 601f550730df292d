// ply.cpp -- PCL-free PLY reader/writer for the lidar cloud (host side of a1).
// Replaces pcl::io::loadPLYFile<LidarPoint> (src/lidar/ply.cc:14): vertex fields x y z and
// normal_x normal_y normal_z (also accepted: nx ny nz), ASCII or binary little-endian.
// Missing normal fields read as 0, like PCL leaves unmatched fields at their default.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <sstream>
#include <fstream>
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
  float* out = (float*)calloc((size_t)(n > 0 ? n : 1) * 8, sizeof(float));
  if (!out) return fail(PCD_E_NOMEM, "pcd_ply_read: out of memory");
  if (ascii) {
    for (int64_t i = 0; i < n; i++) {
      if (!std::getline(f, line)) { free(out); return fail(PCD_E_IO, std::string(path) + ": truncated vertex data"); }
      const char* s = line.c_str(); char* e = nullptr;
      for (size_t k = 0; k < props.size(); k++) {
        double v = strtod(s, &e);
        if (e == s) { free(out); return fail(PCD_E_IO, std::string(path) + ": bad vertex line"); }
        s = e;
        if (slot[k] >= 0) out[8 * i + slot[k]] = (float)v;
      }
    }
  } else {
    const int64_t chunk = 1 << 16;
    std::vector<unsigned char> buf(stride * (size_t)chunk);
    for (int64_t base = 0; base < n; base += chunk) {
      int64_t m = (n - base < chunk) ? (n - base) : chunk;
      f.read((char*)buf.data(), (std::streamsize)(stride * (size_t)m));
      if ((size_t)f.gcount() != stride * (size_t)m) { free(out); return fail(PCD_E_IO, std::string(path) + ": truncated vertex data"); }
      for (int64_t i = 0; i < m; i++) {
        const unsigned char* b = buf.data() + stride * (size_t)i;
        size_t off = 0;
        for (size_t k = 0; k < props.size(); k++) {
          if (slot[k] >= 0) out[8 * (base + i) + slot[k]] = read_bin(b + off, props[k]);
          off += props[k].size;
        }
      }
    }
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
  for (int64_t i = 0; i < n; i++) {
    const float* p = aos8 + 8 * i;
    float row[6] = {p[0], p[1], p[2], p[4], p[5], p[6]};
    if (binary) fwrite(row, sizeof(float), 6, f);
    else fprintf(f, "%.9g %.9g %.9g %.9g %.9g %.9g\n", row[0], row[1], row[2], row[3], row[4], row[5]);
  }
  fclose(f);
  return PCD_OK;
}
