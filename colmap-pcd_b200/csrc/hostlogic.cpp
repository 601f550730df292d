// hostlogic.cpp -- the two host-side pieces either side of the GPU path that are plain control / file logic:
//   a16  BundleAdjustmentConfig::Project2Image's decision which images to project   src/optim/bundle_adjustment.cc:241-281
//   N3   one-time load: a binary sidecar of the parsed, axis-swapped cloud           src/controllers/bundle_adjustment.cc:123-124,
//        so that re-runs skip the PLY parse and PointCloudDirectionTrans             src/lidar/ply.cc:9-57
// Integer and byte work only; no arithmetic of the hot path lives here.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <sys/stat.h>
#include "../../include/pcd_b200.h"

namespace pcd { void set_global_error(const std::string& s); }

static int fail(int code, const std::string& msg) { pcd::set_global_error(msg); return code; }

// One point's track against the reference's rule.  track_image[k]: image of the k-th track element; pair_corrs[k]:
// num_total_corrs of the pair (image_id, track_image[k]) or -1 when reconstruction->ExistsImagePair is false;
// cached[k] != 0: the image is already in lidar_searched_image_ids_.  project[k] = 1 for the elements whose image the
// reference would now hand to SetNewImage -- an image reached twice in one track is projected by its first element only,
// because that call caches it (:263-271).  Returns the number of images selected.
extern "C" int64_t pcd_project2image_select(const uint32_t* track_image, int64_t n, uint32_t image_id, const int64_t* pair_corrs,
                                            int match_features_threshold, const uint8_t* cached, uint8_t* project) {
  if (n < 0 || (n > 0 && (!track_image || !pair_corrs || !project))) return -(int64_t)fail(PCD_E_INVALID, "pcd_project2image_select: bad argument");
  int64_t count = 0;
  for (int64_t k = 0; k < n; k++) {
    project[k] = 0;
    const uint32_t image_id2 = track_image[k];
    if (image_id2 != image_id) {
      if (pair_corrs[k] >= 0) {                                                    // ExistsImagePair (:248)
        const size_t corrs = (size_t)pair_corrs[k];
        if (corrs <= (size_t)match_features_threshold) continue;                    // size_t <= int: the int converts (:250)
      }
    }
    bool have = cached && cached[k];
    for (int64_t j = 0; j < k && !have; j++) have = project[j] && track_image[j] == image_id2;
    if (!have) { project[k] = 1; count++; }
  }
  return count;
}

// ---- N3: sidecar ------------------------------------------------------------------------------------------------
// Layout (little-endian): "PCDB200C" | u32 version | u32 direction_trans | u64 key | i64 n | n x 8 float32.
// key identifies the source PLY without reading all of it: FNV-1a over its size, its modification time and its first
// and last 64 KiB.
static const char MAGIC[8] = {'P', 'C', 'D', 'B', '2', '0', '0', 'C'};

static uint64_t fnv(uint64_t h, const void* p, size_t n) {
  const unsigned char* b = (const unsigned char*)p;
  for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
  return h;
}

extern "C" int pcd_file_key(const char* path, uint64_t* key) {
  if (!path || !key) return fail(PCD_E_INVALID, "pcd_file_key: bad argument");
  struct stat st;
  if (stat(path, &st) != 0) return fail(PCD_E_IO, std::string("cannot stat ") + path);
  FILE* f = fopen(path, "rb");
  if (!f) return fail(PCD_E_IO, std::string("cannot open ") + path);
  uint64_t h = 1469598103934665603ull;
  const int64_t size = (int64_t)st.st_size, mt = (int64_t)st.st_mtim.tv_sec * 1000000000ll + st.st_mtim.tv_nsec;
  h = fnv(h, &size, 8); h = fnv(h, &mt, 8);
  static const size_t W = 65536;
  char* buf = (char*)malloc(W);
  size_t got = fread(buf, 1, W, f);
  h = fnv(h, buf, got);
  if (size > (int64_t)W) {
    fseek(f, (long)(size - (int64_t)W), SEEK_SET);
    got = fread(buf, 1, W, f);
    h = fnv(h, buf, got);
  }
  free(buf);
  fclose(f);
  *key = h;
  return PCD_OK;
}

extern "C" int pcd_sidecar_write(const char* path, uint64_t key, int direction_trans, const float* aos8, int64_t n) {
  if (!path || n < 0 || (n > 0 && !aos8)) return fail(PCD_E_INVALID, "pcd_sidecar_write: bad argument");
  const std::string tmp = std::string(path) + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return fail(PCD_E_IO, "cannot write " + tmp);
  const uint32_t version = 1, dt = (uint32_t)direction_trans;
  bool ok = fwrite(MAGIC, 1, 8, f) == 8 && fwrite(&version, 4, 1, f) == 1 && fwrite(&dt, 4, 1, f) == 1 && fwrite(&key, 8, 1, f) == 1 &&
            fwrite(&n, 8, 1, f) == 1 && (n == 0 || fwrite(aos8, 32, (size_t)n, f) == (size_t)n);
  ok = (fclose(f) == 0) && ok;
  if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return fail(PCD_E_IO, std::string("writing ") + path + " failed"); }
  return PCD_OK;
}

// PCD_E_IO when the sidecar is missing, damaged, of another version or belongs to another source file (key mismatch).
extern "C" int pcd_sidecar_read(const char* path, uint64_t key, int direction_trans, float** aos8, int64_t* n_out) {
  if (!path || !aos8 || !n_out) return fail(PCD_E_INVALID, "pcd_sidecar_read: bad argument");
  *aos8 = nullptr; *n_out = 0;
  FILE* f = fopen(path, "rb");
  if (!f) return fail(PCD_E_IO, std::string("no sidecar ") + path);
  char magic[8]; uint32_t version = 0, dt = 0; uint64_t k = 0; int64_t n = -1;
  bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, MAGIC, 8) == 0 && fread(&version, 4, 1, f) == 1 && version == 1 &&
            fread(&dt, 4, 1, f) == 1 && fread(&k, 8, 1, f) == 1 && fread(&n, 8, 1, f) == 1;
  if (ok && (k != key || dt != (uint32_t)direction_trans || n < 0 || n >= ((int64_t)1 << 31))) ok = false;
  float* buf = nullptr;
  if (ok) {
    struct stat st;
    ok = fstat(fileno(f), &st) == 0 && (int64_t)st.st_size == 32 + 32 * n;          // truncated or padded files are rejected
  }
  if (ok) {
    buf = (float*)malloc(n > 0 ? 32 * (size_t)n : 32);
    ok = buf && (n == 0 || fread(buf, 32, (size_t)n, f) == (size_t)n);
  }
  fclose(f);
  if (!ok) { free(buf); return fail(PCD_E_IO, std::string("sidecar ") + path + " is not usable for this source file"); }
  *aos8 = buf; *n_out = n;
  return PCD_OK;
}
