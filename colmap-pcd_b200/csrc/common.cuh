// common.cuh -- shared host-side state of libpcd_b200 (context, cloud, scratch, error plumbing).
// Product code: never includes or links anything under oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include <mutex>
#include "../../include/pcd_b200.h"

namespace pcd {

void set_global_error(const std::string& s);

// Grow-only device buffer.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
    cudaError_t e = cudaMallocHost(&p, bytes);
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Uniform-grid NN index view passed by value to kernels.
struct GridView {
  const float4* pts;            // cell-sorted: x, y, z, __int_as_float(original index)
  const uint32_t* cell_start;   // n_cells + 1 offsets into pts
  float ox, oy, oz;             // origin = bbox min
  float c, inv_c;               // cell edge along y and z (power of two) and its reciprocal
  float cx, inv_cx;             // cell edge along x = c / xs: rows are contiguous along x, so a finer
  int xs;                       //   x resolution narrows the candidate run at no extra lookup cost
  int dx, dy, dz;               // grid dimensions (x fastest; dx counts x cells of edge cx)
  float margin;                 // conservative slack (metres) for the float cell assignment
  // occupancy pyramid for far queries: one 64-bit word per block of 4 x 4 x 4 cubic cells of edge c
  // (x cell = ix / xs); bit (lz * 4 + ly) * 4 + lx is set iff the cell holds a point
  const unsigned long long* occ;
  int nbx, nby, nbz;
  uint32_t n;
};

// 1 m sub-map of PcdProj::BuildSubMap, cell-sorted.
struct SubmapView {
  const int32_t* keys;          // 3 per cell, lexicographic order
  const uint32_t* start;        // n_cells + 1
  const float4* pts;            // x, y, z, original index; insertion order inside a cell
  int64_t n_cells;
  const uint2* seg;             // work list of the projection: {cell, first point}, at most seg_len points each
  int64_t n_seg;
  uint32_t seg_len;
  float len, height, width;
};
constexpr uint32_t SUBMAP_SEG = 64;    // default points per work item: bounds the serial work of one warp in k_project

}  // namespace pcd

struct pcd_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;       // used by the host-pointer entry points
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  cudaEvent_t ev[8] = {};
  std::string err;
  int64_t launches = 0;
  int64_t counters[4] = {0, 0, 0, 0};
  int sm_count = 148;
  // scratch for the batched search (one batch in flight per context)
  pcd::DevBuf keys, keys_alt, vals, vals_alt, cub_tmp, misc, partial;
  pcd::DevBuf stage_in[2], stage_out[2], stage_aux[2];
  pcd::PinBuf pin_in[2], pin_out[2];
  pcd::DevBuf img_buf, feat_buf, zbuf, sel_buf, far_q, qrec, eval_partial;
  // optional per-kernel CUDA-event timing (pcd_ctx_profile); spans are resolved lazily
  bool prof = false;
  struct Span { cudaEvent_t a, b; int which; };
  std::vector<Span> prof_pending;
  std::vector<cudaEvent_t> prof_free;
  double prof_ms[PCD_K_COUNT] = {};
  int64_t prof_n[PCD_K_COUNT] = {};
};

namespace pcd {
// Brackets the launches of one kernel class with CUDA events on the launching stream.
struct ProfScope {
  pcd_ctx* c; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr; int which;
  static cudaEvent_t get(pcd_ctx* c) {
    if (!c->prof_free.empty()) { cudaEvent_t e = c->prof_free.back(); c->prof_free.pop_back(); return e; }
    cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
  }
  ProfScope(pcd_ctx* c_, int which_, cudaStream_t st_) : c(c_), st(st_), which(which_) {
    if (!c->prof) return;
    a = get(c); b = get(c);
    cudaEventRecord(a, st);
  }
  ~ProfScope() {
    if (!a) return;
    cudaEventRecord(b, st);
    c->prof_pending.push_back({a, b, which});
  }
};
}  // namespace pcd

struct pcd_cloud {
  pcd_ctx* ctx = nullptr;
  int64_t n = 0;
  float* aos = nullptr;                // AoS8, post-swap, index space
  // NN grid
  bool has_index = false;
  float4* grid_pts = nullptr;
  uint32_t* cell_start = nullptr;
  unsigned long long* occ = nullptr;
  pcd::GridView grid = {};
  int64_t n_cells = 0;
  float bb_lo[3] = {0, 0, 0}, bb_hi[3] = {0, 0, 0};
  // sub-map
  bool has_submap = false;
  pcd_proj_options opt = {};
  int32_t* sub_keys = nullptr;
  uint32_t* sub_start = nullptr;
  float4* sub_pts = nullptr;
  uint2* sub_seg = nullptr;
  int64_t sub_cells = 0, sub_nseg = 0;
  uint32_t sub_seglen = pcd::SUBMAP_SEG;
};

// Debug build (make dbg -> libpcd_b200_dbg.so, -DPCD_DEBUG_CHECKS): device-side bounds assertions at every indexed
// access of the search and projection kernels.  compute-sanitizer is closed on this pool, so this build plus
// tools/sanitize_run.py is the memory check of the path (profiles/r02_debug_checks.log).  Compiled out otherwise.
#ifdef PCD_DEBUG_CHECKS
#include <assert.h>
#define PCD_DASSERT(cond_) assert(cond_)
#else
#define PCD_DASSERT(cond_) ((void)0)
#endif

#define PCD_FAIL(ctx_, code_, ...)                                   \
  do {                                                               \
    char _b[512];                                                    \
    snprintf(_b, sizeof _b, __VA_ARGS__);                            \
    if (ctx_) (ctx_)->err = _b;                                      \
    pcd::set_global_error(_b);                                       \
    return (code_);                                                  \
  } while (0)

#define PCD_CUDA(ctx_, expr_)                                                            \
  do {                                                                                   \
    cudaError_t _e = (expr_);                                                            \
    if (_e != cudaSuccess)                                                               \
      PCD_FAIL(ctx_, PCD_E_CUDA, "CUDA error %s at %s:%d (%s)", cudaGetErrorString(_e),  \
               __FILE__, __LINE__, #expr_);                                              \
  } while (0)

#define PCD_LAUNCH_CHECK(ctx_)                                                           \
  do {                                                                                   \
    (ctx_)->launches++;                                                                  \
    PCD_CUDA(ctx_, cudaGetLastError());                                                  \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

static inline int pcd_blocks(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }
