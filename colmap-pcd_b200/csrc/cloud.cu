// cloud.cu -- cloud upload (a1/a2), NN grid index build (a3) and 1 m sub-map build (a9).
#include "common.cuh"
#include <string>
#include <cub/cub.cuh>
#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace {

// ---------------------------------------------------------------------------------------
// a2  PointCloudProcess::PointCloudDirectionTrans (src/lidar/ply.cc:33-57) on the device:
// flag rows without NaN after the swap, exclusive-scan the flags, scatter in order.
// ---------------------------------------------------------------------------------------
__global__ void k_trans_flag(const float4* __restrict__ in, int64_t n, int do_swap, uint32_t* __restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = in[2 * i], b = in[2 * i + 1];
  bool bad = do_swap && (isnan(a.x) || isnan(a.y) || isnan(a.z) || isnan(b.x) || isnan(b.y) || isnan(b.z));
  flag[i] = bad ? 0u : 1u;
}

__global__ void k_trans_scatter(const float4* __restrict__ in, int64_t n, int do_swap, const uint32_t* __restrict__ flag,
                                const uint32_t* __restrict__ pos, float4* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n || !flag[i]) return;
  float4 a = in[2 * i], b = in[2 * i + 1];
  float4 p, q;
  if (do_swap) {
    p = make_float4(-a.y, -a.z, a.x, 0.0f);
    q = make_float4(-b.y, -b.z, b.x, 0.0f);
  } else {
    p = make_float4(a.x, a.y, a.z, 0.0f);
    q = make_float4(b.x, b.y, b.z, 0.0f);
  }
  uint32_t o = pos[i];
  out[2 * (int64_t)o] = p;
  out[2 * (int64_t)o + 1] = q;
}

// ---- bounding box: block reduce then ordered-int atomics --------------------------------
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__host__ __device__ __forceinline__ float ord2f(int i) {
  int j = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  float f; memcpy(&f, &j, 4); return f;
#endif
}

__global__ void k_bbox(const float4* __restrict__ aos, int64_t n, int* __restrict__ bb /* lo[3], hi[3] ordered ints */) {
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float4 a = aos[2 * i];
    lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
    hi[0] = fmaxf(hi[0], a.x); hi[1] = fmaxf(hi[1], a.y); hi[2] = fmaxf(hi[2], a.z);
  }
  for (int d = 0; d < 3; d++) {
    for (int o = 16; o > 0; o >>= 1) {
      lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
      hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
    for (int d = 0; d < 3; d++) { atomicMin(&bb[d], f2ord(lo[d])); atomicMax(&bb[3 + d], f2ord(hi[d])); }
  }
}

// ---- NN grid ------------------------------------------------------------------------------
__device__ __forceinline__ int cell_of(float v, float o, float inv_c, int dim) {
  float t = (v - o) * inv_c;          // inv_c is a power of two: the product is exact
  int c = (int)floorf(t);
  return min(max(c, 0), dim - 1);
}

__global__ void k_grid_keys(const float4* __restrict__ aos, int64_t n, pcd::GridView g, uint32_t* __restrict__ keys,
                            uint32_t* __restrict__ vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = aos[2 * i];
  int cx = cell_of(a.x, g.ox, g.inv_cx, g.dx), cy = cell_of(a.y, g.oy, g.inv_c, g.dy), cz = cell_of(a.z, g.oz, g.inv_c, g.dz);
  keys[i] = (uint32_t)(((int64_t)cz * g.dy + cy) * g.dx + cx);
  vals[i] = (uint32_t)i;
}

__global__ void k_count_unique(const uint32_t* __restrict__ sorted, int64_t n, unsigned long long* __restrict__ cnt) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool head = i < n && (i == 0 || sorted[i] != sorted[i - 1]);
  unsigned m = __ballot_sync(0xffffffffu, head);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(cnt, (unsigned long long)__popc(m));
}

__global__ void k_grid_gather(const float4* __restrict__ aos, const uint32_t* __restrict__ skeys, const uint32_t* __restrict__ svals,
                              int64_t n, float4* __restrict__ pts, uint32_t* __restrict__ counts) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t o = svals[i];
  float4 a = aos[2 * (int64_t)o];
  pts[i] = make_float4(a.x, a.y, a.z, __uint_as_float(o));
  atomicAdd(&counts[skeys[i]], 1u);
}

// occupancy words of the 4x4x4-cell blocks (far-query pruning in nn.cu)
__global__ void k_grid_occupancy(const float4* __restrict__ pts, int64_t n, pcd::GridView g, unsigned long long* __restrict__ occ) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = pts[i];
  int cx = cell_of(a.x, g.ox, g.inv_cx, g.dx) / g.xs, cy = cell_of(a.y, g.oy, g.inv_c, g.dy), cz = cell_of(a.z, g.oz, g.inv_c, g.dz);
  int64_t w = ((int64_t)(cz >> 2) * g.nby + (cy >> 2)) * g.nbx + (cx >> 2);
  unsigned long long bit = 1ull << ((((cz & 3) << 2) | (cy & 3)) << 2 | (cx & 3));
  if (!(occ[w] & bit)) atomicOr(&occ[w], bit);
}

// ---- sub-map (a9): key = round(coord / len), y <-> height, z <-> width (pcd_projection.h:67-73)
__global__ void k_sub_keys(const float4* __restrict__ aos, int64_t n, float len, float height, float width, int kx0, int ky0,
                           int kz0, unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 a = aos[2 * i];
  long long kx = (long long)(int)roundf(__fdiv_rn(a.x, len)) - kx0;
  long long ky = (long long)(int)roundf(__fdiv_rn(a.y, height)) - ky0;
  long long kz = (long long)(int)roundf(__fdiv_rn(a.z, width)) - kz0;
  keys[i] = ((unsigned long long)kx << 42) | ((unsigned long long)ky << 21) | (unsigned long long)kz;
  vals[i] = (uint32_t)i;
}

__global__ void k_sub_gather(const float4* __restrict__ aos, const uint32_t* __restrict__ svals, int64_t n, float4* __restrict__ pts) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t o = svals[i];
  float4 a = aos[2 * (int64_t)o];
  pts[i] = make_float4(a.x, a.y, a.z, __uint_as_float(o));
}

__global__ void k_sub_decode(const unsigned long long* __restrict__ ukeys, int64_t nc, int kx0, int ky0, int kz0,
                             int32_t* __restrict__ keys3) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= nc) return;
  unsigned long long k = ukeys[i];
  keys3[3 * i] = (int)(k >> 42) + kx0;
  keys3[3 * i + 1] = (int)((k >> 21) & 0x1fffff) + ky0;
  keys3[3 * i + 2] = (int)(k & 0x1fffff) + kz0;
}

int upload_common(pcd_ctx* ctx, const float* d_in, int64_t n, int direction_trans, pcd_cloud** out) {
  // d_in: device AoS8 (raw).  Produces the compacted post-swap cloud.
  cudaStream_t st = ctx->stream;
  // released again on every early return below (PCD_CUDA returns on a CUDA error)
  struct Guard { pcd_cloud* c; ~Guard() { if (c) { cudaFree(c->aos); delete c; } } } guard{new pcd_cloud()};
  pcd_cloud* c = guard.c;
  c->ctx = ctx;
  int64_t m = n;
  if (n > 0) {
    PCD_CUDA(ctx, ctx->keys.ensure(sizeof(uint32_t) * (size_t)n));
    PCD_CUDA(ctx, ctx->vals.ensure(sizeof(uint32_t) * (size_t)n));
    uint32_t* flag = ctx->keys.as<uint32_t>();
    uint32_t* pos = ctx->vals.as<uint32_t>();
    k_trans_flag<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)d_in, n, direction_trans, flag);
    PCD_LAUNCH_CHECK(ctx);
    size_t tmp = 0;
    PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, flag, pos, (int)n, st));
    PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
    PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, flag, pos, (int)n, st));
    uint32_t last_pos = 0, last_flag = 0;
    PCD_CUDA(ctx, cudaMemcpyAsync(&last_pos, pos + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    PCD_CUDA(ctx, cudaMemcpyAsync(&last_flag, flag + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    PCD_CUDA(ctx, cudaStreamSynchronize(st));
    m = (int64_t)last_pos + last_flag;
    PCD_CUDA(ctx, cudaMalloc(&c->aos, sizeof(float) * 8 * (size_t)(m > 0 ? m : 1)));
    k_trans_scatter<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)d_in, n, direction_trans, flag, pos, (float4*)c->aos);
    PCD_LAUNCH_CHECK(ctx);
    PCD_CUDA(ctx, cudaStreamSynchronize(st));
  } else {
    PCD_CUDA(ctx, cudaMalloc(&c->aos, 32));
  }
  c->n = m;
  *out = c;
  guard.c = nullptr;
  return PCD_OK;
}

}  // namespace

extern "C" {

int pcd_cloud_from_device(pcd_ctx* ctx, const float* d_aos8, int64_t n, int direction_trans, pcd_cloud** out) {
  if (!ctx || !out || n < 0 || (n > 0 && !d_aos8)) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_cloud_from_device: bad argument");
  if (n >= (int64_t)1 << 31) PCD_FAIL(ctx, PCD_E_INVALID, "cloud too large (%lld points; limit 2^31-1)", (long long)n);
  DeviceGuard g(ctx->device);
  return upload_common(ctx, d_aos8, n, direction_trans, out);
}

int pcd_cloud_from_host(pcd_ctx* ctx, const float* aos8, int64_t n, int direction_trans, pcd_cloud** out) {
  if (!ctx || !out || n < 0 || (n > 0 && !aos8)) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_cloud_from_host: bad argument");
  if (n >= (int64_t)1 << 31) PCD_FAIL(ctx, PCD_E_INVALID, "cloud too large (%lld points; limit 2^31-1)", (long long)n);
  DeviceGuard g(ctx->device);
  float* d_in = nullptr;
  if (n > 0) {
    PCD_CUDA(ctx, cudaMalloc(&d_in, sizeof(float) * 8 * (size_t)n));
    cudaError_t e = cudaMemcpyAsync(d_in, aos8, sizeof(float) * 8 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) { cudaFree(d_in); PCD_CUDA(ctx, e); }
  }
  int rc = upload_common(ctx, d_in, n, direction_trans, out);
  if (d_in) cudaFree(d_in);
  return rc;
}

void pcd_cloud_destroy(pcd_cloud* c) {
  if (!c) return;
  DeviceGuard g(c->ctx->device);
  cudaDeviceSynchronize();
  cudaFree(c->aos); cudaFree(c->grid_pts); cudaFree(c->cell_start); cudaFree(c->occ);
  cudaFree(c->sub_keys); cudaFree(c->sub_start); cudaFree(c->sub_pts); cudaFree(c->sub_seg);
  delete c;
}

int64_t pcd_cloud_size(const pcd_cloud* c) { return c ? c->n : -1; }
const float* pcd_cloud_device_ptr(const pcd_cloud* c) { return c ? c->aos : nullptr; }

int pcd_cloud_download(const pcd_cloud* c, float* aos8) {
  if (!c || (!aos8 && c->n > 0)) return PCD_E_INVALID;
  pcd_ctx* ctx = c->ctx;
  DeviceGuard g(ctx->device);
  if (c->n > 0) {
    PCD_CUDA(ctx, cudaMemcpyAsync(aos8, c->aos, sizeof(float) * 8 * (size_t)c->n, cudaMemcpyDeviceToHost, ctx->stream));
    PCD_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return PCD_OK;
}

// ---------------------------------------------------------------------------------------
// a3  Kdtree::BuildMap (src/lidar/kdtree.cc:5-8) -> uniform grid.  Cell edge is a power of
// two so (p - origin) * inv_c is exact after the single rounding of the subtraction.
// ---------------------------------------------------------------------------------------
int pcd_index_build(pcd_cloud* c, float cell_size) {
  if (!c) return PCD_E_INVALID;
  pcd_ctx* ctx = c->ctx;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  const int64_t n = c->n;
  cudaFree(c->grid_pts); cudaFree(c->cell_start); cudaFree(c->occ);      // also what a failed build left behind
  c->grid_pts = nullptr; c->cell_start = nullptr; c->occ = nullptr; c->has_index = false;
  pcd::GridView& G = c->grid;
  memset(&G, 0, sizeof G);
  if (n == 0) {
    G.c = 1.0f; G.inv_c = 1.0f; G.cx = 1.0f; G.inv_cx = 1.0f; G.xs = 1; G.dx = G.dy = G.dz = 1; G.n = 0; c->n_cells = 1;
    PCD_CUDA(ctx, cudaMalloc(&c->grid_pts, 16));
    PCD_CUDA(ctx, cudaMalloc(&c->cell_start, 2 * sizeof(uint32_t)));
    PCD_CUDA(ctx, cudaMemsetAsync(c->cell_start, 0, 2 * sizeof(uint32_t), st));
    PCD_CUDA(ctx, cudaStreamSynchronize(st));
    G.nbx = G.nby = G.nbz = 1;
    PCD_CUDA(ctx, cudaMalloc(&c->occ, 8));
    PCD_CUDA(ctx, cudaMemsetAsync(c->occ, 0, 8, st));
    PCD_CUDA(ctx, cudaStreamSynchronize(st));
    G.pts = c->grid_pts; G.cell_start = c->cell_start; G.occ = c->occ; c->has_index = true;
    return PCD_OK;
  }
  // bbox
  PCD_CUDA(ctx, ctx->misc.ensure(256));
  int* d_bb = ctx->misc.as<int>();
  int h_bb[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
  PCD_CUDA(ctx, cudaMemcpyAsync(d_bb, h_bb, sizeof h_bb, cudaMemcpyHostToDevice, st));
  int nb = (int)((n + 1023) / 1024); if (nb > ctx->sm_count * 8) nb = ctx->sm_count * 8;
  k_bbox<<<nb, 256, 0, st>>>((const float4*)c->aos, n, d_bb);
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaMemcpyAsync(h_bb, d_bb, sizeof h_bb, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  for (int d = 0; d < 3; d++) { c->bb_lo[d] = ord2f(h_bb[d]); c->bb_hi[d] = ord2f(h_bb[3 + d]); }
  for (int d = 0; d < 3; d++)
    if (!isfinite(c->bb_lo[d]) || !isfinite(c->bb_hi[d]))
      PCD_FAIL(ctx, PCD_E_INVALID, "cloud has non-finite coordinates; cannot index");
  float ext = 0.0f, absmax = 0.0f;
  for (int d = 0; d < 3; d++) {
    ext = fmaxf(ext, c->bb_hi[d] - c->bb_lo[d]);
    absmax = fmaxf(absmax, fmaxf(fabsf(c->bb_lo[d]), fabsf(c->bb_hi[d])));
  }

  PCD_CUDA(ctx, ctx->keys.ensure(sizeof(uint32_t) * (size_t)n));
  PCD_CUDA(ctx, ctx->keys_alt.ensure(sizeof(uint32_t) * (size_t)n));
  PCD_CUDA(ctx, ctx->vals.ensure(sizeof(uint32_t) * (size_t)n));
  PCD_CUDA(ctx, ctx->vals_alt.ensure(sizeof(uint32_t) * (size_t)n));

  int xsub = 4;                                  // x cells per (y, z) cell edge; power of two
  if (const char* e = getenv("PCD_B200_XSUB")) { int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) xsub = v; }
  const bool auto_mode = !(cell_size > 0.0f);
  auto setup = [&](float cs) -> bool {   // fills G dims for cell edge cs; false if too many cells
    G.c = cs; G.inv_c = 1.0f / cs;
    G.xs = xsub; G.cx = cs / (float)xsub; G.inv_cx = G.inv_c * (float)xsub;
    G.ox = c->bb_lo[0]; G.oy = c->bb_lo[1]; G.oz = c->bb_lo[2];
    double dims[3];
    for (int d = 0; d < 3; d++) dims[d] = floor((double)(c->bb_hi[d] - c->bb_lo[d]) * (d == 0 ? G.inv_cx : G.inv_c)) + 1.0;
    double total = dims[0] * dims[1] * dims[2];
    if (auto_mode && total > fmax(6.4e7, 24.0 * (double)n)) return false;      // keep the cell table near the cloud's size
    if (total > 2.0e9 || dims[0] > 2.0e6 || dims[1] > 2.0e6 || dims[2] > 2.0e6) return false;
    G.dx = (int)dims[0]; G.dy = (int)dims[1]; G.dz = (int)dims[2];
    c->n_cells = (int64_t)G.dx * G.dy * G.dz;
    // float cell assignment error: one rounding of (p - o), magnitude <= 2^-24 * |p - o|.
    G.margin = 1e-3f * G.cx + 8.0f * (ext + absmax) * 1.2e-7f;
    G.n = (uint32_t)n;
    G.nbx = ((G.dx + xsub - 1) / xsub + 3) / 4; G.nby = (G.dy + 3) / 4; G.nbz = (G.dz + 3) / 4;
    return true;
  };

  cub::DoubleBuffer<uint32_t> kb(ctx->keys.as<uint32_t>(), ctx->keys_alt.as<uint32_t>());
  cub::DoubleBuffer<uint32_t> vb(ctx->vals.as<uint32_t>(), ctx->vals_alt.as<uint32_t>());
  auto sort_keys = [&]() -> int {
    kb = cub::DoubleBuffer<uint32_t>(ctx->keys.as<uint32_t>(), ctx->keys_alt.as<uint32_t>());
    vb = cub::DoubleBuffer<uint32_t>(ctx->vals.as<uint32_t>(), ctx->vals_alt.as<uint32_t>());
    k_grid_keys<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)c->aos, n, G, kb.Current(), vb.Current());
    PCD_LAUNCH_CHECK(ctx);
    int bits = 1; while (((int64_t)1 << bits) < c->n_cells && bits < 32) bits++;
    size_t tmp = 0;
    PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, kb, vb, (int)n, 0, bits, st));
    PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
    PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp, kb, vb, (int)n, 0, bits, st));
    return PCD_OK;
  };

  float cs = cell_size;
  if (cs > 0.0f) {
    float p2 = exp2f(roundf(log2f(cs)));       // snap to a power of two
    cs = p2;
    while (!setup(cs)) cs *= 2.0f;
    int rc = sort_keys(); if (rc) return rc;
  } else {
    // density-driven: aim for target_occ points per occupied cell (power-of-two steps)
    double lo_occ = 12.0, hi_occ = 48.0;
    if (const char* e = getenv("PCD_B200_OCC_LO")) lo_occ = atof(e);
    if (const char* e = getenv("PCD_B200_OCC_HI")) hi_occ = atof(e);
    cs = 0.125f;
    while (!setup(cs)) cs *= 2.0f;
    int last_dir = 0;
    for (int iter = 0; iter < 24; iter++) {
      int rc = sort_keys(); if (rc) return rc;
      unsigned long long* d_cnt = (unsigned long long*)ctx->misc.p;
      PCD_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, 8, st));
      k_count_unique<<<pcd_blocks(n, 256), 256, 0, st>>>(kb.Current(), n, d_cnt);
      PCD_LAUNCH_CHECK(ctx);
      unsigned long long uniq = 1;
      PCD_CUDA(ctx, cudaMemcpyAsync(&uniq, d_cnt, 8, cudaMemcpyDeviceToHost, st));
      PCD_CUDA(ctx, cudaStreamSynchronize(st));
      double occ = (double)n / (double)(uniq ? uniq : 1) * (double)xsub;   // per (y, z)-sized cell
      int dir = occ < lo_occ ? +1 : (occ >= hi_occ ? -1 : 0);
      if (dir == 0) break;
      if (last_dir != 0 && dir != last_dir) {          // overshoot: settle on the coarser cell
        if (dir < 0) break;                            // current cs is the coarser one
        cs *= 2.0f; setup(cs); rc = sort_keys(); if (rc) return rc; break;
      }
      float next = dir > 0 ? cs * 2.0f : cs * 0.5f;
      if (dir < 0 && (next < 1.0f / 1024.0f)) break;
      pcd::GridView saved = G; int64_t saved_cells = c->n_cells;
      if (!setup(next)) { G = saved; c->n_cells = saved_cells; break; }   // finer grid would not fit: keep current (already sorted)
      if (dir > 0 && uniq <= 1) { cs = next; rc = sort_keys(); if (rc) return rc; break; }
      cs = next; last_dir = dir;
    }
  }

  PCD_CUDA(ctx, cudaMalloc(&c->grid_pts, sizeof(float4) * (size_t)n));
  PCD_CUDA(ctx, cudaMalloc(&c->cell_start, sizeof(uint32_t) * (size_t)(c->n_cells + 1)));
  PCD_CUDA(ctx, cudaMemsetAsync(c->cell_start, 0, sizeof(uint32_t) * (size_t)(c->n_cells + 1), st));
  k_grid_gather<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)c->aos, kb.Current(), vb.Current(), n, c->grid_pts, c->cell_start);
  PCD_LAUNCH_CHECK(ctx);
  {
    size_t tmp = 0;
    // (n_cells + 1) <= 2e9 + 1 fits in int
    PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, c->cell_start, c->cell_start, (int)(c->n_cells + 1), st));
    PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
    PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, c->cell_start, c->cell_start, (int)(c->n_cells + 1), st));
  }
  {
    size_t words = (size_t)G.nbx * G.nby * G.nbz;
    PCD_CUDA(ctx, cudaMalloc(&c->occ, 8 * words));
    PCD_CUDA(ctx, cudaMemsetAsync(c->occ, 0, 8 * words, st));
    k_grid_occupancy<<<pcd_blocks(n, 256), 256, 0, st>>>(c->grid_pts, n, G, c->occ);
    PCD_LAUNCH_CHECK(ctx);
  }
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  G.pts = c->grid_pts; G.cell_start = c->cell_start; G.occ = c->occ;
  c->has_index = true;
  return PCD_OK;
}

int pcd_index_info(const pcd_cloud* c, float* cell_size, int32_t dims[3], float origin[3], int64_t* n_cells) {
  if (!c || !c->has_index) return PCD_E_INVALID;
  if (cell_size) *cell_size = c->grid.c;
  if (dims) { dims[0] = c->grid.dx; dims[1] = c->grid.dy; dims[2] = c->grid.dz; }
  if (origin) { origin[0] = c->grid.ox; origin[1] = c->grid.oy; origin[2] = c->grid.oz; }
  if (n_cells) *n_cells = c->n_cells;
  return PCD_OK;
}

namespace pcd {
// segments per cell (entry `runs` is the zero that closes the exclusive sum)
__global__ void k_seg_count(const uint32_t* __restrict__ start, int runs, uint32_t SEG, uint32_t* __restrict__ nseg) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > runs) return;
  nseg[c] = c == runs ? 0u : (start[c + 1] - start[c] + SEG - 1) / SEG;
}
__global__ void k_seg_fill(const uint32_t* __restrict__ start, const uint32_t* __restrict__ segoff, int runs, uint32_t SEG, uint2* __restrict__ seg) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= runs) return;
  uint32_t o = segoff[c], b = start[c];
  for (uint32_t e = start[c + 1]; b < e; b += SEG) seg[o++] = make_uint2((uint32_t)c, b);
}
}  // namespace pcd

// ---------------------------------------------------------------------------------------
// a9  PcdProj::BuildSubMap (src/lidar/pcd_projection.cc:223-255)
// ---------------------------------------------------------------------------------------
int pcd_submap_build(pcd_cloud* c, const pcd_proj_options* opt) {
  if (!c || !opt) return PCD_E_INVALID;
  pcd_ctx* ctx = c->ctx;
  if (!(opt->submap_length > 0) || !(opt->submap_width > 0) || !(opt->submap_height > 0))
    PCD_FAIL(ctx, PCD_E_INVALID, "sub-map cell sizes must be positive");
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  const int64_t n = c->n;
  // also after a build that failed half way: whatever it allocated is released before the next attempt
  cudaFree(c->sub_keys); cudaFree(c->sub_start); cudaFree(c->sub_pts); cudaFree(c->sub_seg);
  c->sub_keys = nullptr; c->sub_start = nullptr; c->sub_pts = nullptr; c->sub_seg = nullptr;
  c->sub_cells = 0; c->sub_nseg = 0; c->has_submap = false;
  c->opt = *opt;
  if (n == 0) {
    PCD_CUDA(ctx, cudaMalloc(&c->sub_keys, 16)); PCD_CUDA(ctx, cudaMalloc(&c->sub_start, 8)); PCD_CUDA(ctx, cudaMalloc(&c->sub_pts, 16));
    PCD_CUDA(ctx, cudaMemsetAsync(c->sub_start, 0, 8, st));
    c->sub_cells = 0; c->sub_nseg = 0; c->has_submap = true;
    return PCD_OK;
  }
  // bbox (recomputed so the sub-map does not depend on the NN index)
  PCD_CUDA(ctx, ctx->misc.ensure(256));
  int* d_bb = ctx->misc.as<int>();
  int h_bb[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
  PCD_CUDA(ctx, cudaMemcpyAsync(d_bb, h_bb, sizeof h_bb, cudaMemcpyHostToDevice, st));
  int nb = (int)((n + 1023) / 1024); if (nb > ctx->sm_count * 8) nb = ctx->sm_count * 8;
  k_bbox<<<nb, 256, 0, st>>>((const float4*)c->aos, n, d_bb);
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaMemcpyAsync(h_bb, d_bb, sizeof h_bb, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  float lo[3], hi[3];
  for (int d = 0; d < 3; d++) { lo[d] = ord2f(h_bb[d]); hi[d] = ord2f(h_bb[3 + d]); }
  const float lens[3] = {opt->submap_length, opt->submap_height, opt->submap_width};
  int k0[3];
  for (int d = 0; d < 3; d++) {
    if (!isfinite(lo[d]) || !isfinite(hi[d])) PCD_FAIL(ctx, PCD_E_INVALID, "cloud has non-finite coordinates");
    double a = round((double)(lo[d] / lens[d])), b = round((double)(hi[d] / lens[d]));
    if (b - a >= 2097151.0 || fabs(a) > 1e9 || fabs(b) > 1e9) PCD_FAIL(ctx, PCD_E_INVALID, "sub-map key range too large on axis %d", d);
    k0[d] = (int)a;
  }
  struct Scratch : pcd::DevBuf { ~Scratch() { release(); } } k64a, k64b;   // local 8 B keys, freed on every return path
  PCD_CUDA(ctx, k64a.ensure(8 * (size_t)n)); PCD_CUDA(ctx, k64b.ensure(8 * (size_t)n));
  PCD_CUDA(ctx, ctx->vals.ensure(4 * (size_t)n)); PCD_CUDA(ctx, ctx->vals_alt.ensure(4 * (size_t)n));
  cub::DoubleBuffer<unsigned long long> kb(k64a.as<unsigned long long>(), k64b.as<unsigned long long>());
  cub::DoubleBuffer<uint32_t> vb(ctx->vals.as<uint32_t>(), ctx->vals_alt.as<uint32_t>());
  k_sub_keys<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)c->aos, n, lens[0], lens[1], lens[2], k0[0], k0[1], k0[2], kb.Current(), vb.Current());
  PCD_LAUNCH_CHECK(ctx);
  size_t tmp = 0;
  PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, kb, vb, (int)n, 0, 63, st));
  PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
  PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp, kb, vb, (int)n, 0, 63, st));
  PCD_CUDA(ctx, cudaMalloc(&c->sub_pts, sizeof(float4) * (size_t)n));
  k_sub_gather<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)c->aos, vb.Current(), n, c->sub_pts);
  PCD_LAUNCH_CHECK(ctx);
  // run-length encode the sorted keys -> unique cells + counts
  unsigned long long* d_uniq = kb.Alternate();
  PCD_CUDA(ctx, ctx->keys.ensure(4 * (size_t)(n + 1)));
  uint32_t* d_counts = ctx->keys.as<uint32_t>();
  int* d_runs = (int*)ctx->misc.p;
  tmp = 0;
  PCD_CUDA(ctx, cub::DeviceRunLengthEncode::Encode(nullptr, tmp, kb.Current(), d_uniq, d_counts, d_runs, (int)n, st));
  PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
  PCD_CUDA(ctx, cub::DeviceRunLengthEncode::Encode(ctx->cub_tmp.p, tmp, kb.Current(), d_uniq, d_counts, d_runs, (int)n, st));
  int runs = 0;
  PCD_CUDA(ctx, cudaMemcpyAsync(&runs, d_runs, 4, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  c->sub_cells = runs;
  PCD_CUDA(ctx, cudaMalloc(&c->sub_keys, sizeof(int32_t) * 3 * (size_t)runs));
  PCD_CUDA(ctx, cudaMalloc(&c->sub_start, sizeof(uint32_t) * (size_t)(runs + 1)));
  PCD_CUDA(ctx, cudaMemsetAsync(d_counts + runs, 0, 4, st));
  tmp = 0;
  PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, d_counts, c->sub_start, runs + 1, st));
  PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
  PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, d_counts, c->sub_start, runs + 1, st));
  k_sub_decode<<<pcd_blocks(runs, 256), 256, 0, st>>>(d_uniq, runs, k0[0], k0[1], k0[2], c->sub_keys);
  PCD_LAUNCH_CHECK(ctx);
  // work list of the projection: every cell cut into segments of at most SUBMAP_SEG points
  uint32_t* d_nseg = d_counts;                               // the counts are consumed; reuse as segments per cell
  PCD_CUDA(ctx, ctx->vals.ensure(4 * (size_t)(runs + 1)));
  uint32_t* d_segoff = ctx->vals.as<uint32_t>();
  c->sub_seglen = pcd::SUBMAP_SEG;
  if (const char* e = getenv("PCD_B200_SUBMAP_SEG")) { int v = atoi(e); if (v >= 32 && v <= 65536) c->sub_seglen = (uint32_t)v; }   // tuning knob
  pcd::k_seg_count<<<pcd_blocks(runs + 1, 256), 256, 0, st>>>(c->sub_start, runs, c->sub_seglen, d_nseg);
  PCD_LAUNCH_CHECK(ctx);
  tmp = 0;
  PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp, d_nseg, d_segoff, runs + 1, st));
  PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
  PCD_CUDA(ctx, cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tmp, d_nseg, d_segoff, runs + 1, st));
  uint32_t nseg = 0;
  PCD_CUDA(ctx, cudaMemcpyAsync(&nseg, d_segoff + runs, 4, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  c->sub_nseg = nseg;
  PCD_CUDA(ctx, cudaMalloc(&c->sub_seg, sizeof(uint2) * (size_t)(nseg > 0 ? nseg : 1)));
  pcd::k_seg_fill<<<pcd_blocks(runs, 256), 256, 0, st>>>(c->sub_start, d_segoff, runs, c->sub_seglen, c->sub_seg);
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  k64a.release(); k64b.release();
  c->has_submap = true;
  return PCD_OK;
}

int64_t pcd_submap_cells(const pcd_cloud* c) { return (c && c->has_submap) ? c->sub_cells : -1; }

int pcd_cloud_load_ply(pcd_ctx* ctx, const char* path, const pcd_proj_options* opt, pcd_cloud** out) {
  if (!ctx || !path || !out) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_cloud_load_ply: bad argument");
  float* host = nullptr; int64_t n = 0;
  int rc = pcd_ply_read(path, &host, &n);
  if (rc) { ctx->err = pcd_last_error(nullptr); return rc; }
  rc = pcd_cloud_from_host(ctx, host, n, 1, out);
  pcd_free(host);
  if (rc) return rc;
  rc = pcd_index_build(*out, 0.0f);
  if (!rc && opt) rc = pcd_submap_build(*out, opt);
  if (rc) { pcd_cloud_destroy(*out); *out = nullptr; }
  return rc;
}

// N3 (src/controllers/bundle_adjustment.cc:123-124 re-loads and re-indexes the map on every refine): the parsed,
// axis-swapped, NaN-compacted cloud is kept in a binary sidecar next to the PLY (path + ".pcdb200", keyed by the source
// file's size, mtime and head / tail bytes).  A hit skips the PLY parse and PointCloudDirectionTrans; the two GPU
// indices are always rebuilt (tens of milliseconds, far less than reading a table of their size back).
int pcd_cloud_load_ply_cached(pcd_ctx* ctx, const char* path, const pcd_proj_options* opt, pcd_cloud** out, int* cache_hit) {
  if (!ctx || !path || !out) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_cloud_load_ply_cached: bad argument");
  if (cache_hit) *cache_hit = 0;
  *out = nullptr;
  uint64_t key = 0;
  int rc = pcd_file_key(path, &key);
  if (rc) { ctx->err = pcd_last_error(nullptr); return rc; }
  const std::string side = std::string(path) + ".pcdb200";
  float* host = nullptr; int64_t n = 0;
  if (pcd_sidecar_read(side.c_str(), key, 1, &host, &n) == PCD_OK) {
    rc = pcd_cloud_from_host(ctx, host, n, 0, out);            // already in the index space: no swap, no compaction
    pcd_free(host);
    if (rc) return rc;
    if (cache_hit) *cache_hit = 1;
  } else {
    rc = pcd_ply_read(path, &host, &n);
    if (rc) { ctx->err = pcd_last_error(nullptr); return rc; }
    rc = pcd_cloud_from_host(ctx, host, n, 1, out);
    pcd_free(host);
    if (rc) return rc;
    const int64_t m = (*out)->n;                              // best effort: a read-only directory just means no sidecar
    float* swapped = (float*)malloc(m > 0 ? 32 * (size_t)m : 32);
    if (swapped && pcd_cloud_download(*out, swapped) == PCD_OK) pcd_sidecar_write(side.c_str(), key, 1, swapped, m);
    free(swapped);
  }
  rc = pcd_index_build(*out, 0.0f);
  if (!rc && opt) rc = pcd_submap_build(*out, opt);
  if (rc) { pcd_cloud_destroy(*out); *out = nullptr; }
  return rc;
}

}  // extern "C"
