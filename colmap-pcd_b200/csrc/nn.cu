// nn.cu -- batched exact 1-NN on the uniform grid (a4/a5: Kdtree::GetClosestPoint,
// src/lidar/kdtree.cc:10-21; float32 L2 in FLANN's L2_Simple order (dx2+dy2)+dz2 on the
// float-cast query, src/lidar/ply.cc:92; ties -> lowest original index).
//
// Pipeline per batch:  k_query_keys -> cub radix sort (cell id, query id) -> k_nn_search.
// k_nn_search: G lanes cooperate on one query.  A CTA owns a contiguous chunk of the
// cell-sorted queries (neighbouring queries share cells -> L1/L2 reuse).
//   phase 1: the 3x3x3 block around the query cell as 9 x-runs (x is the fastest grid axis, so
//            cells x-1..x+1 of one (y,z) row are ONE contiguous range of the sorted points:
//            2 cell_start loads per row, coalesced float4 point loads).
//   phase 2: queries that the block could not certify (best distance > distance to the block
//            faces) are compacted in shared memory and continue shell by shell (k = 2, 3, ...)
//            until certified, bounded by their search radius, or the grid is exhausted.
// Exactness: a row/shell is skipped only when a conservative lower bound on its distance
// (cell geometry minus a float-rounding margin) exceeds the current best.
#include "nn.cuh"
#include <cub/cub.cuh>
#include <math.h>

namespace pcd {

struct RadiusSrc {
  const double* radius;      // per query or NULL
  const int32_t* opt_num;    // per query GlobalOptNum or NULL
  double max_r, min_r, drop;
  double fixed;              // used when both are NULL (INFINITY = unbounded)
};

__device__ __forceinline__ double query_radius(const RadiusSrc& rs, uint32_t i) {
  if (rs.radius) return rs.radius[i];
  if (rs.opt_num) return sweep_range(rs.opt_num[i], rs.max_r, rs.min_r, rs.drop);
  return rs.fixed;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

__global__ void k_query_keys(const double* __restrict__ q, int64_t n, GridView g, uint32_t* __restrict__ keys,
                             uint32_t* __restrict__ vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float x = (float)q[3 * i], y = (float)q[3 * i + 1], z = (float)q[3 * i + 2];
  int cx = clampi((int)floorf((x - g.ox) * g.inv_c), 0, g.dx - 1);
  int cy = clampi((int)floorf((y - g.oy) * g.inv_c), 0, g.dy - 1);
  int cz = clampi((int)floorf((z - g.oz) * g.inv_c), 0, g.dz - 1);
  keys[i] = (uint32_t)(((int64_t)cz * g.dy + cy) * g.dx + cx);
  vals[i] = (uint32_t)i;
}

constexpr int NN_THREADS = 256;
constexpr int NN_MAX_CHUNK = 1024;

template <int G>
struct Searcher {
  const GridView& g;
  float qx, qy, qz, tx, ty, tz;
  int cx, cy, cz;
  float R2;                 // conservative float bound on radius^2 (INFINITY when unbounded)
  unsigned long long best;  // (float bits of d2 << 32) | original index
  unsigned gmask; int gl, gbase;

  __device__ __forceinline__ Searcher(const GridView& g_) : g(g_) {}

  __device__ __forceinline__ float best_d2() const { return __uint_as_float((unsigned)(best >> 32)); }

  __device__ __forceinline__ void scan(uint32_t b, uint32_t e) {
    unsigned long long bst = best;
    for (uint32_t j = b + gl; j < e; j += G) {
      float4 p = __ldg(&g.pts[j]);
      float ddx = qx - p.x, ddy = qy - p.y, ddz = qz - p.z;
      float d = __fadd_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)), __fmul_rn(ddz, ddz));
      unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | __float_as_uint(p.w);
      bst = key < bst ? key : bst;
    }
    best = bst;
  }

  // lower bound (in cells) on the distance along one axis from the query to cells [lo, hi]
  __device__ __forceinline__ float gap(float t, int lo, int hi) const {
    float a = fmaxf((float)lo - t, t - (float)(hi + 1));
    return fmaxf(a - g.margin_cells, 0.0f);
  }

  __device__ __forceinline__ float group_min_d2() const {
    float d = best_d2();
    if constexpr (G == 1) return d;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) d = fminf(d, __shfl_xor_sync(gmask, d, o, 32));
    return d;
  }

  __device__ __forceinline__ void group_reduce_best() {
    if constexpr (G == 1) return;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
      unsigned long long other = __shfl_xor_sync(gmask, best, o, 32);
      best = other < best ? other : best;
    }
  }

  // range of one entry: cells x0..x1 of row (y, z); returns false when empty/out of grid/pruned
  __device__ __forceinline__ bool entry_range(int x0, int x1, int y, int z, float bound2, uint32_t& b, uint32_t& e) const {
    if (y < 0 || y >= g.dy || z < 0 || z >= g.dz) return false;
    x0 = max(x0, 0); x1 = min(x1, g.dx - 1);
    if (x0 > x1) return false;
    float gx = gap(tx, x0, x1), gy = gap(ty, y, y), gz = gap(tz, z, z);
    float rm = (gx * gx + gy * gy + gz * gz) * (g.c * g.c) * 0.999999f;
    if (rm > bound2) return false;
    uint32_t lin = ((uint32_t)z * (uint32_t)g.dy + (uint32_t)y) * (uint32_t)g.dx;
    b = __ldg(&g.cell_start[lin + x0]);
    e = __ldg(&g.cell_start[lin + x1 + 1]);
    return b < e;
  }

  // scan every entry whose owner lane reported a non-empty range
  __device__ __forceinline__ void run_batch(bool ok, uint32_t b, uint32_t e) {
    if constexpr (G == 1) { if (ok) scan(b, e); return; }
    unsigned has = (__ballot_sync(gmask, ok) & gmask) >> gbase;
    while (has) {
      int l = __ffs(has) - 1; has &= has - 1;
      uint32_t bb = __shfl_sync(gmask, b, gbase + l, 32);
      uint32_t ee = __shfl_sync(gmask, e, gbase + l, 32);
      scan(bb, ee);
    }
  }

  // phase 1: the 3x3x3 block as nine full x-runs, centre row first
  __device__ __forceinline__ void scan_block1() {
    uint32_t b = 0, e = 0;
    // centre row: every lane loads the same two offsets (broadcast)
    if (entry_range(cx - 1, cx + 1, cy, cz, R2, b, e)) scan(b, e);
    float bound2 = fminf(group_min_d2(), R2);
#pragma unroll
    for (int base = 0; base < 8; base += G) {
      int r = base + gl;            // 0..7 -> the eight non-centre rows
      bool ok = false;
      if (r < 8) {
        int rr = r + (r >= 4);      // skip index 4 (dy = dz = 0)
        int dy = rr % 3 - 1, dz = rr / 3 - 1;
        ok = entry_range(cx - 1, cx + 1, cy + dy, cz + dz, bound2, b, e);
      }
      run_batch(ok, b, e);
      if (base + G < 8) bound2 = fminf(group_min_d2(), R2);
    }
    group_reduce_best();
  }

  // shell k >= 2: boundary rows as full runs, interior rows as their two end cells
  __device__ __forceinline__ void scan_shell(int k) {
    const int W = 2 * k + 1, nent = 2 * W * W;
    float bound2 = fminf(group_min_d2(), R2);
    for (int base = 0; base < nent; base += G) {
      int ent = base + gl;
      bool ok = false; uint32_t b = 0, e = 0;
      if (ent < nent) {
        int row = ent >> 1, seg = ent & 1;
        int dy = row % W - k, dz = row / W - k;
        bool boundary = (abs(dy) == k) || (abs(dz) == k);
        if (boundary) {
          if (seg == 0) ok = entry_range(cx - k, cx + k, cy + dy, cz + dz, bound2, b, e);
        } else {
          int x = seg == 0 ? cx - k : cx + k;
          if (x >= 0 && x < g.dx) ok = entry_range(x, x, cy + dy, cz + dz, bound2, b, e);
        }
      }
      unsigned any = (G == 1) ? (unsigned)ok : (__ballot_sync(gmask, ok) & gmask);
      if (any) {
        run_batch(ok, b, e);
        bound2 = fminf(group_min_d2(), R2);
      }
    }
    group_reduce_best();
  }

  // squared radius (metres^2) within which the (2k+1)^3 block is known to hold every point
  __device__ __forceinline__ float certified2(int k) const {
    float rc = INFINITY;
    if (cx - k > 0) rc = fminf(rc, tx - (float)(cx - k));
    if (cx + k < g.dx - 1) rc = fminf(rc, (float)(cx + k + 1) - tx);
    if (cy - k > 0) rc = fminf(rc, ty - (float)(cy - k));
    if (cy + k < g.dy - 1) rc = fminf(rc, (float)(cy + k + 1) - ty);
    if (cz - k > 0) rc = fminf(rc, tz - (float)(cz - k));
    if (cz + k < g.dz - 1) rc = fminf(rc, (float)(cz + k + 1) - tz);
    if (rc == INFINITY) return INFINITY;
    rc = fmaxf(rc - g.margin_cells, 0.0f) * g.c;
    return rc * rc * 0.999999f;
  }


  // loads query `i`, derives its cell, radius bound and validity
  __device__ __forceinline__ bool load_query(const double* __restrict__ q, const RadiusSrc& rs, uint32_t i,
                                             double& px, double& py, double& pz, double& R) {
    px = q[3 * (int64_t)i]; py = q[3 * (int64_t)i + 1]; pz = q[3 * (int64_t)i + 2];
    R = query_radius(rs, i);
    qx = (float)px; qy = (float)py; qz = (float)pz;
    tx = (qx - g.ox) * g.inv_c; ty = (qy - g.oy) * g.inv_c; tz = (qz - g.oz) * g.inv_c;
    cx = clampi((int)floorf(tx), 0, g.dx - 1);
    cy = clampi((int)floorf(ty), 0, g.dy - 1);
    cz = clampi((int)floorf(tz), 0, g.dz - 1);
    if (R == INFINITY) R2 = INFINITY;
    else {
      float Rf = __double2float_ru(fmax(R, 0.0));
      R2 = __fmul_ru(Rf, Rf) * 1.00001f + 1e-12f;
    }
    bool valid = isfinite(qx) && isfinite(qy) && isfinite(qz) && !(R < 0.0) && !isnan(R) && g.n > 0;
    if (valid && R2 != INFINITY) {
      // no cloud point can be within R of a query farther than R from the grid box
      float bx = gap(tx, 0, g.dx - 1), by = gap(ty, 0, g.dy - 1), bz = gap(tz, 0, g.dz - 1);
      if ((bx * bx + by * by + bz * bz) * (g.c * g.c) * 0.999999f > R2) valid = false;
    }
    return valid;
  }

  __device__ __forceinline__ bool done(int k) const {
    float rc2 = certified2(k);
    return best_d2() <= rc2 || R2 <= rc2;
  }
};

// final write: the reference's accept test |l - p| <= range in fp64 (src/optim/bundle_adjustment.cc:398-400,
// Eigen's (a - b).norm() = sqrt(x2 + (y2 + z2))) decides whether a radius-bounded query reports a hit.
__device__ __forceinline__ void nn_write(unsigned long long best, double px, double py, double pz, double R,
                                         const float4* __restrict__ aos, uint32_t i, int32_t* __restrict__ idx_out,
                                         float* __restrict__ d2_out) {
  int32_t idx = -1; float d2 = INFINITY;
  if (best != ~0ull) {
    idx = (int32_t)(uint32_t)(best & 0xffffffffu);
    d2 = __uint_as_float((unsigned)(best >> 32));
    if (R != INFINITY) {
      float4 l = __ldg(&aos[2 * (int64_t)idx]);
      double x = (double)l.x - px, y = (double)l.y - py, z = (double)l.z - pz;
      double dist = sqrt(x * x + (y * y + z * z));
      if (dist > R) { idx = -1; d2 = INFINITY; }
    }
  }
  idx_out[i] = idx;
  if (d2_out) d2_out[i] = d2;
}

template <int G>
__global__ void __launch_bounds__(NN_THREADS) k_nn_search(GridView g, const double* __restrict__ q, RadiusSrc rs,
                                                          const uint32_t* __restrict__ order, int64_t n, int chunk,
                                                          const float4* __restrict__ aos, int32_t* __restrict__ idx_out,
                                                          float* __restrict__ d2_out) {
  constexpr int GPC = NN_THREADS / G;   // groups (queries in flight) per CTA
  __shared__ uint32_t s_slot[NN_MAX_CHUNK];
  __shared__ unsigned long long s_best[NN_MAX_CHUNK];
  __shared__ int s_count;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();

  const int lane = threadIdx.x & 31;
  Searcher<G> S(g);
  S.gl = lane & (G - 1);
  S.gbase = lane & ~(G - 1);
  S.gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << S.gbase);
  const int group = threadIdx.x / G;
  const int64_t cta_lo = (int64_t)blockIdx.x * chunk;
  const int64_t cta_hi = min(cta_lo + (int64_t)chunk, n);

  // ---------------- phase 1 ----------------
  for (int64_t slot = cta_lo + group; slot < cta_hi; slot += GPC) {
    double px, py, pz, R;
    const uint32_t i = order ? order[slot] : (uint32_t)slot;
    bool valid = S.load_query(q, rs, i, px, py, pz, R);
    S.best = ~0ull;
    bool resolved = true;
    if (valid) {
      S.scan_block1();
      resolved = S.done(1);
    }
    if (S.gl == 0) {
      if (resolved) nn_write(S.best, px, py, pz, R, aos, i, idx_out, d2_out);
      else {
        int pos = atomicAdd(&s_count, 1);
        s_slot[pos] = (uint32_t)(slot - cta_lo);
        s_best[pos] = S.best;
      }
    }
  }
  __syncthreads();
  // ---------------- phase 2 ----------------
  const int pending = s_count;
  const int kmax = max(g.dx, max(g.dy, g.dz));
  for (int e = group; e < pending; e += GPC) {
    double px, py, pz, R;
    const int64_t slot = cta_lo + s_slot[e];
    const uint32_t i = order ? order[slot] : (uint32_t)slot;
    S.load_query(q, rs, i, px, py, pz, R);
    S.best = s_best[e];
    for (int k = 2; k <= kmax; k++) {
      S.scan_shell(k);
      if (S.done(k)) break;
    }
    if (S.gl == 0) nn_write(S.best, px, py, pz, R, aos, i, idx_out, d2_out);
  }
}

static int g_nn_group = 8;

template <int G>
static void launch_nn(int blocks, cudaStream_t st, const GridView& g, const double* q, const RadiusSrc& rs,
                      const uint32_t* order, int64_t n, int chunk, const float4* aos, int32_t* idx, float* d2) {
  k_nn_search<G><<<blocks, NN_THREADS, 0, st>>>(g, q, rs, order, n, chunk, aos, idx, d2);
}

int nn_search_dev(pcd_cloud* cloud, const double* d_q, const double* d_radius, double fixed_radius,
                  const int32_t* d_opt_num, const pcd_range_sweep* sweep, int64_t n, int32_t* d_idx, float* d2,
                  cudaStream_t st) {
  pcd_ctx* ctx = cloud->ctx;
  if (!cloud->has_index) PCD_FAIL(ctx, PCD_E_INVALID, "NN index not built (call pcd_index_build)");
  if (n == 0) return PCD_OK;
  if (n >= (int64_t)1 << 31) PCD_FAIL(ctx, PCD_E_INVALID, "batch too large (%lld queries; limit 2^31-1)", (long long)n);
  if (d_opt_num && !sweep) PCD_FAIL(ctx, PCD_E_INVALID, "opt_num given without a range sweep");
  const GridView& g = cloud->grid;
  const uint32_t* order = nullptr;
  const char* genv = getenv("PCD_B200_NN_GROUP");   // tuning knob (tools/explore_nn.py)
  const int group = genv ? atoi(genv) : 0;
  int G = group ? group : g_nn_group;
  if (n > 2048) {
    ProfScope ps(ctx, PCD_K_NN_KEYSORT, st);
    PCD_CUDA(ctx, ctx->keys.ensure(4 * (size_t)n)); PCD_CUDA(ctx, ctx->keys_alt.ensure(4 * (size_t)n));
    PCD_CUDA(ctx, ctx->vals.ensure(4 * (size_t)n)); PCD_CUDA(ctx, ctx->vals_alt.ensure(4 * (size_t)n));
    cub::DoubleBuffer<uint32_t> kb(ctx->keys.as<uint32_t>(), ctx->keys_alt.as<uint32_t>());
    cub::DoubleBuffer<uint32_t> vb(ctx->vals.as<uint32_t>(), ctx->vals_alt.as<uint32_t>());
    k_query_keys<<<pcd_blocks(n, 256), 256, 0, st>>>(d_q, n, g, kb.Current(), vb.Current());
    PCD_LAUNCH_CHECK(ctx);
    int bits = 1; while (((int64_t)1 << bits) < cloud->n_cells && bits < 32) bits++;
    size_t tmp = 0;
    PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, kb, vb, (int)n, 0, bits, st));
    PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
    PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp, kb, vb, (int)n, 0, bits, st));
    order = vb.Current();
  }
  RadiusSrc rs;
  rs.radius = d_radius; rs.opt_num = d_opt_num; rs.fixed = fixed_radius;
  rs.max_r = sweep ? sweep->max_r : 0; rs.min_r = sweep ? sweep->min_r : 0; rs.drop = sweep ? sweep->drop : 0;
  const int gpc = NN_THREADS / G;
  int64_t want = (n + (int64_t)ctx->sm_count * 8 - 1) / ((int64_t)ctx->sm_count * 8);
  int chunk = (int)((want + gpc - 1) / gpc) * gpc;
  if (chunk < gpc) chunk = gpc;
  if (chunk > NN_MAX_CHUNK) chunk = NN_MAX_CHUNK;
  int blocks = (int)((n + chunk - 1) / chunk);
  const float4* aos = (const float4*)cloud->aos;
  ProfScope ps(ctx, PCD_K_NN_SEARCH, st);
  switch (G) {
    case 1: launch_nn<1>(blocks, st, g, d_q, rs, order, n, chunk, aos, d_idx, d2); break;
    case 2: launch_nn<2>(blocks, st, g, d_q, rs, order, n, chunk, aos, d_idx, d2); break;
    case 4: launch_nn<4>(blocks, st, g, d_q, rs, order, n, chunk, aos, d_idx, d2); break;
    case 16: launch_nn<16>(blocks, st, g, d_q, rs, order, n, chunk, aos, d_idx, d2); break;
    case 32: launch_nn<32>(blocks, st, g, d_q, rs, order, n, chunk, aos, d_idx, d2); break;
    default: launch_nn<8>(blocks, st, g, d_q, rs, order, n, chunk, aos, d_idx, d2); break;
  }
  PCD_LAUNCH_CHECK(ctx);
  ctx->counters[0] += n;
  return PCD_OK;
}

}  // namespace pcd

extern "C" {

int pcd_nn_batch_dev(pcd_cloud* cloud, const double* d_q, const double* d_radius, int64_t n, int32_t* d_idx, float* d_d2,
                     void* stream) {
  if (!cloud) return PCD_E_INVALID;
  if (n < 0 || (n > 0 && (!d_q || !d_idx))) PCD_FAIL(cloud->ctx, PCD_E_INVALID, "pcd_nn_batch_dev: bad argument");
  DeviceGuard g(cloud->ctx->device);
  return pcd::nn_search_dev(cloud, d_q, d_radius, INFINITY, nullptr, nullptr, n, d_idx, d_d2, (cudaStream_t)stream);
}

int pcd_nn_batch(pcd_cloud* cloud, const double* q, const double* radius, int64_t n, int32_t* idx, float* d2) {
  if (!cloud) return PCD_E_INVALID;
  pcd_ctx* ctx = cloud->ctx;
  if (n < 0 || (n > 0 && !q)) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_nn_batch: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  size_t in_bytes = 24 * (size_t)n + (radius ? 8 * (size_t)n : 0);
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(in_bytes));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure(8 * (size_t)n));
  double* d_q = ctx->stage_in[0].as<double>();
  double* d_r = radius ? d_q + 3 * n : nullptr;
  int32_t* d_idx = ctx->stage_out[0].as<int32_t>();
  float* d_d2 = (float*)(d_idx + n);
  PCD_CUDA(ctx, cudaMemcpyAsync(d_q, q, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  if (radius) PCD_CUDA(ctx, cudaMemcpyAsync(d_r, radius, 8 * (size_t)n, cudaMemcpyHostToDevice, st));
  int rc = pcd::nn_search_dev(cloud, d_q, d_r, INFINITY, nullptr, nullptr, n, d_idx, d_d2, st);
  if (rc) return rc;
  if (idx) PCD_CUDA(ctx, cudaMemcpyAsync(idx, d_idx, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (d2) PCD_CUDA(ctx, cudaMemcpyAsync(d2, d_d2, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

}  // extern "C"
