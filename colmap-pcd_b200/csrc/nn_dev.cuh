// nn_dev.cuh -- device-side pieces shared by the search kernels (nn.cu: thread-per-query search over global memory and
// the far-query kernels; nn_tile.cu: tile-staged search): the query record, the per-query geometry (cell, gaps, chord,
// certification radius) and the result writer.  a4/a5: src/lidar/kdtree.cc:10-21, src/lidar/ply.cc:90-107.
#pragma once
#include "nn.cuh"
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

// Query record used by the search kernels: the float-cast query (the reference's cast, ply.cc:92) and the
// search radius rounded UP to float (INFINITY = unbounded, NaN = invalid radius).
__device__ __forceinline__ float4 query_record(const double* __restrict__ q, const RadiusSrc& rs, int64_t i) {
  float x = (float)q[3 * i], y = (float)q[3 * i + 1], z = (float)q[3 * i + 2];
  const double R = query_radius(rs, (uint32_t)i);
  float Rf = INFINITY;
  if (R != INFINITY) Rf = (R < 0.0 || isnan(R)) ? __int_as_float(0x7fc00000) : __double2float_ru(R);
  return make_float4(x, y, z, Rf);
}

constexpr float SHRINK = 0.999999f;   // applied to lower bounds before comparing with a float d2
constexpr float GROW = 1.000001f;

// Per-query search state.  All geometry is in metres relative to the grid origin.
struct Query {
  float qx, qy, qz;          // float-cast query (the reference's cast, ply.cc:92)
  float px, py, pz;          // q - origin
  int ix, iy, iz;            // own cell (clamped into the grid)
  float R2;                  // conservative float bound on radius^2 (INFINITY when unbounded)
  float rw;                  // the record's radius word (float radius rounded up, INFINITY, or NaN = invalid)
  unsigned long long best;   // (float bits of d2 << 32) | original index

  __device__ __forceinline__ float best_d2() const { return __uint_as_float((unsigned)(best >> 32)); }
  __device__ __forceinline__ float bound2() const { return fminf(best_d2(), R2); }

  __device__ __forceinline__ void scan(const GridView& g, uint32_t b, uint32_t e, int first, int stride) {
    unsigned long long bst = best;
    for (uint32_t j = b + first; j < e; j += stride) {
      float4 p = __ldg(&g.pts[j]);
      float ddx = qx - p.x, ddy = qy - p.y, ddz = qz - p.z;
      float d = __fadd_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)), __fmul_rn(ddz, ddz));
      unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | __float_as_uint(p.w);
      bst = key < bst ? key : bst;
    }
    best = bst;
  }

  // lower bound (metres) on the distance along one axis from coordinate p to cells [lo, hi] of edge c
  static __device__ __forceinline__ float gap(float p, int lo, int hi, float c, float margin) {
    float a = fmaxf((float)lo * c - p, p - (float)(hi + 1) * c);
    return fmaxf(a - margin, 0.0f);
  }

  // x interval [x0, x1] (cells of edge cx) clipped to the chord of the ball of squared radius b2 at
  // squared (y, z) offset g2, and to the grid
  __device__ __forceinline__ bool chord(const GridView& g, float b2, float g2, int& x0, int& x1) const {
    float hx = sqrtf(b2 - g2) * GROW + g.margin;          // INFINITY stays INFINITY
    x0 = max(max(x0, 0), (int)floorf((px - hx) * g.inv_cx));
    x1 = min(min(x1, g.dx - 1), (int)floorf((px + hx) * g.inv_cx));
    return x0 <= x1;
  }

  // squared radius (metres^2) within which the block x: +-kx cells, y/z: +-k cells around the own
  // cell is known to hold every cloud point; INFINITY when the block covers the grid
  __device__ __forceinline__ float certified2(const GridView& g, int kx, int k) const {
    float rc = INFINITY;
    if (ix - kx > 0) rc = fminf(rc, px - (float)(ix - kx) * g.cx);
    if (ix + kx < g.dx - 1) rc = fminf(rc, (float)(ix + kx + 1) * g.cx - px);
    if (iy - k > 0) rc = fminf(rc, py - (float)(iy - k) * g.c);
    if (iy + k < g.dy - 1) rc = fminf(rc, (float)(iy + k + 1) * g.c - py);
    if (iz - k > 0) rc = fminf(rc, pz - (float)(iz - k) * g.c);
    if (iz + k < g.dz - 1) rc = fminf(rc, (float)(iz + k + 1) * g.c - pz);
    if (rc == INFINITY) return INFINITY;
    rc = fmaxf(rc - g.margin, 0.0f);
    return rc * rc * SHRINK;
  }

  // loads the record of query `i`, derives its cell, radius bound and validity
  __device__ __forceinline__ bool load(const GridView& g, const float4* __restrict__ qrec, uint32_t i) {
    return from_record(g, __ldg(&qrec[i]));
  }
  __device__ __forceinline__ float4 record() const { return make_float4(qx, qy, qz, rw); }
  __device__ __forceinline__ bool from_record(const GridView& g, const float4 r) {
    qx = r.x; qy = r.y; qz = r.z; rw = r.w;
    px = qx - g.ox; py = qy - g.oy; pz = qz - g.oz;
    ix = clampi((int)floorf(px * g.inv_cx), 0, g.dx - 1);
    iy = clampi((int)floorf(py * g.inv_c), 0, g.dy - 1);
    iz = clampi((int)floorf(pz * g.inv_c), 0, g.dz - 1);
    R2 = r.w == INFINITY ? INFINITY : __fmul_ru(r.w, r.w) * 1.00001f + 1e-12f;     // conservative bound on radius^2
    bool valid = isfinite(qx) && isfinite(qy) && isfinite(qz) && !isnan(r.w) && g.n > 0;
    if (valid && R2 != INFINITY) {
      // no cloud point can be within R of a query farther than R from the grid box
      float bx = gap(px, 0, g.dx - 1, g.cx, g.margin), by = gap(py, 0, g.dy - 1, g.c, g.margin),
            bz = gap(pz, 0, g.dz - 1, g.c, g.margin);
      if ((bx * bx + by * by + bz * bz) * SHRINK > R2) valid = false;
    }
    best = ~0ull;
    return valid;
  }

  // One row (iy + DY, iz + DZ) of phase 1.  gy2 / gz2: squared axis gaps, NaN for rows outside the grid.
  __device__ __forceinline__ void row(const GridView& g, int kx, int dy, int dz, float gy2, float gz2) {
    const float g2 = (gy2 + gz2) * SHRINK;
    const float b2 = bound2();
    if (!(g2 <= b2)) return;                               // pruned, or outside the grid (NaN)
    int x0 = ix - kx, x1 = ix + kx;
    if (!chord(g, b2, g2, x0, x1)) return;
    const uint32_t lin = ((uint32_t)(iz + dz) * (uint32_t)g.dy + (uint32_t)(iy + dy)) * (uint32_t)g.dx;
    const uint32_t b = __ldg(&g.cell_start[lin + x0]);
    const uint32_t e = __ldg(&g.cell_start[lin + x1 + 1]);
    scan(g, b, e, 0, 1);
  }

  // phase 1: own cell, rest of the centre row, then rings 1..K1 of rows; returns true when certified
  template <int K1>
  __device__ __forceinline__ bool search_block(const GridView& g) {
    const int kx = K1 * g.xs;
    // squared axis gaps to the rows / slabs at offset -K1..K1 (NaN outside the grid -> never scanned)
    float gy2[2 * K1 + 1], gz2[2 * K1 + 1];
#pragma unroll
    for (int d = -K1; d <= K1; d++) {
      float a = gap(py, iy + d, iy + d, g.c, g.margin), b = gap(pz, iz + d, iz + d, g.c, g.margin);
      gy2[d + K1] = (iy + d < 0 || iy + d >= g.dy) ? __int_as_float(0x7fc00000) : a * a;
      gz2[d + K1] = (iz + d < 0 || iz + d >= g.dz) ? __int_as_float(0x7fc00000) : b * b;
    }
    // own cell, then the rest of the centre row inside the chord of the bound it gave
    const uint32_t lin0 = ((uint32_t)iz * (uint32_t)g.dy + (uint32_t)iy) * (uint32_t)g.dx;
    const uint32_t b0 = __ldg(&g.cell_start[lin0 + ix]), e0 = __ldg(&g.cell_start[lin0 + ix + 1]);
    scan(g, b0, e0, 0, 1);
    {
      int x0 = ix - kx, x1 = ix + kx;
      chord(g, bound2(), 0.0f, x0, x1);                    // always contains the own cell
      x0 = min(x0, ix); x1 = max(x1, ix);
      const uint32_t b = __ldg(&g.cell_start[lin0 + x0]), e = __ldg(&g.cell_start[lin0 + x1 + 1]);
      scan(g, b, b0, 0, 1);
      scan(g, e0, e, 0, 1);
    }
#pragma unroll
    for (int k = 1; k <= K1; k++) {
      // ring k of rows, nearest first: the four axis rows, then the rest of the ring
#pragma unroll
      for (int t = 0; t <= k; t++) {
        // offsets (+-k, +-t) and (+-t, +-k); t = 0 gives the axis rows, t = k the corners
#pragma unroll
        for (int s = 0; s < 8; s++) {
          const int a = (s & 1) ? k : -k, bsgn = (s & 2) ? t : -t;
          const int dy = (s & 4) ? a : bsgn, dz = (s & 4) ? bsgn : a;
          if (t == 0 && (s & 2)) continue;                 // +-0 duplicates
          if (t == k && (s & 4)) continue;                 // corners appear in both orientations
          row(g, kx, dy, dz, gy2[dy + K1], gz2[dz + K1]);
        }
      }
      const float rc2 = certified2(g, kx, k);
      if (best_d2() <= rc2 || R2 <= rc2) return true;
    }
    return false;
  }
};

// The search reports the nearest point inside the conservative float ball.  The reference's accept test
// |l - p| <= range in fp64 (src/optim/bundle_adjustment.cc:398-400, Eigen's (a - b).norm() = sqrt(x2 + (y2 + z2)))
// is applied afterwards in ORIGINAL query order, where p, the radius and idx stream in coalesced: by
// k_assoc_record on the association path, by k_nn_radius_filter for pcd_nn_batch with a radius.
__device__ __forceinline__ void nn_write(unsigned long long best, uint32_t i, int32_t* __restrict__ idx_out,
                                         float* __restrict__ d2_out) {
  int32_t idx = -1; float d2 = INFINITY;
  if (best != ~0ull) {
    idx = (int32_t)(uint32_t)(best & 0xffffffffu);
    d2 = __uint_as_float((unsigned)(best >> 32));
  }
  idx_out[i] = idx;
  if (d2_out) d2_out[i] = d2;
}

// queue entry of an uncertified query: its state after phase 1 and its record, so that phase 2 starts from ONE
// aligned 32-byte load instead of the dependent chain far[e] -> qrec[i]
struct __align__(16) FarItem { unsigned long long best; uint32_t i; uint32_t pad; float4 rec; };

// nn_tile.cu
int nn_tile_search(pcd_cloud* cloud, const double* d_q, const RadiusSrc& rs, int64_t n, int32_t* d_idx, float* d2,
                   FarItem* far, unsigned int* far_count, cudaStream_t st, bool* used);

}  // namespace pcd
