// nn.cu -- batched exact 1-NN on the uniform grid (a4/a5: Kdtree::GetClosestPoint,
// src/lidar/kdtree.cc:10-21; float32 L2 in FLANN's L2_Simple order (dx2+dy2)+dz2 on the
// float-cast query, src/lidar/ply.cc:92; ties -> lowest original index).
//
// Pipeline per batch:  k_query_keys -> cub radix sort (cell id, query id) -> k_nn_search -> k_nn_far.
// The grid is anisotropic: cells are c x c in (y, z) and c / xs along x.  A (y, z) row of cells
// is ONE contiguous range of the cell-sorted points, so any x interval of a row costs two
// cell_start loads, and the finer x resolution only narrows the interval that is scanned.
//
// k_nn_search: ONE THREAD PER QUERY, queries in cell-sorted order (neighbouring threads share
//   cells, so their candidate loads hit the same L1 lines / sectors).  Own cell first (gives a
//   distance bound), then the rows of the (2*K1+1)^2 (y, z) neighbourhood in order of increasing
//   distance; a row is skipped when its distance lower bound exceeds the current best, and its x
//   interval is clipped to the chord of the current best-distance ball.  Queries the K1 block
//   cannot certify (best distance > distance to the block faces: outliers far from every
//   surface) are pushed, warp-aggregated, to a device queue.
// k_nn_far: ONE WARP per queued query, best-first over the occupancy words of 4x4x4-cell blocks in
//   Chebyshev shells: the blocks of a shell are spread over the lanes (empty space costs one 8-byte
//   load per 64 cells), the cells of a block are bounded two per lane and a cell's points are scanned
//   32 at a time; only cells whose box distance beats the warp's current best are ever scanned.
// Exactness: a row / interval is skipped only when a conservative lower bound on its distance
// (cell geometry minus a float-rounding margin) exceeds the current best, so every point that
// could win or tie is examined; the winner is min over (d2 bits << 32 | original index).
#include "nn.cuh"
#include <cub/cub.cuh>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;
#include <math.h>
#include <stdlib.h>

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
// search radius rounded UP to float (INFINITY = unbounded, NaN = invalid radius).  Written in original
// order with coalesced stores; the search gathers it through the sorted order, one aligned 16-byte
// load per query instead of three doubles plus the radius source.
__global__ void k_query_keys(const double* __restrict__ q, RadiusSrc rs, int64_t n, GridView g, uint32_t* __restrict__ keys,
                             uint32_t* __restrict__ vals, float4* __restrict__ qrec) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float x = (float)q[3 * i], y = (float)q[3 * i + 1], z = (float)q[3 * i + 2];
  const double R = query_radius(rs, (uint32_t)i);
  float Rf = INFINITY;
  // The float ball must hold every point the reference's fp64 test |l - p| <= R can accept: the cast above moves the
  // query by at most |x| 2^-24 per axis, so the radius grows by the sum of the three before it is rounded up.
  if (R != INFINITY)
    Rf = (R < 0.0 || isnan(R)) ? __int_as_float(0x7fc00000)
                               : __double2float_ru(R + (fabs(q[3 * i]) + fabs(q[3 * i + 1]) + fabs(q[3 * i + 2])) * 5.9604644775390625e-8);
  qrec[i] = make_float4(x, y, z, Rf);
  if (keys) {
    int cx = clampi((int)floorf((x - g.ox) * g.inv_cx), 0, g.dx - 1);
    int cy = clampi((int)floorf((y - g.oy) * g.inv_c), 0, g.dy - 1);
    int cz = clampi((int)floorf((z - g.oz) * g.inv_c), 0, g.dz - 1);
    keys[i] = (uint32_t)(((int64_t)cz * g.dy + cy) * g.dx + cx);
    vals[i] = (uint32_t)i;
  }
}

// Launch shape and loop knobs, each measured on the bench scene (profiles/r02_nn_variants*.log; rebuild with
// tools/build_nn_variants.sh to repeat): 128-thread blocks at full occupancy (32 registers) for both kernels; the
// candidate loop unrolled by four in the search and rolled in the far kernel.
#ifndef NN_BLOCK
#define NN_BLOCK 128
#endif
constexpr int NN_THREADS = NN_BLOCK;
#ifndef NN_MIN_BLOCKS
#define NN_MIN_BLOCKS 16
#endif
#ifndef FAR_MIN_BLOCKS
#define FAR_MIN_BLOCKS 16  // 32 registers and spills, but the far kernel is latency-bound: 0.999 ms against 1.009 (12) / 1.057 (10) / 1.175 (8)
#endif
#ifndef FAR_DRAW
#define FAR_DRAW 1         // queue entries a group takes per draw (1 measured best: 1.16 ms vs 1.29 ms at 4)
#endif
#ifndef FAR_GROUP
#define FAR_GROUP 16      // lanes per far query (32 or 16)
#endif
#ifndef NN_UNROLL
#define NN_UNROLL 4        // unroll factor of the candidate loop in k_nn_search (8: 2.86 ms against 2.12)
#endif
#ifndef FAR_UNROLL
#define FAR_UNROLL 1       // ... and in k_nn_far (4: 1.08 ms against 1.04)
#endif
constexpr int NN_UNROLL_N = NN_UNROLL;
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

  // one candidate: exact float32 (dx^2 + dy^2) + dz^2 of the reference's metric, packed with the point's index
  __device__ __forceinline__ unsigned long long offer(const float4 p, unsigned long long bst) const {
    float ddx = qx - p.x, ddy = qy - p.y, ddz = qz - p.z;
    float d = __fadd_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)), __fmul_rn(ddz, ddz));
    unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | __float_as_uint(p.w);
    return key < bst ? key : bst;
  }

  // UNROLL of the candidate loop: 4 (with the compiler's predicated remainder) suits the thread-per-query search,
  // whose trip counts are the warp's maximum anyway; 1 (rolled) suits the far groups: a cell's run is split over
  // the group's lanes, most lanes make one or two trips and the remainder code is pure overhead.
  template <int UNROLL = NN_UNROLL_N>
  __device__ __forceinline__ void scan(const GridView& g, uint32_t b, uint32_t e, int first, int stride) {
    unsigned long long bst = best;
#pragma unroll UNROLL
    for (uint32_t j = b + first; j < e; j += stride) {
      PCD_DASSERT(j < g.n);
      bst = offer(__ldg(&g.pts[j]), bst);
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
  template <bool APPROX_SQRT = false>
  __device__ __forceinline__ bool chord(const GridView& g, float b2, float g2, int& x0, int& x1) const {
    float hr;
    if (APPROX_SQRT) {
      // one MUFU.SQRT (relative error <= 2^-23: GROW covers 8 ulp; a subnormal argument flushes to 0 and the margin alone
      // covers it) instead of the correctly rounded sequence with its fix-up branch: 0.99 -> 0.96 ms in k_nn_far, where a
      // chord is cut per visited cell; k_nn_search keeps sqrtf (2.09 -> 2.18 ms with the approximation: more spills)
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(hr) : "f"(b2 - g2));
    } else {
      hr = sqrtf(b2 - g2);
    }
    float hx = hr * GROW + g.margin;                       // INFINITY stays INFINITY
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
    PCD_DASSERT(x0 >= 0 && x1 < g.dx && (uint64_t)lin + (uint64_t)x1 + 1 <= (uint64_t)g.dx * g.dy * g.dz);
    const uint32_t b = __ldg(&g.cell_start[lin + x0]);
    const uint32_t e = __ldg(&g.cell_start[lin + x1 + 1]);
    PCD_DASSERT(b <= e && e <= g.n);
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

// ---- phase 2: far queries, one GROUP of FAR_GROUP lanes per query, best-first over the block occupancy words ----
// Every lane of a group holds the same query; the blocks of a Chebyshev shell are spread over the lanes (one
// occupancy word each), occupied blocks that can still hold a winner are then visited by the whole group in
// order of increasing box distance: the 64 cells of a block are bounded 64 / FAR_GROUP per lane, and a cell's
// points are scanned FAR_GROUP at a time.  Each lane keeps the best of the points IT has seen; the pruning bound is
// the group's (a min over the lanes' distance words, refreshed after every cell that improved some lane), and the
// (distance, index) pairs are reduced over the group once per query.  The kernel is bound by the latency of its dependent loads
// (occupancy word -> cell_start -> points); FAR_GROUP < 32 puts several independent queries into one warp.
template <int G>
struct FarSearch {
  const GridView& g;
  Query& Q;
  const cg::thread_block_tile<G>& tile;
  const int lane;
  float b2;                  // min(R2, best distance so far); group-uniform

  __device__ __forceinline__ FarSearch(const GridView& g_, Query& Q_, const cg::thread_block_tile<G>& t)
      : g(g_), Q(Q_), tile(t), lane((int)t.thread_rank()) { b2 = Q.bound2(); }

  __device__ __forceinline__ unsigned min_u32(unsigned v) const {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = min(v, tile.shfl_xor(v, o));
    return v;
  }
  __device__ __forceinline__ unsigned long long min_u64(unsigned long long v) const {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
      const unsigned long long w = tile.shfl_xor(v, o);
      v = w < v ? w : v;
    }
    return v;
  }

  // squared lower bound to the whole block (X, Y, Z)
  __device__ __forceinline__ float block_lb2(int X, int Y, int Z) const {
    const int sx = 4 * g.xs;
    float a = Query::gap(Q.px, X * sx, X * sx + sx - 1, g.cx, g.margin), b = Query::gap(Q.py, 4 * Y, 4 * Y + 3, g.c, g.margin),
          c = Query::gap(Q.pz, 4 * Z, 4 * Z + 3, g.c, g.margin);
    return (a * a + (b * b + c * c)) * SHRINK;
  }

  // whole group, uniform arguments: the points of one cubic cell, G at a time
  __device__ __forceinline__ void scan_cell(int X, int Y, int Z, int bit) {
    const int cxc = 4 * X + (bit & 3), cy = 4 * Y + ((bit >> 2) & 3), cz = 4 * Z + (bit >> 4);
    float gy = Query::gap(Q.py, cy, cy, g.c, g.margin), gz = Query::gap(Q.pz, cz, cz, g.c, g.margin);
    float g2 = (gy * gy + gz * gz) * SHRINK;
    if (!(g2 <= b2)) return;
    int x0 = cxc * g.xs, x1 = x0 + g.xs - 1;
    if (!Q.template chord<true>(g, b2, g2, x0, x1)) return;
    const uint32_t lin = ((uint32_t)cz * (uint32_t)g.dy + (uint32_t)cy) * (uint32_t)g.dx;
    PCD_DASSERT(x0 >= 0 && x1 < g.dx && cy >= 0 && cy < g.dy && cz >= 0 && cz < g.dz);
    const uint32_t b = __ldg(&g.cell_start[lin + x0]), e = __ldg(&g.cell_start[lin + x1 + 1]);
    PCD_DASSERT(b <= e && e <= g.n);
    if (b == e) return;
    // every lane keeps its own best; only the BOUND is the group's, and it is refreshed (a 32-bit min over the
    // distance words) only after a cell in which some lane improved.  The 64-bit (distance, index) min over the
    // lanes is taken once, at the end of run().
    const unsigned long long before = Q.best;
    Q.template scan<FAR_UNROLL>(g, b, e, lane, G);
    if (tile.any(Q.best != before)) b2 = fminf(Q.R2, __uint_as_float(min_u32((unsigned)(Q.best >> 32))));
  }

  // whole group, uniform arguments: the occupied cells of one block in order of increasing box distance
  // (64 / G cells per lane), stopping at the first cell that can no longer beat the group's bound
  __device__ __forceinline__ void scan_block(int X, int Y, int Z, unsigned long long w) {
    static_assert(G == 16 || G == 32, "cell bit = lane + G * h must keep x and y fixed per lane");
    constexpr int H = 64 / G;
    unsigned key[H];
    // cell `bit` = (z << 4) | (y << 2) | x of the block: a lane's cells share x = lane & 3 and y = (lane >> 2) & 3,
    // so the x and y gaps are computed once per lane and only the z gap changes with h
    const int cxc = 4 * X + (lane & 3), cy = 4 * Y + ((lane >> 2) & 3);
    const float ga = Query::gap(Q.px, cxc * g.xs, cxc * g.xs + g.xs - 1, g.cx, g.margin), gb = Query::gap(Q.py, cy, cy, g.c, g.margin);
    const float gab2 = ga * ga + gb * gb;
#pragma unroll
    for (int h = 0; h < H; h++) {
      const int bit = lane + G * h, cz = 4 * Z + (bit >> 4);
      const float gc = Query::gap(Q.pz, cz, cz, g.c, g.margin);
      // lb >= 0, so its bit pattern orders like the value; the low 6 bits carry the cell (lb rounded DOWN: conservative)
      key[h] = ((w >> bit) & 1ull) ? ((__float_as_uint((gab2 + gc * gc) * SHRINK) & ~63u) | (unsigned)bit) : 0xffffffffu;
    }
    for (;;) {
      unsigned k = key[0];
#pragma unroll
      for (int h = 1; h < H; h++) k = min(k, key[h]);
      k = min_u32(k);
      if (k == 0xffffffffu || !(__uint_as_float(k & ~63u) <= b2)) return;
      const int bit = (int)(k & 63u);
#pragma unroll
      for (int h = 0; h < H; h++) if (key[h] == k) key[h] = 0xffffffffu;
      scan_cell(X, Y, Z, bit);
    }
  }

  // Each lane offers one candidate block (has == false: none).  The group visits them in order of increasing
  // box distance and stops at the first one that can no longer beat the bound.
  __device__ __forceinline__ void visit(bool has, int X, int Y, int Z, unsigned long long w, float lb) {
    unsigned key = has ? ((__float_as_uint(lb) & ~31u) | (unsigned)lane) : 0xffffffffu;
    for (;;) {
      const unsigned k = min_u32(key);
      if (k == 0xffffffffu || !(__uint_as_float(k & ~31u) <= b2)) return;
      const int src = (int)(k & 31u);
      if (lane == src) key = 0xffffffffu;
      const int bX = tile.shfl(X, src), bY = tile.shfl(Y, src), bZ = tile.shfl(Z, src);
      const unsigned long long bw = tile.shfl(w, src);
      scan_block(bX, bY, bZ, bw);
    }
  }

  // this lane's block: bound it and, if it can matter, fetch its occupancy word
  __device__ __forceinline__ void offer(bool in_grid, int X, int Y, int Z) {
    float lb = INFINITY;
    unsigned long long w = 0;
    if (in_grid) {
      lb = block_lb2(X, Y, Z);
      PCD_DASSERT(X >= 0 && X < g.nbx && Y >= 0 && Y < g.nby && Z >= 0 && Z < g.nbz);
      if (lb <= b2) w = __ldg(&g.occ[((int64_t)Z * g.nby + Y) * g.nbx + X]);
    }
    visit(w != 0, X, Y, Z, w, lb);
  }

  // one rectangle of blocks: axis `fix` (0 = x, 1 = y, 2 = z) held at `v`, the other two over [a0, a1] x [b0, b1]
  // (a = the lower-numbered free axis).  Blocks are spread over the lanes, G at a time.
  __device__ __forceinline__ void rect(int fix, int v, int a0, int a1, int b0, int b1) {
    if (a0 > a1 || b0 > b1) return;
    const int na = a1 - a0 + 1, total = na * (b1 - b0 + 1);
    for (int base = 0; base < total; base += G) {
      const int t = base + lane;
      const int a = a0 + t % na, b = b0 + t / na;
      offer(t < total, fix == 0 ? v : a, fix == 1 ? v : (fix == 0 ? a : b), fix == 2 ? v : b);
    }
  }

  // every point outside the cube of shell kb is at least sqrt(returned) away (faces on the grid boundary do not
  // count); INFINITY when the cube covers the grid
  __device__ __forceinline__ float certified2(int bx, int by, int bz, int kb) const {
    const int sx = 4 * g.xs;
    float rc = INFINITY;
    const int x0 = (bx - kb) * sx, x1 = (bx + kb + 1) * sx - 1, y0 = 4 * (by - kb), y1 = 4 * (by + kb) + 3,
              z0 = 4 * (bz - kb), z1 = 4 * (bz + kb) + 3;
    if (x0 > 0) rc = fminf(rc, Q.px - (float)x0 * g.cx);
    if (x1 < g.dx - 1) rc = fminf(rc, (float)(x1 + 1) * g.cx - Q.px);
    if (y0 > 0) rc = fminf(rc, Q.py - (float)y0 * g.c);
    if (y1 < g.dy - 1) rc = fminf(rc, (float)(y1 + 1) * g.c - Q.py);
    if (z0 > 0) rc = fminf(rc, Q.pz - (float)z0 * g.c);
    if (z1 < g.dz - 1) rc = fminf(rc, (float)(z1 + 1) * g.c - Q.pz);
    if (rc == INFINITY) return INFINITY;
    rc = fmaxf(rc - g.margin, 0.0f);
    return rc * rc * SHRINK;
  }

  __device__ __forceinline__ void run() {
    const int bx = (Q.ix / g.xs) >> 2, by = Q.iy >> 2, bz = Q.iz >> 2;
    // shells 0 and 1 together: the 27 blocks around the query's block, G at a time, nearest first within a
    // round; the enumeration starts at the query's own block
#pragma unroll
    for (int base = 0; base < 27; base += G) {
      const int t = base + lane;
      const int u = t < 27 ? (t + 13) % 27 : 0;            // 13 = the query's own block
      const int dz = u / 9 - 1, dy = (u / 3) % 3 - 1, dx = u % 3 - 1;
      const int X = bx + dx, Y = by + dy, Z = bz + dz;
      offer(t < 27 && X >= 0 && X < g.nbx && Y >= 0 && Y < g.nby && Z >= 0 && Z < g.nbz, X, Y, Z);
    }
    const int kmax = max(g.nbx, max(g.nby, g.nbz));
    for (int kb = 1; kb <= kmax; kb++) {
      if (kb > 1) {
        // the shell of blocks at Chebyshev distance kb, clamped to the grid: two z faces (full), two y faces
        // (without the z-face rows), two x faces (without both)
        const int Y0 = max(by - kb, 0), Y1 = min(by + kb, g.nby - 1), X0 = max(bx - kb, 0), X1 = min(bx + kb, g.nbx - 1);
        const int Zi0 = max(bz - kb + 1, 0), Zi1 = min(bz + kb - 1, g.nbz - 1);      // z range without the z faces
        const int Yi0 = max(by - kb + 1, 0), Yi1 = min(by + kb - 1, g.nby - 1);
        if (bz - kb >= 0) rect(2, bz - kb, X0, X1, Y0, Y1);
        if (bz + kb < g.nbz) rect(2, bz + kb, X0, X1, Y0, Y1);
        if (by - kb >= 0) rect(1, by - kb, X0, X1, Zi0, Zi1);
        if (by + kb < g.nby) rect(1, by + kb, X0, X1, Zi0, Zi1);
        if (bx - kb >= 0) rect(0, bx - kb, Yi0, Yi1, Zi0, Zi1);
        if (bx + kb < g.nbx) rect(0, bx + kb, Yi0, Yi1, Zi0, Zi1);
      }
      const float rc2 = certified2(bx, by, bz, kb);
      if (rc2 == INFINITY || b2 <= rc2) break;
    }
    Q.best = min_u64(Q.best);
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

__global__ void k_nn_radius_filter(const double* __restrict__ q, RadiusSrc rs, int64_t n, const float4* __restrict__ aos,
                                   int32_t* __restrict__ idx, float* __restrict__ d2) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t k = idx[i];
  if (k < 0) return;
  const double R = query_radius(rs, (uint32_t)i);
  if (R == INFINITY) return;
  const float4 l = __ldg(&aos[2 * (int64_t)k]);
  const double x = (double)l.x - q[3 * i], y = (double)l.y - q[3 * i + 1], z = (double)l.z - q[3 * i + 2];
  const double dist = sqrt(__dadd_rn(__dmul_rn(x, x), __dadd_rn(__dmul_rn(y, y), __dmul_rn(z, z))));
  if (dist > R) { idx[i] = -1; if (d2) d2[i] = INFINITY; }
}

// phase 1: one thread per (cell-sorted) query; uncertified queries go to the far queue
// queue entry of an uncertified query: its state after phase 1 and its record, so that phase 2 starts from ONE
// aligned 32-byte load instead of the dependent chain far[e] -> qrec[i]
struct __align__(16) FarItem { unsigned long long best; uint32_t i; uint32_t pad; float4 rec; };

template <int K1>
__global__ void __launch_bounds__(NN_THREADS, NN_MIN_BLOCKS) k_nn_search(GridView g, const float4* __restrict__ qrec,
                                                          const uint32_t* __restrict__ order, int64_t n,
                                                          int32_t* __restrict__ idx_out, float* __restrict__ d2_out,
                                                          FarItem* __restrict__ far, unsigned int* __restrict__ far_count) {
  const int64_t slot = (int64_t)blockIdx.x * NN_THREADS + threadIdx.x;
  if (slot >= n) return;
  const uint32_t i = order ? order[slot] : (uint32_t)slot;
  Query Q;
  const bool valid = Q.load(g, qrec, i);
  const bool resolved = valid ? Q.search_block<K1>(g) : true;
  if (resolved) nn_write(Q.best, i, idx_out, d2_out);
  // warp-aggregated push of the uncertified queries
  const unsigned act = __activemask();
  const unsigned m = __ballot_sync(act, !resolved);
  if (m) {
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(far_count, (unsigned)__popc(m));
    base = __shfl_sync(act, base, leader);
    if (!resolved) {
      FarItem it; it.best = Q.best; it.i = i; it.pad = 0; it.rec = Q.record();
      PCD_DASSERT((int64_t)base + __popc(m) <= n);                  // the queue holds one entry per query at most
      far[base + __popc(m & ((1u << lane) - 1u))] = it;
    }
  }
}

// phase 2: one group of FAR_GROUP lanes per far query.  Persistent blocks (exactly one resident wave); the groups
// draw queue entries from an atomic counter, because the cost of a far query varies by an order of magnitude
// (a query with nothing inside its range walks every block of the ball) and a static split leaves a long tail.
__global__ void __launch_bounds__(NN_THREADS, FAR_MIN_BLOCKS) k_nn_far(GridView g, int32_t* __restrict__ idx_out, float* __restrict__ d2_out,
                                                       const FarItem* __restrict__ far, const unsigned int* __restrict__ far_count,
                                                       unsigned int* __restrict__ next) {
  const unsigned count = *far_count;
  const cg::thread_block_tile<FAR_GROUP> tile = cg::tiled_partition<FAR_GROUP>(cg::this_thread_block());
  const uint4* fq = reinterpret_cast<const uint4*>(far);
  for (;;) {
    // FAR_DRAW queue entries per draw
    unsigned e0 = 0;
    if (tile.thread_rank() == 0) e0 = atomicAdd(next, (unsigned)FAR_DRAW);
    e0 = tile.shfl(e0, 0);
    if (e0 >= count) return;
    const unsigned e1 = min(e0 + (unsigned)FAR_DRAW, count);
    for (unsigned e = e0; e < e1; e++) {
      const uint4 c0 = __ldg(&fq[2 * (size_t)e]), c1 = __ldg(&fq[2 * (size_t)e + 1]);
      const uint32_t i = c0.z;
      PCD_DASSERT(e < count);
      Query Q;
      Q.from_record(g, make_float4(__uint_as_float(c1.x), __uint_as_float(c1.y), __uint_as_float(c1.z), __uint_as_float(c1.w)));
      Q.best = ((unsigned long long)c0.y << 32) | c0.x;
      FarSearch<FAR_GROUP> F(g, Q, tile);
      F.run();
      if (tile.thread_rank() == 0) nn_write(Q.best, i, idx_out, d2_out);
    }
  }
}

int nn_search_dev(pcd_cloud* cloud, const double* d_q, const double* d_radius, double fixed_radius,
                  const int32_t* d_opt_num, const pcd_range_sweep* sweep, int64_t n, int32_t* d_idx, float* d2,
                  bool exact_radius, cudaStream_t st) {
  pcd_ctx* ctx = cloud->ctx;
  if (!cloud->has_index) PCD_FAIL(ctx, PCD_E_INVALID, "NN index not built (call pcd_index_build)");
  if (n == 0) return PCD_OK;
  if (n >= (int64_t)1 << 31) PCD_FAIL(ctx, PCD_E_INVALID, "batch too large (%lld queries; limit 2^31-1)", (long long)n);
  if (d_opt_num && !sweep) PCD_FAIL(ctx, PCD_E_INVALID, "opt_num given without a range sweep");
  const GridView& g = cloud->grid;
  const uint32_t* order = nullptr;
  const char* kenv = getenv("PCD_B200_NN_K1");      // tuning knobs (tools/explore_nn.py)
  const int K1 = kenv ? atoi(kenv) : 1;
  RadiusSrc rs;
  rs.radius = d_radius; rs.opt_num = d_opt_num; rs.fixed = fixed_radius;
  rs.max_r = sweep ? sweep->max_r : 0; rs.min_r = sweep ? sweep->min_r : 0; rs.drop = sweep ? sweep->drop : 0;
  PCD_CUDA(ctx, ctx->qrec.ensure(16 * (size_t)n));
  float4* qrec = ctx->qrec.as<float4>();
  {
    ProfScope ps(ctx, PCD_K_NN_KEYSORT, st);
    const bool sorted = n > 2048;
    if (sorted) {
      PCD_CUDA(ctx, ctx->keys.ensure(4 * (size_t)n)); PCD_CUDA(ctx, ctx->keys_alt.ensure(4 * (size_t)n));
      PCD_CUDA(ctx, ctx->vals.ensure(4 * (size_t)n)); PCD_CUDA(ctx, ctx->vals_alt.ensure(4 * (size_t)n));
    }
    cub::DoubleBuffer<uint32_t> kb(ctx->keys.as<uint32_t>(), ctx->keys_alt.as<uint32_t>());
    cub::DoubleBuffer<uint32_t> vb(ctx->vals.as<uint32_t>(), ctx->vals_alt.as<uint32_t>());
    k_query_keys<<<pcd_blocks(n, 256), 256, 0, st>>>(d_q, rs, n, g, sorted ? kb.Current() : nullptr, sorted ? vb.Current() : nullptr, qrec);
    PCD_LAUNCH_CHECK(ctx);
    if (sorted) {
      int bits = 1; while (((int64_t)1 << bits) < cloud->n_cells && bits < 32) bits++;
      // The order only has to make neighbouring lanes touch neighbouring cells: the lowest key bits (position
      // inside a short x stretch) may stay unsorted when that saves a whole 8-bit radix pass.
      int skip = 0;
      if (const char* e = getenv("PCD_B200_SORT_SKIP")) skip = atoi(e);
      else if (bits > 24 && bits <= 28) skip = bits - 24;
      if (skip < 0 || skip >= bits) skip = 0;
      size_t tmp = 0;
      PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp, kb, vb, (int)n, skip, bits, st));
      PCD_CUDA(ctx, ctx->cub_tmp.ensure(tmp));
      PCD_CUDA(ctx, cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tmp, kb, vb, (int)n, skip, bits, st));
      order = vb.Current();
    }
  }
  PCD_CUDA(ctx, ctx->far_q.ensure(sizeof(FarItem) * (size_t)n + 16));
  unsigned int* far_count = ctx->far_q.as<unsigned int>();
  FarItem* far = (FarItem*)(ctx->far_q.as<char>() + 16);
  const int blocks = (int)((n + NN_THREADS - 1) / NN_THREADS);
  {
    ProfScope ps(ctx, PCD_K_NN_SEARCH, st);
    PCD_CUDA(ctx, cudaMemsetAsync(far_count, 0, 8, st));    // queue length and the work counter of phase 2
    if (K1 == 2) k_nn_search<2><<<blocks, NN_THREADS, 0, st>>>(g, qrec, order, n, d_idx, d2, far, far_count);
    else k_nn_search<1><<<blocks, NN_THREADS, 0, st>>>(g, qrec, order, n, d_idx, d2, far, far_count);
    PCD_LAUNCH_CHECK(ctx);
  }
  {
    static int far_occ = 0;                                  // resident blocks per SM of k_nn_far
    if (far_occ == 0) {
      PCD_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&far_occ, k_nn_far, NN_THREADS, 0));
      if (far_occ < 1) far_occ = 1;
    }
    int far_blocks = ctx->sm_count * far_occ;                // one resident wave; groups pull work dynamically
    if ((int64_t)far_blocks * (NN_THREADS / FAR_GROUP) > n) far_blocks = (int)((n + NN_THREADS / FAR_GROUP - 1) / (NN_THREADS / FAR_GROUP));
    ProfScope psf(ctx, PCD_K_NN_FAR, st);
    k_nn_far<<<far_blocks, NN_THREADS, 0, st>>>(g, d_idx, d2, far, far_count, far_count + 1);
    PCD_LAUNCH_CHECK(ctx);
  }
  if (exact_radius && (d_radius || d_opt_num || fixed_radius != INFINITY)) {
    k_nn_radius_filter<<<pcd_blocks(n, 256), 256, 0, st>>>(d_q, rs, n, (const float4*)cloud->aos, d_idx, d2);
    PCD_LAUNCH_CHECK(ctx);
  }
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
  return pcd::nn_search_dev(cloud, d_q, d_radius, INFINITY, nullptr, nullptr, n, d_idx, d_d2, true, (cudaStream_t)stream);
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
  int rc = pcd::nn_search_dev(cloud, d_q, d_r, INFINITY, nullptr, nullptr, n, d_idx, d_d2, true, st);
  if (rc) return rc;
  if (idx) PCD_CUDA(ctx, cudaMemcpyAsync(idx, d_idx, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (d2) PCD_CUDA(ctx, cudaMemcpyAsync(d2, d_d2, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

}  // extern "C"
