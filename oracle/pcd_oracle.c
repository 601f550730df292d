/*
 * pcd_oracle.c -- CPU restatement of Colmap-PCD's image-to-point-cloud association path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * PARITY UNPINNED: the reference (Wangshihu12/colmap-pcd @ 0eecc21) has no test, fixture or
 * golden vector for src/lidar, and its own sources cannot be compiled in this
 * image (PCL, FLANN, Eigen, Ceres, Boost, OpenCV are all absent).  This file is
 * therefore a hand restatement of the reference's arithmetic, each function citing the
 * reference file:line it follows.  NN arithmetic lives in un-vendored, un-pinned PCL->FLANN
 * (CMakeLists.txt:101,106); the published algorithm restated here is FLANN 1.9.1's
 * KDTreeSingleIndex (leaf_max_size 15, L2_Simple<float>, eps = 0).
 *
 * Conventions
 *  - a cloud is the reference's 32-byte AoS (src/lidar/pt_type.h:14-20): 8 floats per point,
 *    [x y z pad nx ny nz pad].
 *  - float code is compiled with -ffp-contract=off (no FMA), matching baseline x86-64.
 *  - Eigen fixed-size 3-vector reductions (dot, norm, matrix*vector coefficients) are
 *    evaluated as a0 + (a1 + a2): Eigen's redux_novec_unroller splits Length=3 as 1 + 2.
 *    This only matters for the float32 paths; fp64 results carry a 1e-5 tolerance.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * a2  PointCloudProcess::PointCloudDirectionTrans      src/lidar/ply.cc:33-57
 * (x,y,z) <- (-y,-z,x), same for the normal; rows with any NaN field are dropped, which
 * defines the index space every later index refers to.  Returns the surviving count.
 * ---------------------------------------------------------------------------------------- */
ORC_API int64_t orc_direction_trans(const float* in, int64_t n, float* out) {
  int64_t m = 0;
  for (int64_t i = 0; i < n; i++) {
    const float* p = in + 8 * i;
    float x = -p[1], y = -p[2], z = p[0];
    float nx = -p[5], ny = -p[6], nz = p[4];
    if (isnan(x) || isnan(y) || isnan(z) || isnan(nx) || isnan(ny) || isnan(nz)) continue;
    float* o = out + 8 * m;
    o[0] = x; o[1] = y; o[2] = z; o[3] = 0.0f;
    o[4] = nx; o[5] = ny; o[6] = nz; o[7] = 0.0f;
    m++;
  }
  return m;
}

/* ------------------------------------------------------------------------------------------
 * a4  Kdtree::GetClosestPoint          src/lidar/kdtree.cc:10-21  (k = 1 nearestKSearch)
 * Distance is FLANN L2_Simple<float>: result += diff*diff over x,y,z, i.e. (dx2+dy2)+dz2 in
 * float32 on the float-cast query (src/lidar/ply.cc:92).  Ties go to the lowest point index
 * (north_star's rule; FLANN keeps the first encountered).
 * ---------------------------------------------------------------------------------------- */
static inline float orc_d2(const float* p, float qx, float qy, float qz) {
  float dx = qx - p[0], dy = qy - p[1], dz = qz - p[2];
  float r = dx * dx;
  r = r + dy * dy;
  r = r + dz * dz;
  return r;
}

/* Exhaustive search: the definition of the NN result.  O(n*nq); OpenMP over queries. */
ORC_API void orc_nn_brute(const float* cloud, int64_t n, const double* q, int64_t nq,
                          int32_t* idx, float* d2) {
#pragma omp parallel for schedule(static)
  for (int64_t j = 0; j < nq; j++) {
    float qx = (float)q[3 * j], qy = (float)q[3 * j + 1], qz = (float)q[3 * j + 2];
    float best = INFINITY; int32_t bi = -1;
    if (!(isnan(qx) || isnan(qy) || isnan(qz))) {
      for (int64_t i = 0; i < n; i++) {
        float d = orc_d2(cloud + 8 * i, qx, qy, qz);
        if (d < best) { best = d; bi = (int32_t)i; }   /* ascending i: first hit = lowest index */
      }
    }
    idx[j] = bi; d2[j] = best;
  }
}

/* ---- exact single kd-tree, restating FLANN 1.9.1 KDTreeSingleIndex (leaf <= 15, middle split
 *      on the widest box side, balanced fallback).  a3 Kdtree::BuildMap src/lidar/kdtree.cc:5-8 */
typedef struct {
  int32_t left, right;      /* children (node ids), -1 for a leaf                           */
  int32_t lo, hi;           /* leaf: range into vind                                         */
  int32_t cut_dim;
  float div_low, div_high;
} orc_node;

typedef struct orc_kdtree {
  const float* cloud;       /* borrowed, AoS8 */
  int64_t n;
  int32_t* vind;
  float* reord;             /* xyz reordered by vind (FLANN reorder=true), 3 floats per point */
  orc_node* nodes;
  int32_t n_nodes, cap_nodes;
  int leaf_max;
  float bb_lo[3], bb_hi[3];
} orc_kdtree;

static int32_t orc_new_node(orc_kdtree* t) {
  if (t->n_nodes == t->cap_nodes) {
    t->cap_nodes = t->cap_nodes ? t->cap_nodes * 2 : 1024;
    t->nodes = (orc_node*)realloc(t->nodes, sizeof(orc_node) * (size_t)t->cap_nodes);
  }
  return t->n_nodes++;
}

static int32_t orc_divide(orc_kdtree* t, int32_t lo, int32_t hi, float* blo, float* bhi) {
  int32_t id = orc_new_node(t);
  int32_t count = hi - lo;
  if (count <= t->leaf_max) {
    orc_node nd; nd.left = nd.right = -1; nd.lo = lo; nd.hi = hi; nd.cut_dim = 0;
    nd.div_low = nd.div_high = 0.0f;
    /* leaf bounding box = exact extent of its points */
    for (int d = 0; d < 3; d++) { blo[d] = INFINITY; bhi[d] = -INFINITY; }
    for (int32_t k = lo; k < hi; k++) {
      const float* p = t->cloud + 8 * (int64_t)t->vind[k];
      for (int d = 0; d < 3; d++) { if (p[d] < blo[d]) blo[d] = p[d]; if (p[d] > bhi[d]) bhi[d] = p[d]; }
    }
    t->nodes[id] = nd;
    return id;
  }
  /* middleSplit_: widest side of the box, refined by actual data spread */
  float max_span = bhi[0] - blo[0]; int cut = 0;
  for (int d = 1; d < 3; d++) { float s = bhi[d] - blo[d]; if (s > max_span) { max_span = s; cut = d; } }
  float max_spread = -1.0f;
  for (int d = 0; d < 3; d++) {
    float span = bhi[d] - blo[d];
    if (span > (1.0f - 0.00001f) * max_span) {
      float mn = INFINITY, mx = -INFINITY;
      for (int32_t k = lo; k < hi; k++) {
        float v = t->cloud[8 * (int64_t)t->vind[k] + d];
        if (v < mn) mn = v; if (v > mx) mx = v;
      }
      if (mx - mn > max_spread) { cut = d; max_spread = mx - mn; }
    }
  }
  float split = (blo[cut] + bhi[cut]) / 2;
  {
    float mn = INFINITY, mx = -INFINITY;
    for (int32_t k = lo; k < hi; k++) {
      float v = t->cloud[8 * (int64_t)t->vind[k] + cut];
      if (v < mn) mn = v; if (v > mx) mx = v;
    }
    if (split < mn) split = mn; else if (split > mx) split = mx;
  }
  /* planeSplit: [lo,lim1) < split, [lim1,lim2) == split, [lim2,hi) > split */
  int32_t l = lo, r = hi - 1;
  for (;;) {
    while (l <= r && t->cloud[8 * (int64_t)t->vind[l] + cut] < split) l++;
    while (l <= r && t->cloud[8 * (int64_t)t->vind[r] + cut] >= split) r--;
    if (l > r) break;
    int32_t tmp = t->vind[l]; t->vind[l] = t->vind[r]; t->vind[r] = tmp; l++; r--;
  }
  int32_t lim1 = l; r = hi - 1;
  for (;;) {
    while (l <= r && t->cloud[8 * (int64_t)t->vind[l] + cut] <= split) l++;
    while (l <= r && t->cloud[8 * (int64_t)t->vind[r] + cut] > split) r--;
    if (l > r) break;
    int32_t tmp = t->vind[l]; t->vind[l] = t->vind[r]; t->vind[r] = tmp; l++; r--;
  }
  int32_t lim2 = l;
  int32_t half = lo + count / 2, mid;
  if (lim1 > half) mid = lim1; else if (lim2 < half) mid = lim2; else mid = half;
  if (mid == lo || mid == hi) mid = half;   /* degenerate (all equal on this axis) */

  float llo[3], lhi[3], rlo[3], rhi[3];
  memcpy(llo, blo, sizeof llo); memcpy(lhi, bhi, sizeof lhi); lhi[cut] = split;
  memcpy(rlo, blo, sizeof rlo); memcpy(rhi, bhi, sizeof rhi); rlo[cut] = split;
  int32_t L = orc_divide(t, lo, mid, llo, lhi);
  int32_t R = orc_divide(t, mid, hi, rlo, rhi);
  orc_node nd; nd.left = L; nd.right = R; nd.lo = lo; nd.hi = hi; nd.cut_dim = cut;
  nd.div_low = lhi[cut]; nd.div_high = rlo[cut];
  for (int d = 0; d < 3; d++) { blo[d] = llo[d] < rlo[d] ? llo[d] : rlo[d]; bhi[d] = lhi[d] > rhi[d] ? lhi[d] : rhi[d]; }
  t->nodes[id] = nd;
  return id;
}

ORC_API orc_kdtree* orc_kdtree_build(const float* cloud, int64_t n, int leaf_max) {
  orc_kdtree* t = (orc_kdtree*)calloc(1, sizeof(orc_kdtree));
  t->cloud = cloud; t->n = n; t->leaf_max = leaf_max > 0 ? leaf_max : 15;
  t->vind = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  for (int64_t i = 0; i < n; i++) t->vind[i] = (int32_t)i;
  for (int d = 0; d < 3; d++) { t->bb_lo[d] = INFINITY; t->bb_hi[d] = -INFINITY; }
  for (int64_t i = 0; i < n; i++)
    for (int d = 0; d < 3; d++) {
      float v = cloud[8 * i + d];
      if (v < t->bb_lo[d]) t->bb_lo[d] = v; if (v > t->bb_hi[d]) t->bb_hi[d] = v;
    }
  if (n > 0) {
    float lo[3], hi[3]; memcpy(lo, t->bb_lo, sizeof lo); memcpy(hi, t->bb_hi, sizeof hi);
    orc_divide(t, 0, (int32_t)n, lo, hi);
    memcpy(t->bb_lo, lo, sizeof lo); memcpy(t->bb_hi, hi, sizeof hi);
  }
  t->reord = (float*)malloc(sizeof(float) * 3 * (size_t)(n > 0 ? n : 1));
  for (int64_t k = 0; k < n; k++) {
    const float* p = cloud + 8 * (int64_t)t->vind[k];
    t->reord[3 * k] = p[0]; t->reord[3 * k + 1] = p[1]; t->reord[3 * k + 2] = p[2];
  }
  return t;
}

ORC_API void orc_kdtree_free(orc_kdtree* t) {
  if (!t) return;
  free(t->vind); free(t->reord); free(t->nodes); free(t);
}

typedef struct { float best; int32_t bi; } orc_best;

/* searchLevel: descend into the nearer child first, visit the other only if the box distance
 * can still beat (or tie) the current best.  The cut distances are accumulated in double and
 * compared with "<=", so no candidate that could win the (d2, index) order is ever pruned
 * (FLANN's own float bound may drop a <1 ulp closer point, SURVEY.md 8c divergence (i)). */
static void orc_search(const orc_kdtree* t, int32_t id, const float* q, double mind, double* dd, orc_best* b) {
  const orc_node* nd = &t->nodes[id];
  if (nd->left < 0) {
    for (int32_t k = nd->lo; k < nd->hi; k++) {
      const float* p = t->reord + 3 * (int64_t)k;
      float dx = q[0] - p[0], dy = q[1] - p[1], dz = q[2] - p[2];
      float r = dx * dx; r = r + dy * dy; r = r + dz * dz;
      int32_t gi = t->vind[k];
      if (r < b->best || (r == b->best && gi < b->bi)) { b->best = r; b->bi = gi; }
    }
    return;
  }
  int cd = nd->cut_dim;
  double v = q[cd];
  double diff1 = v - (double)nd->div_low, diff2 = v - (double)nd->div_high;
  int32_t near, far; double cut;
  if (diff1 + diff2 < 0) { near = nd->left; far = nd->right; cut = diff2 * diff2; }
  else { near = nd->right; far = nd->left; cut = diff1 * diff1; }
  orc_search(t, near, q, mind, dd, b);
  double save = dd[cd];
  double nm = mind + cut - save;
  dd[cd] = cut;
  if (nm * (1.0 - 1e-6) <= (double)b->best) orc_search(t, far, q, nm, dd, b);
  dd[cd] = save;
}

static void orc_kdtree_nn1(const orc_kdtree* t, const float* q, int32_t* idx, float* d2) {
  orc_best b; b.best = INFINITY; b.bi = -1;
  if (t->n > 0 && !(isnan(q[0]) || isnan(q[1]) || isnan(q[2]))) {
    double dd[3] = {0, 0, 0}, mind = 0;
    for (int d = 0; d < 3; d++) {
      if (q[d] < t->bb_lo[d]) { double e = (double)q[d] - t->bb_lo[d]; dd[d] = e * e; mind += dd[d]; }
      if (q[d] > t->bb_hi[d]) { double e = (double)q[d] - t->bb_hi[d]; dd[d] = e * e; mind += dd[d]; }
    }
    orc_search(t, 0, q, mind, dd, &b);
  }
  *idx = b.bi; *d2 = b.best;
}

ORC_API void orc_kdtree_nn(const orc_kdtree* t, const double* q, int64_t nq, int32_t* idx, float* d2, int nthreads) {
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
  for (int64_t j = 0; j < nq; j++) {
    float qf[3] = {(float)q[3 * j], (float)q[3 * j + 1], (float)q[3 * j + 2]};
    orc_kdtree_nn1(t, qf, idx + j, d2 + j);
  }
}

/* ------------------------------------------------------------------------------------------
 * a5  PointCloudProcess::SearchNearestNeiborByKdtree          src/lidar/ply.cc:90-107
 * double query -> float, NN, widen to [x y z nx ny nz] doubles; false on NaN or |n| < 1e-6.
 * ---------------------------------------------------------------------------------------- */
static int orc_widen(const float* cloud, int32_t i, double l[6]) {
  if (i < 0) return 0;
  const float* p = cloud + 8 * (int64_t)i;
  l[0] = p[0]; l[1] = p[1]; l[2] = p[2]; l[3] = p[4]; l[4] = p[5]; l[5] = p[6];
  for (int k = 0; k < 6; k++) if (isnan(l[k])) return 0;
  double nn = sqrt(l[3] * l[3] + (l[4] * l[4] + l[5] * l[5]));
  if (nn < 1e-6) return 0;
  return 1;
}

ORC_API int orc_search_nn(const orc_kdtree* t, const double* p, double* l_pt, int32_t* idx) {
  float qf[3] = {(float)p[0], (float)p[1], (float)p[2]}; float d2;
  orc_kdtree_nn1(t, qf, idx, &d2);
  return orc_widen(t->cloud, *idx, l_pt);
}

/* ------------------------------------------------------------------------------------------
 * a6-a8 association record.
 *   variant 0: BundleAdjustmentConfig::MatchClosestLidarPoint   src/optim/bundle_adjustment.cc:358-410
 *   variant 1: AdjustGlobalBundleByLidar loop body              src/sfm/incremental_mapper.cc:1430-1468
 *   variant 2: BundleAdjustmentController::Run loop body        src/controllers/bundle_adjustment.cc:141-183
 *   LidarPoint::Normalize / ComputeDist / ComputePointToPointDist / ComputeAngle
 *                                                               src/lidar/lidar_point.cc:21-50
 * type: 0 Proj, 1 Icp, 2 IcpGround (enum order of src/lidar/lidar_point.h:9).
 * accepted = 1 iff the reference would have stored the record.
 * ---------------------------------------------------------------------------------------- */
typedef struct {
  int32_t idx; uint8_t accepted, type, rgb[3];
  double xyz[3], abcd[4], dist, angle;
} orc_record;

static void orc_normalize(const double l[3], const double n[3], double abcd[4]) {
  double a = n[0], b = n[1], c = n[2];
  double norm = sqrt(pow(a, 2) + pow(b, 2) + pow(c, 2));   /* lidar_point.cc:43 */
  a = a / norm; b = b / norm; c = c / norm;
  double d = 0 - a * l[0] - b * l[1] - c * l[2];          /* :48, left to right */
  abcd[0] = a; abcd[1] = b; abcd[2] = c; abcd[3] = d;
}
static double orc_dist_plane(const double abcd[4], const double p[3]) {      /* :21-25 */
  return fabs((p[0] * abcd[0] + (p[1] * abcd[1] + p[2] * abcd[2])) + abcd[3]);
}
static double orc_dist_point(const double l[3], const double p[3]) {          /* :27-30 */
  double x = l[0] - p[0], y = l[1] - p[1], z = l[2] - p[2];
  return sqrt(x * x + (y * y + z * z));
}
static double orc_angle(const double abcd[4], const double l[3], const double p[3]) { /* :32-37 */
  double x = p[0] - l[0], y = p[1] - l[1], z = p[2] - l[2];
  double nrm = sqrt(x * x + (y * y + z * z));
  return fabs((abcd[0] * x + (abcd[1] * y + abcd[2] * z)) / nrm);
}

static void orc_make_record(const float* cloud, int32_t idx, const double* p, double range, int variant, orc_record* r) {
  memset(r, 0, sizeof *r);
  r->idx = idx; r->type = 255;
  double l[6];
  if (!orc_widen(cloud, idx, l)) { r->idx = idx; return; }
  orc_normalize(l, l + 3, r->abcd);
  r->xyz[0] = l[0]; r->xyz[1] = l[1]; r->xyz[2] = l[2];
  int ground = (fabs(l[4] / l[3]) > 10) && (fabs(l[4] / l[5]) > 10);   /* raw normal, bundle_adjustment.cc:381 */
  r->type = ground ? 2 : 1;
  if (ground) { r->rgb[0] = 255; r->rgb[1] = 255; r->rgb[2] = 0; }
  else if (variant == 0) { r->rgb[0] = 0; r->rgb[1] = 255; r->rgb[2] = 0; }
  else { r->rgb[0] = 0; r->rgb[1] = 0; r->rgb[2] = 255; }
  double dist = orc_dist_point(l, p);
  if (variant == 2) {
    double d2plane = orc_dist_plane(r->abcd, p);
    if (d2plane > 1 || dist > 2) return;          /* controllers/bundle_adjustment.cc:160 */
    r->accepted = 1;                               /* dist_/angle_ stay unset in the reference */
    r->dist = dist; r->angle = orc_angle(r->abcd, l, p);
    return;
  }
  if (dist > range) return;                        /* NaN dist is accepted, as in the reference */
  r->accepted = 1;
  r->dist = dist;
  r->angle = orc_angle(r->abcd, l, p);             /* set only by variant 0; reported for both */
}

/* range = max(min, max - opt_num*drop)      src/sfm/incremental_mapper.cc:1159-1163, :1423-1427 */
ORC_API void orc_search_range(const int32_t* opt_num, int64_t n, double max_r, double min_r, double drop, double* range) {
  for (int64_t i = 0; i < n; i++) {
    double r = max_r - opt_num[i] * drop;
    if (r <= min_r) r = min_r;
    range[i] = r;
  }
}

ORC_API void orc_associate(const orc_kdtree* t, const double* p, const double* range, int64_t nq, int variant,
                           int32_t* idx, uint8_t* accepted, uint8_t* type, uint8_t* rgb,
                           double* xyz, double* abcd, double* dist, double* angle, int nthreads) {
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
  for (int64_t j = 0; j < nq; j++) {
    float qf[3] = {(float)p[3 * j], (float)p[3 * j + 1], (float)p[3 * j + 2]};
    int32_t i; float d2;
    orc_kdtree_nn1(t, qf, &i, &d2);
    orc_record r;
    orc_make_record(t->cloud, i, p + 3 * j, range ? range[j] : INFINITY, variant, &r);
    if (idx) idx[j] = r.idx;
    if (accepted) accepted[j] = r.accepted;
    if (type) type[j] = r.type;
    if (rgb) memcpy(rgb + 3 * j, r.rgb, 3);
    if (xyz) memcpy(xyz + 3 * j, r.xyz, sizeof r.xyz);
    if (abcd) memcpy(abcd + 4 * j, r.abcd, sizeof r.abcd);
    if (dist) dist[j] = r.dist;
    if (angle) angle[j] = r.angle;
  }
}

/* Same record builder on externally supplied NN indices (used to check the fused GPU stage
 * independently of the search). */
ORC_API void orc_records_from_idx(const float* cloud, const int32_t* nn_idx, const double* p, const double* range,
                                  int64_t nq, int variant, uint8_t* accepted, uint8_t* type, uint8_t* rgb,
                                  double* xyz, double* abcd, double* dist, double* angle) {
  for (int64_t j = 0; j < nq; j++) {
    orc_record r;
    orc_make_record(cloud, nn_idx[j], p + 3 * j, range ? range[j] : INFINITY, variant, &r);
    if (accepted) accepted[j] = r.accepted;
    if (type) type[j] = r.type;
    if (rgb) memcpy(rgb + 3 * j, r.rgb, 3);
    if (xyz) memcpy(xyz + 3 * j, r.xyz, sizeof r.xyz);
    if (abcd) memcpy(abcd + 4 * j, r.abcd, sizeof r.abcd);
    if (dist) dist[j] = r.dist;
    if (angle) angle[j] = r.angle;
  }
}

/* ------------------------------------------------------------------------------------------
 * a17 BundleAdjustmentConfig::MatchVariablePoint2LidarPoint   src/optim/bundle_adjustment.cc:288-350
 * cand: n_cand x 6 doubles [xyz n] in track order.  Picks min |v.n|/(|n||v|), strict < from 360.
 * Returns 1 and fills the Proj record when a candidate was chosen.
 * ---------------------------------------------------------------------------------------- */
ORC_API int orc_match_proj(const double* p, const double* cand, int n_cand,
                           double* xyz, double* abcd, double* dist, double* angle_out) {
  double angle = 360; const double* best = 0;
  for (int k = 0; k < n_cand; k++) {
    const double* c = cand + 6 * k;
    double vx = p[0] - c[0], vy = p[1] - c[1], vz = p[2] - c[2];
    double nn = sqrt(c[3] * c[3] + (c[4] * c[4] + c[5] * c[5]));
    double vn = sqrt(vx * vx + (vy * vy + vz * vz));
    double t = fabs((vx * c[3] + (vy * c[4] + vz * c[5])) / (nn * vn));
    if (t < angle) { angle = t; best = c; }
  }
  if (angle == 360) return 0;
  orc_normalize(best, best + 3, abcd);
  xyz[0] = best[0]; xyz[1] = best[1]; xyz[2] = best[2];
  *dist = orc_dist_plane(abcd, p);
  *angle_out = orc_angle(abcd, best, p);
  return 1;
}

/* ------------------------------------------------------------------------------------------
 * a19 BundleAdjustmentLidarCostFunction::operator()       src/base/cost_functions.h:204-232
 * evaluated with ceres::Jet<double,3> rules (Ceres 2.1 jet.h): (f*g)' = f.a*g.v + f.v*g.a,
 * sqrt(f)' = f.v * (1/(2*sqrt(f.a))), T(w)*g = (w*g.a, w*g.v + 0*g.a).
 * J is NaN at u == 0 (0 * inf), reproduced on purpose.
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_resjac(const double* p, const double* abcd, const double* w, int64_t n, double* r, double* J) {
  for (int64_t i = 0; i < n; i++) {
    const double* P = p + 3 * i; const double* A = abcd + 4 * i;
    if (isnan(A[0]) || isnan(A[1]) || isnan(A[2]) || isnan(A[3])) {   /* no residual block: bundle_adjustment.cc:1003-1007 */
      r[i] = 0.0;
      if (J) J[3 * i] = J[3 * i + 1] = J[3 * i + 2] = 0.0;
      continue;
    }
    double s = ((P[0] * A[0] + P[1] * A[1]) + P[2] * A[2]) + A[3];
    double u = 0.0 - s;
    double f = u * u;
    double sq = sqrt(f);
    double inv = 1.0 / (2.0 * sq);
    r[i] = w[i] * sq;
    if (J) {
      for (int k = 0; k < 3; k++) {
        double du = 0.0 - A[k];
        double fv = u * du + du * u;
        double gv = fv * inv;
        J[3 * i + k] = w[i] * gv + 0.0 * sq;
      }
    }
  }
}

/* Same evaluation spread over host threads (what Ceres does with num_threads > 1,
 * src/optim/bundle_adjustment.cc:515-530); used only by the timed CPU baseline. */
ORC_API void orc_resjac_mt(const double* p, const double* abcd, const double* w, int64_t n, double* r, double* J, int nthreads) {
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
  const int64_t blk = 4096;
  const int64_t nb = (n + blk - 1) / blk;
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (int64_t b = 0; b < nb; b++) {
    int64_t lo = b * blk, m = (n - lo < blk) ? (n - lo) : blk;
    orc_resjac(p + 3 * lo, abcd + 4 * lo, w + lo, m, r + lo, J ? J + 3 * lo : 0);
  }
}

/* N2 Reconstruction::FilterLidarOutlier distance test     src/base/reconstruction.cc:771-804 */
ORC_API int64_t orc_filter_outliers(const double* p, const double* lidar_xyz, const uint8_t* type, int64_t n,
                                    double max_proj, double max_icp, uint8_t* erase) {
  int64_t cnt = 0;
  for (int64_t i = 0; i < n; i++) {
    double d = orc_dist_point(lidar_xyz + 3 * i, p + 3 * i);
    int e = (type[i] == 0) ? (d > max_proj) : (d > max_icp);
    erase[i] = (uint8_t)e; cnt += e;
  }
  return cnt;
}

/* ==========================================================================================
 * Depth projection ("z-buffer") -- src/lidar/pcd_projection.{h,cc}
 * ======================================================================================== */
typedef struct {
  double depth_image_scale;      /* pcd_projection.h:32 */
  int max_proj_scale, min_proj_scale;
  double min_proj_dist;
  float submap_length, submap_width, submap_height, choose_meter;
  double min_lidar_proj_dist;
} orc_proj_options;

typedef struct orc_submap {
  int64_t n; int64_t n_cells;
  int32_t* keys;      /* 3 per cell, lexicographic (x, y, z) order = std::map order (pcd_projection.h:75-85) */
  int64_t* start;     /* n_cells + 1 */
  int32_t* order;     /* point indices, insertion order inside a cell */
  float mn[3], mx[3]; /* global_map_{min,max}_* incl. the numeric_limits<float>::min() quirk (:173-178) */
  orc_proj_options opt;
  const float* cloud;
} orc_submap;

typedef struct { int32_t k[3]; int32_t i; } orc_keyed;
static int orc_keyed_cmp(const void* a, const void* b) {
  const orc_keyed* A = (const orc_keyed*)a; const orc_keyed* B = (const orc_keyed*)b;
  for (int d = 0; d < 3; d++) if (A->k[d] != B->k[d]) return A->k[d] < B->k[d] ? -1 : 1;
  return A->i < B->i ? -1 : (A->i > B->i);
}

/* a9 PcdProj::BuildSubMap + GetKeyType      src/lidar/pcd_projection.cc:223-255, pcd_projection.h:67-73
 * key = (round(x/length), round(y/height), round(z/width)) -- y pairs with height, z with width. */
ORC_API orc_submap* orc_submap_build(const float* cloud, int64_t n, const orc_proj_options* opt) {
  orc_submap* s = (orc_submap*)calloc(1, sizeof *s);
  s->n = n; s->opt = *opt; s->cloud = cloud;
  for (int d = 0; d < 3; d++) { s->mn[d] = FLT_MAX; s->mx[d] = FLT_MIN; }
  orc_keyed* kk = (orc_keyed*)malloc(sizeof(orc_keyed) * (size_t)(n > 0 ? n : 1));
  for (int64_t i = 0; i < n; i++) {
    const float* p = cloud + 8 * i;
    for (int d = 0; d < 3; d++) { if (p[d] < s->mn[d]) s->mn[d] = p[d]; if (p[d] > s->mx[d]) s->mx[d] = p[d]; }
    kk[i].k[0] = (int32_t)round(p[0] / opt->submap_length);
    kk[i].k[1] = (int32_t)round(p[1] / opt->submap_height);
    kk[i].k[2] = (int32_t)round(p[2] / opt->submap_width);
    kk[i].i = (int32_t)i;
  }
  qsort(kk, (size_t)n, sizeof(orc_keyed), orc_keyed_cmp);
  int64_t nc = 0;
  for (int64_t i = 0; i < n; i++) if (i == 0 || memcmp(kk[i].k, kk[i - 1].k, sizeof kk[i].k)) nc++;
  s->n_cells = nc;
  s->keys = (int32_t*)malloc(sizeof(int32_t) * 3 * (size_t)(nc > 0 ? nc : 1));
  s->start = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nc + 1));
  s->order = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  int64_t c = -1;
  for (int64_t i = 0; i < n; i++) {
    if (i == 0 || memcmp(kk[i].k, kk[i - 1].k, sizeof kk[i].k)) { c++; memcpy(s->keys + 3 * c, kk[i].k, sizeof kk[i].k); s->start[c] = i; }
    s->order[i] = kk[i].i;
  }
  s->start[nc] = n;
  free(kk);
  return s;
}
ORC_API void orc_submap_free(orc_submap* s) { if (!s) return; free(s->keys); free(s->start); free(s->order); free(s); }
ORC_API int64_t orc_submap_cells(const orc_submap* s) { return s->n_cells; }

/* Eigen::Quaterniond(w,x,y,z).toRotationMatrix(), not normalised  (pcd_projection.cc:17-18) */
static void orc_quat_to_rot(const double q[4], double R[9]) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  double twx = tx * w, twy = ty * w, twz = tz * w;
  double txx = tx * x, txy = ty * x, txz = tz * x;
  double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
  R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
}

static void orc_plane(const float a[3], const float b[3], const float c[3], float out[4]) {  /* pcd_projection.h:140-147 */
  float ab[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
  float ac[3] = {a[0] - c[0], a[1] - c[1], a[2] - c[2]};
  float n0 = ab[1] * ac[2] - ab[2] * ac[1];
  float n1 = ab[2] * ac[0] - ab[0] * ac[2];
  float n2 = ab[0] * ac[1] - ab[1] * ac[0];
  float d = -(n0 * a[0] + n1 * a[1] + n2 * a[2]);
  out[0] = n0; out[1] = n1; out[2] = n2; out[3] = d;
}

/* a10 SearchSubMap + QuadPyramid      src/lidar/pcd_projection.cc:258-297, pcd_projection.h:124-158
 * Output: planes[5][4], verts[5][3] (vertex, corner_1..4), all float32. */
ORC_API void orc_pyramid(const double qvec[4], const double tvec[3], const double cam[8], int cam_w, int cam_h,
                         const orc_proj_options* opt, float* planes, float* verts) {
  double R[9]; orc_quat_to_rot(qvec, R);
  double scale = opt->depth_image_scale;
  int img_h = (int)(cam_h * scale), img_w = (int)(cam_w * scale);
  double fx = cam[0] * scale, fy = cam[1] * scale, cx = cam[2] * scale, cy = cam[3] * scale;
  float Rwc[9];   /* rot_cw.cast<float>().transpose() */
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Rwc[3 * i + j] = (float)R[3 * j + i];
  float tc[3] = {(float)tvec[0], (float)tvec[1], (float)tvec[2]};
  float twc[3];
  for (int i = 0; i < 3; i++) twc[i] = (-Rwc[3 * i]) * tc[0] + ((-Rwc[3 * i + 1]) * tc[1] + (-Rwc[3 * i + 2]) * tc[2]);
  float xmin = (float)(-cx / fx), xmax = (float)((img_w - cx) / fx);
  float ymin = (float)(-cy / fy), ymax = (float)((img_h - cy) / fy);
  float cxy[4][2] = {{xmax, ymax}, {xmax, ymin}, {xmin, ymin}, {xmin, ymax}};
  float V[5][3];
  memcpy(V[0], twc, sizeof twc);
  for (int c = 0; c < 4; c++) {
    /* corner = x_bar_v*x + y_bar_v*y  (0*y, 0*x terms are exact), then center_v + corner */
    float v[3] = {cxy[c][0] + 0.0f * cxy[c][1], 0.0f * cxy[c][0] + cxy[c][1], 0.0f * cxy[c][0] + 0.0f * cxy[c][1]};
    v[0] = 0.0f + v[0]; v[1] = 0.0f + v[1]; v[2] = 1.0f + v[2];
    for (int i = 0; i < 3; i++) {
      float rv = Rwc[3 * i] * v[0] + (Rwc[3 * i + 1] * v[1] + Rwc[3 * i + 2] * v[2]);
      V[c + 1][i] = twc[i] + rv * opt->choose_meter;
    }
  }
  float* P = planes;
  orc_plane(V[1], V[4], V[3], P + 0);
  orc_plane(V[0], V[1], V[2], P + 4);
  orc_plane(V[0], V[2], V[3], P + 8);
  orc_plane(V[0], V[3], V[4], P + 12);
  orc_plane(V[0], V[4], V[1], P + 16);
  memcpy(verts, V, sizeof V);
}

/* a11 SearchImageMap     src/lidar/pcd_projection.cc:499-559
 * The reference walks every integer cell of bbox(frustum U map) +-1 and keeps occupied cells
 * whose centre passes the five plane tests; since that box always contains every occupied
 * key, this is "occupied cells passing the tests", emitted in key order.  sel: n_cells bytes. */
ORC_API int64_t orc_select_cells(const orc_submap* s, const float* planes, uint8_t* sel) {
  int64_t cnt = 0;
  for (int64_t c = 0; c < s->n_cells; c++) {
    float x = s->keys[3 * c] * s->opt.submap_length;
    float y = s->keys[3 * c + 1] * s->opt.submap_height;
    float z = s->keys[3 * c + 2] * s->opt.submap_width;
    int ok = 1;
    for (int k = 0; k < 5 && ok; k++) {
      const float* P = planes + 4 * k;
      float v = P[0] * x + P[1] * y; v = v + P[2] * z; v = v + P[3];
      ok = (v <= 0.0);
    }
    sel[c] = (uint8_t)ok; cnt += ok;
  }
  return cnt;
}

/* a13 PcdProj::DistortOpenCV     src/lidar/pcd_projection.cc:561-594 */
static void orc_distort(const double cam[8], double u, double v, double* ud, double* vd) {
  double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3], k1 = cam[4], k2 = cam[5], p1 = cam[6], p2 = cam[7];
  double x = (u - cx) / fx, y = (v - cy) / fy;
  double r2 = x * x + y * y;
  double Ra = 1. + k1 * r2 + k2 * r2 * r2;
  double Rb = 1;
  double Tx = 2. * p1 * x * y + p2 * (r2 + 2. * x * x);
  double Ty = p1 * (r2 + 2. * y * y) + 2. * p2 * x * y;
  double du = x * Ra * Rb + Tx, dv = y * Ra * Rb + Ty;
  *ud = du * fx + cx; *vd = dv * fy + cy;
}

/* The function-local statics of pcd_projection.cc:391-397, computed from the FIRST camera the
 * process projects (quirk F7; b_y uses raw min_proj_scale).  coef = [a_x b_x a_y b_y]. */
ORC_API void orc_splat_coef(const double cam[8], const orc_proj_options* opt, double* coef) {
  double s = opt->depth_image_scale, fx = cam[0], fy = cam[1];
  double max_x = (double)opt->max_proj_scale * (fx / 3039.0) * (s / 0.2);
  double max_y = (double)opt->max_proj_scale * (fy / 3039.0) * (s / 0.2);
  double min_x = (double)opt->min_proj_scale * (fx / 3039.0) * (s / 0.2);
  double min_y = (double)opt->min_proj_scale * (fy / 3039.0) * (s / 0.2);
  double a_x = (max_x - min_x) / (opt->min_proj_dist - (double)opt->choose_meter);
  double b_x = min_x - a_x * (double)opt->choose_meter;
  double a_y = (max_y - min_y) / (opt->min_proj_dist - (double)opt->choose_meter);
  double b_y = (double)opt->min_proj_scale - a_y * (double)opt->choose_meter;
  coef[0] = a_x; coef[1] = b_x; coef[2] = a_y; coef[3] = b_y;
}

/* a12 PcdProj::ImageMapProj      src/lidar/pcd_projection.cc:305-468, run SEQUENTIALLY (the
 * reference's OpenMP region races, F5): per pixel keep min (range, point index).
 * zbuf: img_w*img_h uint64 = (float bits of range << 32) | index, UINT64_MAX = empty.
 * feat_mask: NULL (dense: every covered pixel, the if_save_depth_image path) or img_w*img_h bytes.
 * counters[0] += points projected, [1] += pixel updates attempted (in-bounds covered pixels). */
ORC_API void orc_project(const orc_submap* s, const uint8_t* sel, const double qvec[4], const double tvec[3],
                         const double cam[8], int cam_w, int cam_h, const double coef[4],
                         const uint8_t* feat_mask, uint64_t* zbuf, int64_t* counters) {
  const orc_proj_options* o = &s->opt;
  double scale = o->depth_image_scale;
  int img_h = (int)(cam_h * scale), img_w = (int)(cam_w * scale);
  double Rd[9]; orc_quat_to_rot(qvec, Rd);
  float R[9]; for (int i = 0; i < 9; i++) R[i] = (float)Rd[i];
  float t[3] = {(float)tvec[0], (float)tvec[1], (float)tvec[2]};
  double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
  double max_x = (double)o->max_proj_scale * (fx / 3039.0) * (scale / 0.2);
  double max_y = (double)o->max_proj_scale * (fy / 3039.0) * (scale / 0.2);
  for (int64_t c = 0; c < s->n_cells; c++) {
    if (!sel[c]) continue;
    for (int64_t k = s->start[c]; k < s->start[c + 1]; k++) {
      int32_t gi = s->order[k];
      const float* pw = s->cloud + 8 * (int64_t)gi;
      float pc[3];
      for (int i = 0; i < 3; i++) pc[i] = (R[3 * i] * pw[0] + (R[3 * i + 1] * pw[1] + R[3 * i + 2] * pw[2])) + t[i];
      if (pc[2] < 0) continue;
      float dist = pc[2];
      if (dist < o->min_lidar_proj_dist) continue;   /* hoisted: u0/v0 are unused for these points */
      double u_ori = fx * (double)(pc[0] / pc[2]) + cx;
      double v_ori = fy * (double)(pc[1] / pc[2]) + cy;
      double ud, vd; orc_distort(cam, u_ori, v_ori, &ud, &vd);
      double ru = round(ud * scale), rv = round(vd * scale);
      if (!(fabs(ru) < 1e9) || !(fabs(rv) < 1e9)) continue;  /* int(round()) would be UB in the reference */
      int u0 = (int)ru, v0 = (int)rv;
      int sx, sy;
      if (dist <= o->min_proj_dist) { sx = (int)max_x; sy = (int)max_y; }
      else if (dist > o->min_proj_dist) { sx = (int)(coef[0] * dist + coef[1]); sy = (int)(coef[2] * dist + coef[3]); }
      else continue;
      float range = sqrtf(pc[0] * pc[0] + (pc[1] * pc[1] + pc[2] * pc[2]));
      uint32_t rb; memcpy(&rb, &range, 4);
      uint64_t packed = ((uint64_t)rb << 32) | (uint32_t)gi;
      if (counters) counters[0]++;
      for (int u = u0 - sx; u <= u0 + sx; u++)
        for (int v = v0 - sy; v <= v0 + sy; v++) {
          if (u < 0 || u >= img_w || v < 0 || v >= img_h) continue;
          int64_t px = (int64_t)v * img_w + u;
          if (counters) counters[1]++;
          if (feat_mask && !feat_mask[px]) continue;
          if (packed < zbuf[px]) zbuf[px] = packed;
        }
    }
  }
}

/* a14 feature pixel of a keypoint: (xy*scale).cast<int>() (truncation)   pcd_projection.cc:35,:122
 * returns -1 when out of bounds, else v*img_w+u. */
ORC_API void orc_feature_pixels(const double* xy, int64_t n, double scale, int cam_w, int cam_h, int64_t* px) {
  int img_h = (int)(cam_h * scale), img_w = (int)(cam_w * scale);
  for (int64_t i = 0; i < n; i++) {
    int u = (int)(xy[2 * i] * scale), v = (int)(xy[2 * i + 1] * scale);
    px[i] = (u < 0 || u >= img_w || v < 0 || v >= img_h) ? -1 : (int64_t)v * img_w + u;
  }
}

/* a15 ray/plane intersection of the init overload   src/lidar/pcd_projection.cc:183-205 */
ORC_API void orc_ray_plane(const float* cloud, int32_t gi, const double cam[8], double u, double v, double* xyz) {
  const float* p = cloud + 8 * (int64_t)gi;
  double a = p[4], b = p[5], c = p[6];
  double d = 0 - a * (double)p[0] - b * (double)p[1] - c * (double)p[2];
  double fx = cam[0], fy = cam[1], cx = cam[2], cy = cam[3];
  double z = -d / (a * (u - cx) / fx + b * (v - cy) / fy + c);
  xyz[0] = z * (u - cx) / fx; xyz[1] = z * (v - cy) / fy; xyz[2] = z;
}

ORC_API int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* ------------------------------------------------------------------------------------------
 * a16 BundleAdjustmentConfig::Project2Image, decision logic    src/optim/bundle_adjustment.cc:241-281
 * Walks one point's track the way the reference does, with lidar_searched_image_ids_ modelled as a growing list:
 * `cached` lists the image ids cached before the call (n_cached of them); selected[k] = 1 when element k's image is
 * handed to SetNewImage (and thereby cached for the elements after it).  pair_corrs[k] < 0: no such image pair.
 * ---------------------------------------------------------------------------------------- */
ORC_API int64_t orc_project2image(const uint32_t* track_image, int64_t n, uint32_t image_id, const int64_t* pair_corrs,
                                  int match_features_threshold, const uint32_t* cached, int64_t n_cached, uint8_t* selected) {
  uint32_t* cache = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n_cached + n + 1));
  int64_t nc = n_cached, count = 0;
  for (int64_t i = 0; i < n_cached; i++) cache[i] = cached[i];
  for (int64_t k = 0; k < n; k++) {
    selected[k] = 0;
    uint32_t image_id2 = track_image[k];
    if (image_id2 != image_id) {
      if (pair_corrs[k] >= 0) {                         /* reconstruction->ExistsImagePair(image_id, image_id2) */
        size_t corrs = (size_t)pair_corrs[k];
        if (corrs <= (size_t)match_features_threshold) continue;
      }
    }
    int found = 0;                                       /* lidar_searched_image_ids_.find(track_el.image_id) */
    for (int64_t i = 0; i < nc; i++) if (cache[i] == image_id2) { found = 1; break; }
    if (!found) { selected[k] = 1; count++; cache[nc++] = image_id2; }
  }
  free(cache);
  return count;
}
