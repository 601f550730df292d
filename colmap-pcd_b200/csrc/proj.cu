// proj.cu -- depth projection "z-buffer" of PcdProj (a10-a15, src/lidar/pcd_projection.cc).
//   SearchSubMap / QuadPyramid       :258-297, pcd_projection.h:124-158   (host, float32, no FMA)
//   SearchImageMap                   :499-559   cell centre inside 5 planes (device, fused)
//   ImageMapProj                     :305-468   project + distort + splat + per-pixel min range
//   DistortOpenCV                    :561-594   fp64 Brown-Conrady
//   SetNewImage (both overloads)     :13-220    feature pixels, winner gather, ray/plane
// Per pixel the winner is min over (range float bits << 32 | point index) by 64-bit atomicMin:
// min Euclidean range first, ties to the lowest point index -- the race-free reading of the
// reference's unsynchronised OpenMP loop.
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <stdlib.h>

namespace pcd {

struct ImgParams {
  float R[9], t[3];        // rot_cw.cast<float>(), t_cw.cast<float>()
  float planes[20];        // plane_0 .. plane_4
  double cam[8];           // unscaled OPENCV params
  double scale;
  double coef[4];          // a_x b_x a_y b_y (process-wide statics of the reference)
  double max_x, max_y;     // max_proj_scale_{x,y}
  double min_lidar, min_proj_dist;
  int w, h;                // scaled image size
  int words;               // mask words per row
};

// Eigen::Quaterniond(w,x,y,z).toRotationMatrix(), not normalised
static void quat_to_rot(const double q[4], double R[9]) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  double twx = tx * w, twy = ty * w, twz = tz * w;
  double txx = tx * x, txy = ty * x, txz = tz * x;
  double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

static void plane_of(const float a[3], const float b[3], const float c[3], float out[4]) {   // pcd_projection.h:140-147
  float ab[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
  float ac[3] = {a[0] - c[0], a[1] - c[1], a[2] - c[2]};
  float n0 = ab[1] * ac[2] - ab[2] * ac[1];
  float n1 = ab[2] * ac[0] - ab[0] * ac[2];
  float n2 = ab[0] * ac[1] - ab[1] * ac[0];
  float d = -(n0 * a[0] + n1 * a[1] + n2 * a[2]);
  out[0] = n0; out[1] = n1; out[2] = n2; out[3] = d;
}

// Host-side per-image setup.  Built with -ffp-contract=off; Eigen's 3-term reductions are a0 + (a1 + a2).
static void make_params(const pcd_image& im, const pcd_proj_options& o, const double coef[4], ImgParams& P) {
  double Rd[9]; quat_to_rot(im.qvec, Rd);
  for (int i = 0; i < 9; i++) P.R[i] = (float)Rd[i];
  for (int i = 0; i < 3; i++) P.t[i] = (float)im.tvec[i];
  for (int i = 0; i < 8; i++) P.cam[i] = im.cam[i];
  double scale = o.depth_image_scale;
  P.scale = scale;
  P.h = (int)(im.height * scale); P.w = (int)(im.width * scale);
  P.words = (P.w + 31) / 32;
  double fx = im.cam[0] * scale, fy = im.cam[1] * scale, cx = im.cam[2] * scale, cy = im.cam[3] * scale;
  float Rwc[9];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Rwc[3 * i + j] = P.R[3 * j + i];
  float twc[3];
  for (int i = 0; i < 3; i++) twc[i] = (-Rwc[3 * i]) * P.t[0] + ((-Rwc[3 * i + 1]) * P.t[1] + (-Rwc[3 * i + 2]) * P.t[2]);
  float xmin = (float)(-cx / fx), xmax = (float)((P.w - cx) / fx);
  float ymin = (float)(-cy / fy), ymax = (float)((P.h - cy) / fy);
  float cxy[4][2] = {{xmax, ymax}, {xmax, ymin}, {xmin, ymin}, {xmin, ymax}};
  float V[5][3];
  memcpy(V[0], twc, sizeof twc);
  for (int c = 0; c < 4; c++) {
    float v[3] = {cxy[c][0] + 0.0f * cxy[c][1], 0.0f * cxy[c][0] + cxy[c][1], 0.0f * cxy[c][0] + 0.0f * cxy[c][1]};
    v[0] = 0.0f + v[0]; v[1] = 0.0f + v[1]; v[2] = 1.0f + v[2];
    for (int i = 0; i < 3; i++) {
      float rv = Rwc[3 * i] * v[0] + (Rwc[3 * i + 1] * v[1] + Rwc[3 * i + 2] * v[2]);
      V[c + 1][i] = twc[i] + rv * o.choose_meter;
    }
  }
  plane_of(V[1], V[4], V[3], P.planes + 0);
  plane_of(V[0], V[1], V[2], P.planes + 4);
  plane_of(V[0], V[2], V[3], P.planes + 8);
  plane_of(V[0], V[3], V[4], P.planes + 12);
  plane_of(V[0], V[4], V[1], P.planes + 16);
  for (int i = 0; i < 4; i++) P.coef[i] = coef[i];
  P.max_x = (double)o.max_proj_scale * (im.cam[0] / 3039.0) * (scale / 0.2);
  P.max_y = (double)o.max_proj_scale * (im.cam[1] / 3039.0) * (scale / 0.2);
  P.min_lidar = o.min_lidar_proj_dist; P.min_proj_dist = o.min_proj_dist;
}

// feature pixel = (xy * scale).cast<int>() (truncation), bounds-checked (pcd_projection.cc:35-36)
// mask: one bit per pixel; summ: the OR of MASK_ROWS consecutive mask rows, so that a splat skips feature-free
// row groups with one load per word instead of MASK_ROWS.
constexpr int MASK_ROWS_LOG2 = 3, MASK_ROWS = 1 << MASK_ROWS_LOG2;
__global__ void k_feature_mask(const double* __restrict__ xy, int64_t n, ImgParams P, uint32_t* __restrict__ mask,
                               uint32_t* __restrict__ summ, int64_t* __restrict__ px_out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int u = (int)(xy[2 * i] * P.scale), v = (int)(xy[2 * i + 1] * P.scale);
  int64_t px = -1;
  if (!(u < 0 || u >= P.w || v < 0 || v >= P.h)) {
    px = (int64_t)v * P.w + u;
    atomicOr(&mask[(int64_t)v * P.words + (u >> 5)], 1u << (u & 31));
    atomicOr(&summ[(int64_t)(v >> MASK_ROWS_LOG2) * P.words + (u >> 5)], 1u << (u & 31));
  }
  px_out[i] = px;
}

// a12/a13 for one cloud point: camera frame, pinhole + DistortOpenCV, splat rectangle clipped to the image.
// Returns true when the rectangle is not empty; n_pts / n_px are the reference's counters.
__device__ __forceinline__ bool point_rect(const ImgParams& P, const float4 pw, unsigned long long& packed, int& ua, int& ub,
                                           int& va, int& vb, unsigned long long& n_pts, unsigned long long& n_px) {
  float pc[3];
#pragma unroll
  for (int i = 0; i < 3; i++) pc[i] = (P.R[3 * i] * pw.x + (P.R[3 * i + 1] * pw.y + P.R[3 * i + 2] * pw.z)) + P.t[i];
  if (pc[2] < 0) return false;
  float dist = pc[2];
  if ((double)dist < P.min_lidar) return false;
  double fx = P.cam[0], fy = P.cam[1], cx = P.cam[2], cy = P.cam[3];
  double u_ori = fx * (double)__fdiv_rn(pc[0], pc[2]) + cx;
  double v_ori = fy * (double)__fdiv_rn(pc[1], pc[2]) + cy;
  // DistortOpenCV
  double xn = (u_ori - cx) / fx, yn = (v_ori - cy) / fy;
  double r2 = xn * xn + yn * yn;
  double Ra = 1. + P.cam[4] * r2 + P.cam[5] * r2 * r2;
  double Tx = 2. * P.cam[6] * xn * yn + P.cam[7] * (r2 + 2. * xn * xn);
  double Ty = P.cam[6] * (r2 + 2. * yn * yn) + 2. * P.cam[7] * xn * yn;
  double ud = (xn * Ra * 1.0 + Tx) * fx + cx, vd = (yn * Ra * 1.0 + Ty) * fy + cy;
  double ru = round(ud * P.scale), rv = round(vd * P.scale);
  if (!(fabs(ru) < 1e9) || !(fabs(rv) < 1e9)) return false;
  int u0 = (int)ru, v0 = (int)rv;
  int sx, sy;
  if ((double)dist <= P.min_proj_dist) { sx = (int)P.max_x; sy = (int)P.max_y; }
  else if ((double)dist > P.min_proj_dist) { sx = (int)(P.coef[0] * (double)dist + P.coef[1]); sy = (int)(P.coef[2] * (double)dist + P.coef[3]); }
  else return false;
  float range = sqrtf(pc[0] * pc[0] + (pc[1] * pc[1] + pc[2] * pc[2]));
  packed = ((unsigned long long)__float_as_uint(range) << 32) | __float_as_uint(pw.w);
  n_pts++;
  ua = max(u0 - sx, 0); ub = min(u0 + sx, P.w - 1); va = max(v0 - sy, 0); vb = min(v0 + sy, P.h - 1);
  if (ua > ub || va > vb) return false;
  n_px += (unsigned long long)(ub - ua + 1) * (unsigned long long)(vb - va + 1);
  return true;
}

template <bool DENSE>
__global__ void __launch_bounds__(256) k_project(SubmapView sm, ImgParams P, const uint32_t* __restrict__ mask,
                                                 const uint32_t* __restrict__ summ, unsigned long long* __restrict__ zbuf, unsigned long long* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  unsigned long long n_pts = 0, n_px = 0, n_cells = 0;
  for (int64_t s = warp; s < sm.n_seg; s += nwarps) {
    const uint2 sg = __ldg(&sm.seg[s]);
    const int64_t c = sg.x;
    PCD_DASSERT(c >= 0 && c < sm.n_cells && sg.y >= sm.start[c] && sg.y < sm.start[c + 1]);
    // a11: cell centre (idx*len, idy*height, idz*width) against the five planes, float32, left to right
    float x = (float)sm.keys[3 * c] * sm.len, y = (float)sm.keys[3 * c + 1] * sm.height, z = (float)sm.keys[3 * c + 2] * sm.width;
    bool in = true;
#pragma unroll
    for (int k = 0; k < 5; k++) {
      float v = P.planes[4 * k] * x + P.planes[4 * k + 1] * y;
      v = v + P.planes[4 * k + 2] * z;
      v = v + P.planes[4 * k + 3];
      in = in && (v <= 0.0f);
    }
    if (!in) continue;
    const uint32_t b = sg.y, e = min(b + sm.seg_len, sm.start[c + 1]);
    if (lane == 0 && b == sm.start[c]) n_cells++;            // a cell counts once, with its first segment
    for (uint32_t j0 = b; j0 < e; j0 += 32) {                // warp-uniform trip count
      const uint32_t j = j0 + lane;
      unsigned long long packed = 0;
      int ua = 0, ub = -1, va = 0, vb = -1;
      const bool rect = j < e && point_rect(P, __ldg(&sm.pts[j]), packed, ua, ub, va, vb, n_pts, n_px);
      if (DENSE) {
        // the warp splats its points one after the other, lanes side by side along the pixel rows: one warp-wide
        // atomic touches 32 consecutive 8-byte words instead of 32 rows
        unsigned todo = __ballot_sync(0xffffffffu, rect);
        while (todo) {
          const int src = __ffs(todo) - 1; todo &= todo - 1;
          const unsigned long long pk = __shfl_sync(0xffffffffu, packed, src);
          const int a0 = __shfl_sync(0xffffffffu, ua, src), a1 = __shfl_sync(0xffffffffu, ub, src);
          const int r0 = __shfl_sync(0xffffffffu, va, src), r1 = __shfl_sync(0xffffffffu, vb, src);
          const int W = a1 - a0 + 1, tot = W * (r1 - r0 + 1);
          for (int t = lane; t < tot; t += 32) {
            const int dv = t / W, du = t - dv * W;
            PCD_DASSERT(r0 + dv >= 0 && r0 + dv < P.h && a0 + du >= 0 && a0 + du < P.w);
            unsigned long long* px = zbuf + (int64_t)(r0 + dv) * P.w + (a0 + du);
            if (pk < *px) atomicMin(px, pk);               // plain pre-read: a stale value is never below the true one (values only fall)
          }
        }
      } else if (rect) {
        const int wa = ua >> 5, wb = ub >> 5;
        for (int vg = va >> MASK_ROWS_LOG2; vg <= (vb >> MASK_ROWS_LOG2); vg++) {
          const uint32_t* srow = summ + (int64_t)vg * P.words;
          for (int wd = wa; wd <= wb; wd++) {
            uint32_t any = __ldg(&srow[wd]);
            if (!any) continue;
            int lo = max(ua - (wd << 5), 0), hi = min(ub - (wd << 5), 31);
            const uint32_t m = (hi == 31 ? 0xffffffffu : ((1u << (hi + 1)) - 1u)) & ~((1u << lo) - 1u);
            if (!(any & m)) continue;
            const int v1 = min(vb, (vg << MASK_ROWS_LOG2) + MASK_ROWS - 1);
            for (int v = max(va, vg << MASK_ROWS_LOG2); v <= v1; v++) {
              uint32_t bits = __ldg(&mask[(int64_t)v * P.words + wd]) & m;
              while (bits) {
                int bpos = __ffs(bits) - 1; bits &= bits - 1;
                PCD_DASSERT(v >= 0 && v < P.h && (wd << 5) + bpos >= 0 && (wd << 5) + bpos < P.w);
                unsigned long long* px = zbuf + (int64_t)v * P.w + (wd << 5) + bpos;
                if (packed < *px) atomicMin(px, packed);
              }
            }
          }
        }
      }
    }
  }
  if (stats) {
    for (int o = 16; o > 0; o >>= 1) {
      n_pts += __shfl_down_sync(0xffffffffu, n_pts, o);
      n_px += __shfl_down_sync(0xffffffffu, n_px, o);
    }
    if (lane == 0) {
      if (n_cells) atomicAdd(&stats[0], n_cells);
      if (n_pts) atomicAdd(&stats[1], n_pts);
      if (n_px) atomicAdd(&stats[2], n_px);
    }
  }
}

// Reference point for the z-buffer's roofline: 64-bit atomicMin on a buffer the size of one z-buffer, no two
// threads on one word at a time, lanes of a warp on 32 consecutive words (the access pattern of the dense splat),
// every operation a real update (values fall with each sweep).
__global__ void k_atomic_sweep(unsigned long long* __restrict__ buf, int64_t words, int sweeps) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int s = 0; s < sweeps; s++) {
    const unsigned long long v = ((unsigned long long)(0x7f000000u - (unsigned)s) << 32) | (unsigned)s;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < words; i += stride) atomicMin(&buf[i], v);
  }
}

__global__ void k_gather_winners(const int64_t* __restrict__ px, int64_t n, const unsigned long long* __restrict__ zbuf, int64_t npx,
                                 int32_t* __restrict__ widx, float* __restrict__ wrange, unsigned long long* __restrict__ stats) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool hit = false;
  if (i < n) {
    int32_t idx = -1; float r = INFINITY;
    int64_t p = px[i];
    if (p >= 0) {
      PCD_DASSERT(p < npx);
      unsigned long long z = zbuf[p];
      if (z != ~0ull) { idx = (int32_t)(uint32_t)(z & 0xffffffffu); r = __uint_as_float((unsigned)(z >> 32)); hit = true; }
    }
    widx[i] = idx;
    if (wrange) wrange[i] = r;
  }
  unsigned m = __ballot_sync(0xffffffffu, hit);
  if (stats && (threadIdx.x & 31) == 0 && m) atomicAdd(&stats[3], (unsigned long long)__popc(m));
}

// a15: z = -d / (a(u-cx)/fx + b(v-cy)/fy + c), unscaled undistorted pinhole (pcd_projection.cc:183-205)
__global__ void k_ray_plane(const float4* __restrict__ aos, const double* __restrict__ xy, const int32_t* __restrict__ widx, int64_t n,
                            double fx, double fy, double cx, double cy, uint8_t* __restrict__ matched, double* __restrict__ xyz) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t gi = widx[i];
  double X = 0, Y = 0, Z = 0; uint8_t ok = 0;
  if (gi >= 0) {
    float4 p = __ldg(&aos[2 * (int64_t)gi]), nn = __ldg(&aos[2 * (int64_t)gi + 1]);
    double a = nn.x, b = nn.y, c = nn.z;
    double d = 0 - a * (double)p.x - b * (double)p.y - c * (double)p.z;
    double u = xy[2 * i], v = xy[2 * i + 1];
    Z = -d / (a * (u - cx) / fx + b * (v - cy) / fy + c);
    X = Z * (u - cx) / fx; Y = Z * (v - cy) / fy;
    ok = 1;
  }
  matched[i] = ok;
  xyz[3 * i] = X; xyz[3 * i + 1] = Y; xyz[3 * i + 2] = Z;
}

static int project_one(pcd_cloud* c, const pcd_image* img, const double coef[4], const double* d_feat, int64_t n_feat,
                       int32_t* d_widx, float* d_wrange, bool dense, unsigned long long* d_stats, cudaStream_t st,
                       unsigned long long** zbuf_out, ImgParams* P_out) {
  pcd_ctx* ctx = c->ctx;
  ImgParams P; make_params(*img, c->opt, coef, P);
  if (P.w <= 0 || P.h <= 0 || (int64_t)P.w * P.h > ((int64_t)1 << 28)) PCD_FAIL(ctx, PCD_E_INVALID, "bad scaled image size %d x %d", P.w, P.h);
  const int64_t npx = (int64_t)P.w * P.h;
  PCD_CUDA(ctx, ctx->zbuf.ensure(8 * (size_t)npx));
  const size_t mask_words = (size_t)P.words * ((size_t)P.h + (size_t)((P.h + MASK_ROWS - 1) >> MASK_ROWS_LOG2));   // mask + row-group summary
  PCD_CUDA(ctx, ctx->sel_buf.ensure(4 * mask_words + 8 * (size_t)(n_feat > 0 ? n_feat : 1) + 64));
  unsigned long long* zb = ctx->zbuf.as<unsigned long long>();
  uint32_t* mask = ctx->sel_buf.as<uint32_t>();
  uint32_t* summ = mask + (size_t)P.words * P.h;
  int64_t* d_px = (int64_t*)((char*)ctx->sel_buf.p + ((4 * mask_words + 63) & ~(size_t)63));
  PCD_CUDA(ctx, cudaMemsetAsync(zb, 0xff, 8 * (size_t)npx, st));
  PCD_CUDA(ctx, cudaMemsetAsync(mask, 0, 4 * mask_words, st));
  if (n_feat > 0) {
    k_feature_mask<<<pcd_blocks(n_feat, 256), 256, 0, st>>>(d_feat, n_feat, P, mask, summ, d_px);
    PCD_LAUNCH_CHECK(ctx);
  }
  SubmapView sm;
  sm.keys = c->sub_keys; sm.start = c->sub_start; sm.pts = c->sub_pts; sm.n_cells = c->sub_cells;
  sm.seg = c->sub_seg; sm.n_seg = c->sub_nseg; sm.seg_len = c->sub_seglen;
  sm.len = c->opt.submap_length; sm.height = c->opt.submap_height; sm.width = c->opt.submap_width;
  if (c->sub_cells > 0 && (dense || n_feat > 0)) {
    int64_t warps = c->sub_nseg;
    int blocks = (int)((warps + 7) / 8);
    int cap = ctx->sm_count * 16; if (blocks > cap) blocks = cap;
    ProfScope ps(ctx, PCD_K_PROJECT, st);
    if (dense) k_project<true><<<blocks, 256, 0, st>>>(sm, P, mask, summ, zb, d_stats);
    else k_project<false><<<blocks, 256, 0, st>>>(sm, P, mask, summ, zb, d_stats);
    PCD_LAUNCH_CHECK(ctx);
  }
  if (n_feat > 0 && d_widx) {
    k_gather_winners<<<pcd_blocks(n_feat, 256), 256, 0, st>>>(d_px, n_feat, zb, npx, d_widx, d_wrange, d_stats);
    PCD_LAUNCH_CHECK(ctx);
  }
  if (zbuf_out) *zbuf_out = zb;
  if (P_out) *P_out = P;
  return PCD_OK;
}

}  // namespace pcd

extern "C" {

int pcd_splat_coef(const double cam[8], const pcd_proj_options* o, double coef[4]) {
  if (!cam || !o || !coef) return PCD_E_INVALID;
  double s = o->depth_image_scale, fx = cam[0], fy = cam[1];
  double max_x = (double)o->max_proj_scale * (fx / 3039.0) * (s / 0.2);
  double max_y = (double)o->max_proj_scale * (fy / 3039.0) * (s / 0.2);
  double min_x = (double)o->min_proj_scale * (fx / 3039.0) * (s / 0.2);
  double min_y = (double)o->min_proj_scale * (fy / 3039.0) * (s / 0.2);
  double a_x = (max_x - min_x) / (o->min_proj_dist - (double)o->choose_meter);
  double b_x = min_x - a_x * (double)o->choose_meter;
  double a_y = (max_y - min_y) / (o->min_proj_dist - (double)o->choose_meter);
  double b_y = (double)o->min_proj_scale - a_y * (double)o->choose_meter;   // quirk F7: raw min_proj_scale
  coef[0] = a_x; coef[1] = b_x; coef[2] = a_y; coef[3] = b_y;
  return PCD_OK;
}

int pcd_atomic_peak(pcd_ctx* ctx, int64_t words, int sweeps, double* atomics_per_s) {
  if (!ctx || !atomics_per_s || words <= 0 || sweeps <= 0) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_atomic_peak: bad argument");
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->zbuf.ensure(8 * (size_t)words));
  unsigned long long* buf = ctx->zbuf.as<unsigned long long>();
  cudaEvent_t a = nullptr, b = nullptr;
  PCD_CUDA(ctx, cudaEventCreate(&a));
  cudaError_t e = cudaEventCreate(&b);
  if (e != cudaSuccess) { cudaEventDestroy(a); PCD_CUDA(ctx, e); }
  const int blocks = ctx->sm_count * 8;
  float best = 0.0f;
  for (int rep = 0; rep < 4 && e == cudaSuccess; rep++) {       // first repetition warms up; best of the rest
    e = cudaMemsetAsync(buf, 0xff, 8 * (size_t)words, st);
    if (e == cudaSuccess) e = cudaEventRecord(a, st);
    pcd::k_atomic_sweep<<<blocks, 256, 0, st>>>(buf, words, sweeps);
    ctx->launches++;
    if (e == cudaSuccess) e = cudaEventRecord(b, st);
    if (e == cudaSuccess) e = cudaEventSynchronize(b);
    float ms = 0.0f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, a, b);
    if (e == cudaSuccess && rep > 0 && (best == 0.0f || ms < best)) best = ms;
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  cudaEventDestroy(a); cudaEventDestroy(b);
  PCD_CUDA(ctx, e);
  *atomics_per_s = (double)words * sweeps / (1e-3 * (double)best);
  return PCD_OK;
}

int pcd_project_image(pcd_cloud* c, const pcd_image* img, const double coef[4], const double* feat_xy, int64_t n_feat,
                      int32_t* winner_idx, float* winner_range, uint64_t* dense_zbuf, int64_t* stats) {
  if (!c) return PCD_E_INVALID;
  pcd_ctx* ctx = c->ctx;
  if (!img || !coef || n_feat < 0 || (n_feat > 0 && !feat_xy)) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_project_image: bad argument");
  if (!c->has_submap) PCD_FAIL(ctx, PCD_E_INVALID, "sub-map not built (call pcd_submap_build)");
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->feat_buf.ensure(16 * (size_t)(n_feat > 0 ? n_feat : 1) + 8 * (size_t)(n_feat > 0 ? n_feat : 1) + 128));
  double* d_feat = ctx->feat_buf.as<double>();
  int32_t* d_widx = (int32_t*)(d_feat + 2 * (n_feat > 0 ? n_feat : 1));
  float* d_wr = (float*)(d_widx + (n_feat > 0 ? n_feat : 1));
  PCD_CUDA(ctx, ctx->misc.ensure(256));
  unsigned long long* d_stats = (unsigned long long*)ctx->misc.p;
  PCD_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 32, st));
  if (n_feat > 0) PCD_CUDA(ctx, cudaMemcpyAsync(d_feat, feat_xy, 16 * (size_t)n_feat, cudaMemcpyHostToDevice, st));
  unsigned long long* zb = nullptr; pcd::ImgParams P;
  int rc = pcd::project_one(c, img, coef, d_feat, n_feat, d_widx, d_wr, dense_zbuf != nullptr, d_stats, st, &zb, &P);
  if (rc) return rc;
  if (n_feat > 0 && winner_idx) PCD_CUDA(ctx, cudaMemcpyAsync(winner_idx, d_widx, 4 * (size_t)n_feat, cudaMemcpyDeviceToHost, st));
  if (n_feat > 0 && winner_range) PCD_CUDA(ctx, cudaMemcpyAsync(winner_range, d_wr, 4 * (size_t)n_feat, cudaMemcpyDeviceToHost, st));
  if (dense_zbuf) PCD_CUDA(ctx, cudaMemcpyAsync(dense_zbuf, zb, 8 * (size_t)P.w * P.h, cudaMemcpyDeviceToHost, st));
  unsigned long long h_stats[4] = {0, 0, 0, 0};
  PCD_CUDA(ctx, cudaMemcpyAsync(h_stats, d_stats, 32, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  if (stats) for (int i = 0; i < 4; i++) stats[i] = (int64_t)h_stats[i];
  return PCD_OK;
}

int pcd_project_images(pcd_cloud* c, const pcd_image* imgs, int64_t n_img, const double coef[4], const double* feat_xy,
                       const int64_t* feat_start, int32_t* winner_idx, float* winner_range, int64_t* stats) {
  if (!c) return PCD_E_INVALID;
  pcd_ctx* ctx = c->ctx;
  if (n_img < 0 || (n_img > 0 && (!imgs || !feat_start)) || !coef) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_project_images: bad argument");
  if (!c->has_submap) PCD_FAIL(ctx, PCD_E_INVALID, "sub-map not built (call pcd_submap_build)");
  if (n_img == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  const int64_t nf = feat_start[n_img];
  if (nf < 0 || (nf > 0 && !feat_xy)) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_project_images: bad feature list");
  const int64_t nfa = nf > 0 ? nf : 1;
  PCD_CUDA(ctx, ctx->feat_buf.ensure(24 * (size_t)nfa + 128));
  double* d_feat = ctx->feat_buf.as<double>();
  int32_t* d_widx = (int32_t*)(d_feat + 2 * nfa);
  float* d_wr = (float*)(d_widx + nfa);
  PCD_CUDA(ctx, ctx->misc.ensure(256));
  unsigned long long* d_stats = (unsigned long long*)ctx->misc.p;
  PCD_CUDA(ctx, cudaMemsetAsync(d_stats, 0, 32, st));
  if (nf > 0) PCD_CUDA(ctx, cudaMemcpyAsync(d_feat, feat_xy, 16 * (size_t)nf, cudaMemcpyHostToDevice, st));
  for (int64_t k = 0; k < n_img; k++) {
    int64_t lo = feat_start[k], m = feat_start[k + 1] - lo;
    if (m < 0) PCD_FAIL(ctx, PCD_E_INVALID, "feat_start must be non-decreasing");
    int rc = pcd::project_one(c, imgs + k, coef, d_feat + 2 * lo, m, d_widx + lo, d_wr + lo, false, d_stats, st, nullptr, nullptr);
    if (rc) return rc;
  }
  if (nf > 0 && winner_idx) PCD_CUDA(ctx, cudaMemcpyAsync(winner_idx, d_widx, 4 * (size_t)nf, cudaMemcpyDeviceToHost, st));
  if (nf > 0 && winner_range) PCD_CUDA(ctx, cudaMemcpyAsync(winner_range, d_wr, 4 * (size_t)nf, cudaMemcpyDeviceToHost, st));
  unsigned long long h_stats[4] = {0, 0, 0, 0};
  PCD_CUDA(ctx, cudaMemcpyAsync(h_stats, d_stats, 32, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  if (stats) for (int i = 0; i < 4; i++) stats[i] = (int64_t)h_stats[i];
  return PCD_OK;
}

int pcd_ray_plane_batch(pcd_cloud* c, const pcd_image* img, const double* feat_xy, const int32_t* winner_idx, int64_t n,
                        uint8_t* matched, double* xyz) {
  if (!c) return PCD_E_INVALID;
  pcd_ctx* ctx = c->ctx;
  if (!img || n < 0 || (n > 0 && (!feat_xy || !winner_idx || !matched || !xyz))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_ray_plane_batch: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(20 * (size_t)n + 64));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure(25 * (size_t)n + 64));
  double* d_xy = ctx->stage_in[0].as<double>(); int32_t* d_w = (int32_t*)(d_xy + 2 * n);
  double* d_xyz = ctx->stage_out[0].as<double>(); uint8_t* d_m = (uint8_t*)(d_xyz + 3 * n);
  PCD_CUDA(ctx, cudaMemcpyAsync(d_xy, feat_xy, 16 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_w, winner_idx, 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  pcd::k_ray_plane<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)c->aos, d_xy, d_w, n, img->cam[0], img->cam[1], img->cam[2], img->cam[3], d_m, d_xyz);
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaMemcpyAsync(xyz, d_xyz, 24 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(matched, d_m, (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

}  // extern "C"
