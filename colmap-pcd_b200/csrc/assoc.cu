// assoc.cu -- fused post-NN stage: the LidarPoint record the three reference loops build.
//   a5  PointCloudProcess::SearchNearestNeiborByKdtree        src/lidar/ply.cc:90-107
//   a6  BundleAdjustmentConfig::MatchClosestLidarPoint        src/optim/bundle_adjustment.cc:358-410
//       twins: src/sfm/incremental_mapper.cc:1430-1468, src/controllers/bundle_adjustment.cc:141-183
//   a7/a8 LidarPoint::Normalize / ComputeDist / ComputePointToPointDist / ComputeAngle
//                                                             src/lidar/lidar_point.cc:21-50
//   a17 BundleAdjustmentConfig::MatchVariablePoint2LidarPoint src/optim/bundle_adjustment.cc:288-350
// One thread per query in ORIGINAL order: coalesced reads of p / idx, one 32-byte gather of
// the winner (xyz + normal share a sector in the reference's AoS), coalesced SoA writes.
// All arithmetic is fp64 in the reference's operation order; the library is built with
// -fmad=false so no product-sum is contracted.
#include "nn.cuh"
#include <math.h>
#include <vector>

namespace pcd {

struct AssocOut { pcd_assoc_out o; };

__device__ __forceinline__ double norm3(double x, double y, double z) { return sqrt(x * x + (y * y + z * z)); }  // Eigen redux order

// widen + validity of ply.cc:94-101
__device__ __forceinline__ bool widen(const float4* __restrict__ aos, int32_t idx, double l[6]) {
  if (idx < 0) return false;
  float4 a = __ldg(&aos[2 * (int64_t)idx]), b = __ldg(&aos[2 * (int64_t)idx + 1]);
  l[0] = a.x; l[1] = a.y; l[2] = a.z; l[3] = b.x; l[4] = b.y; l[5] = b.z;
  bool nan = false;
#pragma unroll
  for (int k = 0; k < 6; k++) nan |= isnan(l[k]);
  if (nan) return false;
  if (norm3(l[3], l[4], l[5]) < 1e-6) return false;
  return true;
}

__device__ __forceinline__ void normalize_plane(const double l[6], double abcd[4]) {
  double a = l[3], b = l[4], c = l[5];
  double norm = sqrt(a * a + b * b + c * c);      // pow(a,2)+pow(b,2)+pow(c,2), left to right (lidar_point.cc:43)
  a = a / norm; b = b / norm; c = c / norm;
  double d = 0 - a * l[0] - b * l[1] - c * l[2];  // :48
  abcd[0] = a; abcd[1] = b; abcd[2] = c; abcd[3] = d;
}

// Optional fused a18 + a19 (FUSED_EVAL): weight by type (src/optim/bundle_adjustment.cc:1013-1028) and the Jet<double,3>
// evaluation of the cost function (src/base/cost_functions.h:226-229) at the SfM point itself for ACCEPTED matches, in
// the operation order of k_resjac (resjac.cu); a rejected query has no residual block in the reference and reports
// r = 0, J = 0.  The point, plane and type are already in registers here, so the fused pass saves re-reading them.
struct EvalDev { double wp, wi, wg; double* r; double* J; double* partial; };

template <bool FUSED_EVAL>
__global__ void __launch_bounds__(256) k_assoc_record(const float4* __restrict__ aos, const double* __restrict__ p, const int32_t* nn_idx /* may alias o.idx */,
                               const double* __restrict__ range, const int32_t* __restrict__ opt_num, pcd_range_sweep sw,
                               int64_t n, int variant, pcd_assoc_out o, unsigned long long* __restrict__ counters, EvalDev ev) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool acc = false, ground = false;
  double rr = 0.0;
  if (i < n) {
    double px = p[3 * i], py = p[3 * i + 1], pz = p[3 * i + 2];
    int32_t idx = nn_idx[i];
    double l[6];
    uint8_t type = PCD_TYPE_NONE; uint8_t r = 0, g = 0, b = 0;
    double abcd[4] = {0, 0, 0, 0}, dist = 0, angle = 0, lx = 0, ly = 0, lz = 0;
    // The search kernels report the nearest point inside the float ball of the search range; the reference's
    // own test |l - p| <= range (fp64, :398-400 and twins) decides here.  A nearest point beyond the range is
    // reported exactly like "no neighbour" (idx -1, type NONE): the reference would find it and drop it.
    double R = INFINITY;
    if (variant == PCD_ASSOC_WHOLE_MAP) R = 2.0;
    else if (range) R = range[i];
    else if (opt_num) R = sweep_range(opt_num[i], sw.max_r, sw.min_r, sw.drop);
    if (idx >= 0 && R != INFINITY) {
      float4 a = __ldg(&aos[2 * (int64_t)idx]);
      double vx = (double)a.x - px, vy = (double)a.y - py, vz = (double)a.z - pz;
      if (norm3(vx, vy, vz) > R) idx = -1;
    }
    if (widen(aos, idx, l)) {
      normalize_plane(l, abcd);
      lx = l[0]; ly = l[1]; lz = l[2];
      ground = (fabs(l[4] / l[3]) > 10) && (fabs(l[4] / l[5]) > 10);   // raw normal (bundle_adjustment.cc:381)
      type = ground ? PCD_TYPE_ICP_GROUND : PCD_TYPE_ICP;
      if (ground) { r = 255; g = 255; b = 0; }
      else if (variant == PCD_ASSOC_LOCAL) { r = 0; g = 255; b = 0; }
      else { r = 0; g = 0; b = 255; }
      double vx = lx - px, vy = ly - py, vz = lz - pz;
      double d = norm3(vx, vy, vz);                                  // ComputePointToPointDist
      bool ok;
      if (variant == PCD_ASSOC_WHOLE_MAP) {
        double d2plane = fabs((px * abcd[0] + (py * abcd[1] + pz * abcd[2])) + abcd[3]);   // ComputeDist
        ok = !(d2plane > 1 || d > 2);
      } else {
        ok = !(d > R);
      }
      if (ok) {
        acc = true;
        dist = d;
        double wx = px - lx, wy = py - ly, wz = pz - lz;
        angle = fabs((abcd[0] * wx + (abcd[1] * wy + abcd[2] * wz)) / norm3(wx, wy, wz));   // ComputeAngle
      }
    }
    if (o.idx) o.idx[i] = idx;
    if (o.accepted) o.accepted[i] = acc ? 1 : 0;
    if (o.type) o.type[i] = type;
    if (o.rgb) { o.rgb[3 * i] = r; o.rgb[3 * i + 1] = g; o.rgb[3 * i + 2] = b; }
    if (o.xyz) { o.xyz[3 * i] = lx; o.xyz[3 * i + 1] = ly; o.xyz[3 * i + 2] = lz; }
    if (o.abcd) { o.abcd[4 * i] = abcd[0]; o.abcd[4 * i + 1] = abcd[1]; o.abcd[4 * i + 2] = abcd[2]; o.abcd[4 * i + 3] = abcd[3]; }
    if (o.dist) o.dist[i] = dist;
    if (o.angle) o.angle[i] = angle;
    if (FUSED_EVAL) {
      double J[3] = {0.0, 0.0, 0.0};
      // a NaN plane coefficient (e.g. an infinite normal) adds no residual block: bundle_adjustment.cc:1003-1007
      if (acc && !(isnan(abcd[0]) || isnan(abcd[1]) || isnan(abcd[2]) || isnan(abcd[3]))) {
        const double w = type == PCD_TYPE_PROJ ? ev.wp : (type == PCD_TYPE_ICP ? ev.wi : (type == PCD_TYPE_ICP_GROUND ? ev.wg : 0.0));
        double s = ((px * abcd[0] + py * abcd[1]) + pz * abcd[2]) + abcd[3];
        double u = 0.0 - s;
        double f = u * u;
        double sq = sqrt(f);
        double inv = 1.0 / (2.0 * sq);
        rr = w * sq;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          double du = 0.0 - abcd[k];
          double fv = u * du + du * u;
          J[k] = w * (fv * inv) + 0.0 * sq;
        }
      }
      ev.r[i] = rr;
      if (ev.J) { ev.J[3 * i] = J[0]; ev.J[3 * i + 1] = J[1]; ev.J[3 * i + 2] = J[2]; }
    }
  }
  if (FUSED_EVAL && ev.partial) {      // deterministic: fixed tree inside the block, fixed order across blocks
    double s2 = rr * rr;
    for (int o2 = 16; o2 > 0; o2 >>= 1) s2 += __shfl_down_sync(0xffffffffu, s2, o2);
    __shared__ double sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s2;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int k = 0; k < 8; k++) t += sh[k];
      ev.partial[blockIdx.x] = t;
    }
  }
  if (counters) {
    unsigned ma = __ballot_sync(0xffffffffu, acc), mg = __ballot_sync(0xffffffffu, acc && ground);
    if ((threadIdx.x & 31) == 0) {
      if (ma) atomicAdd(&counters[0], (unsigned long long)__popc(ma));
      if (mg) atomicAdd(&counters[1], (unsigned long long)__popc(mg));
    }
  }
}

__global__ void k_search_nn_widen(const float4* __restrict__ aos, const int32_t* __restrict__ nn_idx, int64_t n,
                                  double* __restrict__ l_pt, uint8_t* __restrict__ ok) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double l[6] = {0, 0, 0, 0, 0, 0};
  int32_t idx = nn_idx[i];
  bool good = widen(aos, idx, l);
  if (idx < 0) { for (int k = 0; k < 6; k++) l[k] = 0; }
#pragma unroll
  for (int k = 0; k < 6; k++) l_pt[6 * i + k] = l[k];
  ok[i] = good ? 1 : 0;
}

// a17: one thread per SfM point, candidates in track order, strict < from 360.
__global__ void k_match_proj(const double* __restrict__ p, const double* __restrict__ cand, const int64_t* __restrict__ cs,
                             int64_t n, uint8_t* __restrict__ found, double* __restrict__ xyz, double* __restrict__ abcd,
                             double* __restrict__ dist, double* __restrict__ angle_out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double px = p[3 * i], py = p[3 * i + 1], pz = p[3 * i + 2];
  double angle = 360; int64_t bestk = -1;
  for (int64_t k = cs[i]; k < cs[i + 1]; k++) {
    const double* c = cand + 6 * k;
    double vx = px - c[0], vy = py - c[1], vz = pz - c[2];
    double t = fabs((vx * c[3] + (vy * c[4] + vz * c[5])) / (norm3(c[3], c[4], c[5]) * norm3(vx, vy, vz)));
    if (t < angle) { angle = t; bestk = k; }
  }
  double l[6] = {0, 0, 0, 0, 0, 0}, pl[4] = {0, 0, 0, 0}, d = 0, ang = 0;
  bool ok = (angle != 360) && bestk >= 0;
  if (ok) {
    for (int k = 0; k < 6; k++) l[k] = cand[6 * bestk + k];
    normalize_plane(l, pl);
    d = fabs((px * pl[0] + (py * pl[1] + pz * pl[2])) + pl[3]);
    double wx = px - l[0], wy = py - l[1], wz = pz - l[2];
    ang = fabs((pl[0] * wx + (pl[1] * wy + pl[2] * wz)) / norm3(wx, wy, wz));
  }
  if (found) found[i] = ok ? 1 : 0;
  if (xyz) { xyz[3 * i] = l[0]; xyz[3 * i + 1] = l[1]; xyz[3 * i + 2] = l[2]; }
  if (abcd) { abcd[4 * i] = pl[0]; abcd[4 * i + 1] = pl[1]; abcd[4 * i + 2] = pl[2]; abcd[4 * i + 3] = pl[3]; }
  if (dist) dist[i] = d;
  if (angle_out) angle_out[i] = ang;
}

// The arithmetic of the match record as the reference's class methods expose it (src/lidar/lidar_point.cc:21-50):
// Normalize, ComputeDist, ComputePointToPointDist, ComputeAngle, same device functions as k_assoc_record.  The shim's
// colmap::LidarPoint calls this with n = 1 where a call site is left serial; batched callers never need it.
__global__ void k_lidar_point_ops(const double* __restrict__ xyz, const double* __restrict__ abcd_in, const double* __restrict__ pt, int64_t n,
                                  int normalize, double* __restrict__ abcd_out, double* __restrict__ d_plane, double* __restrict__ d_point,
                                  double* __restrict__ angle) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double l[6] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], abcd_in[4 * i], abcd_in[4 * i + 1], abcd_in[4 * i + 2]};
  double pl[4] = {abcd_in[4 * i], abcd_in[4 * i + 1], abcd_in[4 * i + 2], abcd_in[4 * i + 3]};
  if (normalize) normalize_plane(l, pl);
  if (abcd_out) { abcd_out[4 * i] = pl[0]; abcd_out[4 * i + 1] = pl[1]; abcd_out[4 * i + 2] = pl[2]; abcd_out[4 * i + 3] = pl[3]; }
  if (!pt) return;
  const double px = pt[3 * i], py = pt[3 * i + 1], pz = pt[3 * i + 2];
  if (d_plane) d_plane[i] = fabs((px * pl[0] + (py * pl[1] + pz * pl[2])) + pl[3]);                 // ComputeDist
  if (d_point) d_point[i] = norm3(l[0] - px, l[1] - py, l[2] - pz);                                 // ComputePointToPointDist
  if (angle) {
    const double wx = px - l[0], wy = py - l[1], wz = pz - l[2];
    angle[i] = fabs((pl[0] * wx + (pl[1] * wy + pl[2] * wz)) / norm3(wx, wy, wz));                  // ComputeAngle
  }
}

__global__ void k_sum_blocks(const double* __restrict__ partial, int nb, double* __restrict__ out) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < nb; i += 256) s += partial[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sh[0];
}

struct EvalArgs { double wp, wi, wg; double* r; double* J; double* sum_sq; };

}  // namespace pcd

extern "C" {

static int associate_dev_impl(pcd_cloud* cloud, const double* d_p, const double* d_range, const int32_t* d_opt_num,
                              const pcd_range_sweep* sweep, int64_t n, int variant, const pcd_assoc_out* d_out,
                              const pcd::EvalDev* ev, double* d_sum_sq, cudaStream_t st) {
  pcd_ctx* ctx = cloud->ctx;
  int32_t* d_idx = d_out->idx;
  if (!d_idx) { PCD_CUDA(ctx, ctx->partial.ensure(4 * (size_t)n)); d_idx = ctx->partial.as<int32_t>(); }
  // search radius: the accept test bounds how far a useful neighbour can be
  double fixed = variant == PCD_ASSOC_WHOLE_MAP ? 2.0 : INFINITY;
  const double* rr = variant == PCD_ASSOC_WHOLE_MAP ? nullptr : d_range;
  const int32_t* on = variant == PCD_ASSOC_WHOLE_MAP ? nullptr : (d_range ? nullptr : d_opt_num);
  int rc = pcd::nn_search_dev(cloud, d_p, rr, fixed, on, sweep, n, d_idx, nullptr, false, st);   // k_assoc_record applies the fp64 range test
  if (rc) return rc;
  pcd_range_sweep sw = sweep ? *sweep : pcd_range_sweep{0, 0, 0};
  pcd_assoc_out o = *d_out;   // o.idx (= the search's output array when given) is rewritten: -1 for a neighbour beyond the range
  const int nb = pcd_blocks(n, 256);
  pcd::ProfScope ps(ctx, PCD_K_ASSOC_RECORD, st);
  if (ev) {
    pcd::EvalDev e = *ev;
    if (d_sum_sq) { PCD_CUDA(ctx, ctx->eval_partial.ensure(8 * (size_t)nb)); e.partial = ctx->eval_partial.as<double>(); }
    pcd::k_assoc_record<true><<<nb, 256, 0, st>>>((const float4*)cloud->aos, d_p, d_idx, rr, on, sw, n, variant, o, nullptr, e);
    PCD_LAUNCH_CHECK(ctx);
    if (d_sum_sq) { pcd::k_sum_blocks<<<1, 256, 0, st>>>(e.partial, nb, d_sum_sq); PCD_LAUNCH_CHECK(ctx); }
  } else {
    pcd::k_assoc_record<false><<<nb, 256, 0, st>>>((const float4*)cloud->aos, d_p, d_idx, rr, on, sw, n, variant, o, nullptr, pcd::EvalDev{});
    PCD_LAUNCH_CHECK(ctx);
  }
  return PCD_OK;
}

int pcd_associate_batch_dev(pcd_cloud* cloud, const double* d_p, const double* d_range, const int32_t* d_opt_num,
                            const pcd_range_sweep* sweep, int64_t n, int variant, const pcd_assoc_out* d_out, void* stream) {
  if (!cloud) return PCD_E_INVALID;
  pcd_ctx* ctx = cloud->ctx;
  if (n < 0 || !d_out || (n > 0 && !d_p) || variant < 0 || variant > 2) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_associate_batch_dev: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  return associate_dev_impl(cloud, d_p, d_range, d_opt_num, sweep, n, variant, d_out, nullptr, nullptr, (cudaStream_t)stream);
}

int pcd_associate_evaluate_batch_dev(pcd_cloud* cloud, const double* d_p, const double* d_range, const int32_t* d_opt_num,
                                     const pcd_range_sweep* sweep, int64_t n, int variant, double w_proj, double w_icp,
                                     double w_ground, const pcd_assoc_out* d_out, double* d_r, double* d_J, double* d_sum_sq,
                                     void* stream) {
  if (!cloud) return PCD_E_INVALID;
  pcd_ctx* ctx = cloud->ctx;
  if (n < 0 || !d_out || (n > 0 && (!d_p || !d_r)) || variant < 0 || variant > 2)
    PCD_FAIL(ctx, PCD_E_INVALID, "pcd_associate_evaluate_batch_dev: bad argument");
  DeviceGuard g(ctx->device);
  if (n == 0) { if (d_sum_sq) PCD_CUDA(ctx, cudaMemsetAsync(d_sum_sq, 0, 8, (cudaStream_t)stream)); return PCD_OK; }
  pcd::EvalDev ev = {w_proj, w_icp, w_ground, d_r, d_J, nullptr};
  return associate_dev_impl(cloud, d_p, d_range, d_opt_num, sweep, n, variant, d_out, &ev, d_sum_sq, (cudaStream_t)stream);
}

static int associate_pipeline(pcd_cloud* cloud, const double* p, const double* range, const int32_t* opt_num,
                              const pcd_range_sweep* sweep, int64_t n, int variant, const pcd_assoc_out* out,
                              const pcd::EvalArgs* ev) {
  pcd_ctx* ctx = cloud->ctx;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  // Chunked pipeline: H2D of chunk k+1 and D2H of chunk k-1 overlap the kernels of chunk k.
  const int64_t CH = (int64_t)1 << 20;
  const size_t per_in = 24 + 8 + 4, per_out = 4 + 1 + 1 + 3 + 24 + 32 + 8 + 8 + 16 + (ev ? 8 + 24 + 1 : 0);
  for (int b = 0; b < 2; b++) {
    PCD_CUDA(ctx, ctx->stage_in[b].ensure(per_in * (size_t)(n < CH ? n : CH) + 64));
    PCD_CUDA(ctx, ctx->stage_out[b].ensure(per_out * (size_t)(n < CH ? n : CH) + 256));
  }
  int64_t nchunks = (n + CH - 1) / CH;
  double* d_chunk_sums = nullptr;
  if (ev && ev->sum_sq) {
    PCD_CUDA(ctx, ctx->stage_aux[0].ensure(8 * (size_t)nchunks));
    d_chunk_sums = ctx->stage_aux[0].as<double>();
  }
  for (int64_t c = 0; c < nchunks; c++) {
    int b = (int)(c & 1);
    int64_t lo = c * CH, m = (n - lo < CH) ? (n - lo) : CH;
    if (c >= 2) PCD_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_in, ctx->ev[2 + b], 0));   // buffers of chunk c-2 drained
    char* din = (char*)ctx->stage_in[b].p;
    double* d_p = (double*)din;
    double* d_range = range ? (double*)(din + 24 * (size_t)m) : nullptr;
    int32_t* d_on = (!range && opt_num) ? (int32_t*)(din + 32 * (size_t)m) : nullptr;
    PCD_CUDA(ctx, cudaMemcpyAsync(d_p, p + 3 * lo, 24 * (size_t)m, cudaMemcpyHostToDevice, ctx->copy_in));
    if (d_range) PCD_CUDA(ctx, cudaMemcpyAsync(d_range, range + lo, 8 * (size_t)m, cudaMemcpyHostToDevice, ctx->copy_in));
    if (d_on) PCD_CUDA(ctx, cudaMemcpyAsync(d_on, opt_num + lo, 4 * (size_t)m, cudaMemcpyHostToDevice, ctx->copy_in));
    PCD_CUDA(ctx, cudaEventRecord(ctx->ev[b], ctx->copy_in));
    PCD_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev[b], 0));
    // device output layout (8-byte fields first)
    char* dout = (char*)ctx->stage_out[b].p;
    pcd_assoc_out d = {};
    size_t off = 0;
    auto take = [&](size_t bytes) { char* q = dout + off; off += (bytes + 15) & ~(size_t)15; return q; };
    d.abcd = (out->abcd || ev) ? (double*)take(32 * (size_t)m) : nullptr;
    double* d_r = ev ? (double*)take(8 * (size_t)m) : nullptr;
    double* d_J = (ev && ev->J) ? (double*)take(24 * (size_t)m) : nullptr;
    d.xyz = out->xyz ? (double*)take(24 * (size_t)m) : nullptr;
    d.dist = out->dist ? (double*)take(8 * (size_t)m) : nullptr;
    d.angle = out->angle ? (double*)take(8 * (size_t)m) : nullptr;
    d.idx = (int32_t*)take(4 * (size_t)m);
    d.rgb = out->rgb ? (uint8_t*)take(3 * (size_t)m) : nullptr;
    d.accepted = (out->accepted || ev) ? (uint8_t*)take((size_t)m) : nullptr;
    d.type = (out->type || ev) ? (uint8_t*)take((size_t)m) : nullptr;
    if (c >= 2) PCD_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev[4 + b], 0));             // D2H of chunk c-2 finished
    int rc;
    if (ev) {
      pcd::EvalDev e = {ev->wp, ev->wi, ev->wg, d_r, d_J, nullptr};
      rc = associate_dev_impl(cloud, d_p, d_range, d_on, sweep, m, variant, &d, &e, d_chunk_sums ? d_chunk_sums + c : nullptr, st);
    } else {
      rc = associate_dev_impl(cloud, d_p, d_range, d_on, sweep, m, variant, &d, nullptr, nullptr, st);
    }
    if (rc) return rc;
    PCD_CUDA(ctx, cudaEventRecord(ctx->ev[2 + b], st));
    PCD_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_out, ctx->ev[2 + b], 0));
    cudaStream_t co = ctx->copy_out;
    if (out->abcd) PCD_CUDA(ctx, cudaMemcpyAsync(out->abcd + 4 * lo, d.abcd, 32 * (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->xyz) PCD_CUDA(ctx, cudaMemcpyAsync(out->xyz + 3 * lo, d.xyz, 24 * (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->dist) PCD_CUDA(ctx, cudaMemcpyAsync(out->dist + lo, d.dist, 8 * (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->angle) PCD_CUDA(ctx, cudaMemcpyAsync(out->angle + lo, d.angle, 8 * (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->idx) PCD_CUDA(ctx, cudaMemcpyAsync(out->idx + lo, d.idx, 4 * (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->rgb) PCD_CUDA(ctx, cudaMemcpyAsync(out->rgb + 3 * lo, d.rgb, 3 * (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->accepted) PCD_CUDA(ctx, cudaMemcpyAsync(out->accepted + lo, d.accepted, (size_t)m, cudaMemcpyDeviceToHost, co));
    if (out->type) PCD_CUDA(ctx, cudaMemcpyAsync(out->type + lo, d.type, (size_t)m, cudaMemcpyDeviceToHost, co));
    if (ev) {
      PCD_CUDA(ctx, cudaMemcpyAsync(ev->r + lo, d_r, 8 * (size_t)m, cudaMemcpyDeviceToHost, co));
      if (ev->J) PCD_CUDA(ctx, cudaMemcpyAsync(ev->J + 3 * lo, d_J, 24 * (size_t)m, cudaMemcpyDeviceToHost, co));
    }
    PCD_CUDA(ctx, cudaEventRecord(ctx->ev[4 + b], co));
  }
  PCD_CUDA(ctx, cudaStreamSynchronize(ctx->copy_out));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  if (d_chunk_sums) {     // chunk sums added on the host in chunk order: deterministic
    std::vector<double> hs((size_t)nchunks);
    PCD_CUDA(ctx, cudaMemcpy(hs.data(), d_chunk_sums, 8 * (size_t)nchunks, cudaMemcpyDeviceToHost));
    double t = 0.0;
    for (double v : hs) t += v;
    *ev->sum_sq = t;
  }
  return PCD_OK;
}

int pcd_associate_batch(pcd_cloud* cloud, const double* p, const double* range, const int32_t* opt_num,
                        const pcd_range_sweep* sweep, int64_t n, int variant, const pcd_assoc_out* out) {
  if (!cloud) return PCD_E_INVALID;
  pcd_ctx* ctx = cloud->ctx;
  if (n < 0 || !out || (n > 0 && !p) || variant < 0 || variant > 2) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_associate_batch: bad argument");
  if (n == 0) return PCD_OK;
  return associate_pipeline(cloud, p, range, opt_num, sweep, n, variant, out, nullptr);
}

int pcd_associate_evaluate_batch(pcd_cloud* cloud, const double* p, const double* range, const int32_t* opt_num,
                                 const pcd_range_sweep* sweep, int64_t n, int variant, double w_proj, double w_icp,
                                 double w_ground, const pcd_assoc_out* out, double* r, double* J, double* sum_sq) {
  if (!cloud) return PCD_E_INVALID;
  pcd_ctx* ctx = cloud->ctx;
  if (n < 0 || !out || (n > 0 && (!p || !r)) || variant < 0 || variant > 2)
    PCD_FAIL(ctx, PCD_E_INVALID, "pcd_associate_evaluate_batch: bad argument");
  if (n == 0) { if (sum_sq) *sum_sq = 0.0; return PCD_OK; }
  pcd::EvalArgs ev = {w_proj, w_icp, w_ground, r, J, sum_sq};
  return associate_pipeline(cloud, p, range, opt_num, sweep, n, variant, out, &ev);
}

int pcd_search_nn_batch(pcd_cloud* cloud, const double* q, int64_t n, double* l_pt, uint8_t* ok) {
  if (!cloud) return PCD_E_INVALID;
  pcd_ctx* ctx = cloud->ctx;
  if (n < 0 || (n > 0 && (!q || !l_pt || !ok))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_search_nn_batch: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(24 * (size_t)n));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure((48 + 4 + 1) * (size_t)n + 64));
  double* d_q = ctx->stage_in[0].as<double>();
  double* d_l = ctx->stage_out[0].as<double>();
  int32_t* d_idx = (int32_t*)(d_l + 6 * n);
  uint8_t* d_ok = (uint8_t*)(d_idx + n);
  PCD_CUDA(ctx, cudaMemcpyAsync(d_q, q, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  int rc = pcd::nn_search_dev(cloud, d_q, nullptr, INFINITY, nullptr, nullptr, n, d_idx, nullptr, false, st);
  if (rc) return rc;
  pcd::k_search_nn_widen<<<pcd_blocks(n, 256), 256, 0, st>>>((const float4*)cloud->aos, d_idx, n, d_l, d_ok);
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaMemcpyAsync(l_pt, d_l, 48 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(ok, d_ok, (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

int pcd_lidar_point_ops(pcd_ctx* ctx, const double* xyz, const double* abcd, const double* pt, int64_t n, int normalize,
                        double* abcd_out, double* dist_plane, double* dist_point, double* angle) {
  if (!ctx || n < 0 || (n > 0 && (!xyz || !abcd))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_lidar_point_ops: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(80 * (size_t)n + 64));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure(56 * (size_t)n + 64));
  double* d_xyz = ctx->stage_in[0].as<double>();
  double* d_abcd = d_xyz + 3 * n; double* d_pt = pt ? d_abcd + 4 * n : nullptr;
  double* d_out = ctx->stage_out[0].as<double>();
  double* d_dp = d_out + 4 * n; double* d_dd = d_dp + n; double* d_an = d_dd + n;
  PCD_CUDA(ctx, cudaMemcpyAsync(d_xyz, xyz, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_abcd, abcd, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
  if (pt) PCD_CUDA(ctx, cudaMemcpyAsync(d_pt, pt, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  pcd::k_lidar_point_ops<<<pcd_blocks(n, 128), 128, 0, st>>>(d_xyz, d_abcd, d_pt, n, normalize, d_out, d_dp, d_dd, d_an);
  PCD_LAUNCH_CHECK(ctx);
  if (abcd_out) PCD_CUDA(ctx, cudaMemcpyAsync(abcd_out, d_out, 32 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (pt && dist_plane) PCD_CUDA(ctx, cudaMemcpyAsync(dist_plane, d_dp, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (pt && dist_point) PCD_CUDA(ctx, cudaMemcpyAsync(dist_point, d_dd, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (pt && angle) PCD_CUDA(ctx, cudaMemcpyAsync(angle, d_an, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

int pcd_match_proj_batch(pcd_ctx* ctx, const double* p, const double* cand, const int64_t* cand_start, int64_t n,
                         uint8_t* found, double* xyz, double* abcd, double* dist, double* angle) {
  if (!ctx || n < 0 || (n > 0 && (!p || !cand_start))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_match_proj_batch: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  int64_t nc = cand_start[n];
  if (nc < 0 || (nc > 0 && !cand)) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_match_proj_batch: bad candidate list");
  size_t in_bytes = 24 * (size_t)n + 8 * (size_t)(n + 1) + 48 * (size_t)nc;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(in_bytes + 64));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure((24 + 32 + 8 + 8 + 1) * (size_t)n + 64));
  double* d_p = ctx->stage_in[0].as<double>();
  int64_t* d_cs = (int64_t*)(d_p + 3 * n);
  double* d_c = (double*)(d_cs + n + 1);
  double* d_xyz = ctx->stage_out[0].as<double>();
  double* d_abcd = d_xyz + 3 * n; double* d_dist = d_abcd + 4 * n; double* d_ang = d_dist + n;
  uint8_t* d_found = (uint8_t*)(d_ang + n);
  PCD_CUDA(ctx, cudaMemcpyAsync(d_p, p, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_cs, cand_start, 8 * (size_t)(n + 1), cudaMemcpyHostToDevice, st));
  if (nc) PCD_CUDA(ctx, cudaMemcpyAsync(d_c, cand, 48 * (size_t)nc, cudaMemcpyHostToDevice, st));
  pcd::k_match_proj<<<pcd_blocks(n, 128), 128, 0, st>>>(d_p, d_c, d_cs, n, d_found, d_xyz, d_abcd, d_dist, d_ang);
  PCD_LAUNCH_CHECK(ctx);
  if (found) PCD_CUDA(ctx, cudaMemcpyAsync(found, d_found, (size_t)n, cudaMemcpyDeviceToHost, st));
  if (xyz) PCD_CUDA(ctx, cudaMemcpyAsync(xyz, d_xyz, 24 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (abcd) PCD_CUDA(ctx, cudaMemcpyAsync(abcd, d_abcd, 32 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (dist) PCD_CUDA(ctx, cudaMemcpyAsync(dist, d_dist, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (angle) PCD_CUDA(ctx, cudaMemcpyAsync(angle, d_ang, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

}  // extern "C"
