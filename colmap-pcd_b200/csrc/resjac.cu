// resjac.cu -- batched BundleAdjustmentLidarCostFunction (a19, src/base/cost_functions.h:204-232)
// evaluated with ceres::Jet<double,3> rules, weight selection (a18, src/optim/bundle_adjustment.cc:1013-1028)
// and FilterLidarOutlier's distance test (N2, src/base/reconstruction.cc:771-804).
// Streaming kernels, HBM-bound: 96 B per evaluation (p 24, abcd 32, w 8 in; r 8, J 24 out).
#include "common.cuh"
#include <math.h>

namespace pcd {

// residual + Jet derivative for one constraint. J = NaN at u == 0 (0 * inf), as autodiff gives.
// A plane with a NaN coefficient never becomes a residual block (AddLidarToProblem returns before Create,
// src/optim/bundle_adjustment.cc:1003-1007): its row is reported like a rejected match, r = 0 and J = 0.
__device__ __forceinline__ void eval_lidar(double p0, double p1, double p2, double a, double b, double c, double d,
                                           double w, double& r, double J[3]) {
  if (isnan(a) || isnan(b) || isnan(c) || isnan(d)) { r = 0.0; J[0] = J[1] = J[2] = 0.0; return; }
  double s = ((p0 * a + p1 * b) + p2 * c) + d;
  double u = 0.0 - s;
  double f = u * u;
  double sq = sqrt(f);
  double inv = 1.0 / (2.0 * sq);
  r = w * sq;
  double n[3] = {a, b, c};
#pragma unroll
  for (int k = 0; k < 3; k++) {
    double du = 0.0 - n[k];
    double fv = u * du + du * u;
    J[k] = w * (fv * inv) + 0.0 * sq;
  }
}

constexpr int RJ_THREADS = 256;

__global__ void __launch_bounds__(RJ_THREADS) k_resjac(const double* __restrict__ p, const double4* __restrict__ abcd,
                                                       const double* __restrict__ w, int64_t n, double* __restrict__ r,
                                                       double* __restrict__ J, double* __restrict__ partial) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double rr = 0.0;
  if (i < n) {
    double4 pl = abcd[i];
    double Jl[3];
    eval_lidar(p[3 * i], p[3 * i + 1], p[3 * i + 2], pl.x, pl.y, pl.z, pl.w, w[i], rr, Jl);
    r[i] = rr;
    if (J) { J[3 * i] = Jl[0]; J[3 * i + 1] = Jl[1]; J[3 * i + 2] = Jl[2]; }
  }
  if (partial) {   // deterministic: fixed tree inside the block, fixed order across blocks
    double s = rr * rr;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    __shared__ double sh[RJ_THREADS / 32];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int k = 0; k < RJ_THREADS / 32; k++) t += sh[k];
      partial[blockIdx.x] = t;
    }
  }
}

__global__ void k_sum_partials(const double* __restrict__ partial, int nb, double* __restrict__ out) {
  // one block; strided accumulation + fixed tree => run-to-run identical
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

__global__ void k_pose_jac(const double* __restrict__ pc, const double* __restrict__ qv, const double* __restrict__ tv,
                           const double4* __restrict__ abcd, const double* __restrict__ w, int64_t n,
                           double* __restrict__ r, double* __restrict__ Jp, double* __restrict__ Jpose) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double qw = qv[4 * i], x = qv[4 * i + 1], y = qv[4 * i + 2], z = qv[4 * i + 3];
  double v0 = pc[3 * i] - tv[3 * i], v1 = pc[3 * i + 1] - tv[3 * i + 1], v2 = pc[3 * i + 2] - tv[3 * i + 2];
  // Eigen::Quaterniond::toRotationMatrix (not normalised), as the reference builds R (pcd_projection.cc:17-18)
  double R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y - qw * z), 2 * (x * z + qw * y)},
                    {2 * (x * y + qw * z), 1 - 2 * (x * x + z * z), 2 * (y * z - qw * x)},
                    {2 * (x * z - qw * y), 2 * (y * z + qw * x), 1 - 2 * (x * x + y * y)}};
  double pw[3];
  for (int k = 0; k < 3; k++) pw[k] = R[0][k] * v0 + R[1][k] * v1 + R[2][k] * v2;   // R^T v
  double4 pl = abcd[i];
  double rr, J[3];
  eval_lidar(pw[0], pw[1], pw[2], pl.x, pl.y, pl.z, pl.w, w[i], rr, J);
  r[i] = rr;
  if (Jp) { Jp[3 * i] = J[0]; Jp[3 * i + 1] = J[1]; Jp[3 * i + 2] = J[2]; }
  if (Jpose) {
    double dw[3] = {2 * z * v1 - 2 * y * v2, -2 * z * v0 + 2 * x * v2, 2 * y * v0 - 2 * x * v1};
    double dx[3] = {2 * y * v1 + 2 * z * v2, 2 * y * v0 - 4 * x * v1 + 2 * qw * v2, 2 * z * v0 - 2 * qw * v1 - 4 * x * v2};
    double dy[3] = {-4 * y * v0 + 2 * x * v1 - 2 * qw * v2, 2 * x * v0 + 2 * z * v2, 2 * qw * v0 + 2 * z * v1 - 4 * y * v2};
    double dz[3] = {-4 * z * v0 + 2 * qw * v1 + 2 * x * v2, -2 * qw * v0 - 4 * z * v1 + 2 * y * v2, 2 * x * v0 + 2 * y * v1};
    double* o = Jpose + 7 * i;
    o[0] = J[0] * dw[0] + J[1] * dw[1] + J[2] * dw[2];
    o[1] = J[0] * dx[0] + J[1] * dx[1] + J[2] * dx[2];
    o[2] = J[0] * dy[0] + J[1] * dy[1] + J[2] * dy[2];
    o[3] = J[0] * dz[0] + J[1] * dz[1] + J[2] * dz[2];
    for (int j = 0; j < 3; j++) o[4 + j] = -(J[0] * R[j][0] + J[1] * R[j][1] + J[2] * R[j][2]);
  }
}

__global__ void k_weights(const uint8_t* __restrict__ type, int64_t n, double wp, double wi, double wg, double* __restrict__ w) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint8_t t = type[i];
  w[i] = t == PCD_TYPE_PROJ ? wp : (t == PCD_TYPE_ICP ? wi : (t == PCD_TYPE_ICP_GROUND ? wg : 0.0));
}

// fp32 wire format of one residual block for the multi-GPU exchange: {r, J0, J1, J2} rounded to nearest float
// (relative error <= 2^-24, inside the path's 1e-5 tolerance; the fp64 blocks stay on the producing GPU).
__global__ void k_pack_blocks_f32(const double* __restrict__ r, const double* __restrict__ J, int64_t n, float4* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = make_float4((float)r[i], (float)J[3 * i], (float)J[3 * i + 1], (float)J[3 * i + 2]);
}

__global__ void k_filter(const double* __restrict__ p, const double* __restrict__ l, const uint8_t* __restrict__ type, int64_t n,
                         double max_proj, double max_icp, uint8_t* __restrict__ erase, unsigned long long* __restrict__ cnt) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool e = false;
  if (i < n) {
    double x = l[3 * i] - p[3 * i], y = l[3 * i + 1] - p[3 * i + 1], z = l[3 * i + 2] - p[3 * i + 2];
    double d = sqrt(x * x + (y * y + z * z));
    e = type[i] == PCD_TYPE_PROJ ? (d > max_proj) : (d > max_icp);
    erase[i] = e ? 1 : 0;
  }
  unsigned m = __ballot_sync(0xffffffffu, e);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(cnt, (unsigned long long)__popc(m));
}

}  // namespace pcd

extern "C" {

int pcd_residual_jacobian_batch_dev(pcd_ctx* ctx, const double* d_p, const double* d_abcd, const double* d_w, int64_t n,
                                    double* d_r, double* d_J, double* d_sum_sq, void* stream) {
  if (!ctx || n < 0 || (n > 0 && (!d_p || !d_abcd || !d_w || !d_r))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_residual_jacobian_batch_dev: bad argument");
  DeviceGuard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { if (d_sum_sq) PCD_CUDA(ctx, cudaMemsetAsync(d_sum_sq, 0, 8, st)); return PCD_OK; }
  int nb = pcd_blocks(n, pcd::RJ_THREADS);
  double* partial = nullptr;
  if (d_sum_sq) { PCD_CUDA(ctx, ctx->partial.ensure(8 * (size_t)nb)); partial = ctx->partial.as<double>(); }
  pcd::ProfScope ps(ctx, PCD_K_RESJAC, st);
  pcd::k_resjac<<<nb, pcd::RJ_THREADS, 0, st>>>(d_p, (const double4*)d_abcd, d_w, n, d_r, d_J, partial);
  PCD_LAUNCH_CHECK(ctx);
  if (d_sum_sq) { pcd::k_sum_partials<<<1, 256, 0, st>>>(partial, nb, d_sum_sq); PCD_LAUNCH_CHECK(ctx); }
  return PCD_OK;
}

int pcd_pack_blocks_f32_dev(pcd_ctx* ctx, const double* d_r, const double* d_J, int64_t n, float* d_out, void* stream) {
  if (!ctx || n < 0 || (n > 0 && (!d_r || !d_J || !d_out))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_pack_blocks_f32_dev: bad argument");
  if (((uintptr_t)d_out & 15) != 0) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_pack_blocks_f32_dev: d_out must be 16-byte aligned");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  pcd::k_pack_blocks_f32<<<pcd_blocks(n, 256), 256, 0, st>>>(d_r, d_J, n, (float4*)d_out);
  PCD_LAUNCH_CHECK(ctx);
  return PCD_OK;
}

int pcd_residual_jacobian_batch(pcd_ctx* ctx, const double* p, const double* abcd, const double* w, int64_t n, double* r,
                                double* J, double* sum_sq) {
  if (!ctx || n < 0 || (n > 0 && (!p || !abcd || !w || !r))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_residual_jacobian_batch: bad argument");
  if (n == 0) { if (sum_sq) *sum_sq = 0.0; return PCD_OK; }
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(64 * (size_t)n));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure(32 * (size_t)n + 64));
  double* d_abcd = ctx->stage_in[0].as<double>();       // 32-byte aligned rows first
  double* d_p = d_abcd + 4 * n; double* d_w = d_p + 3 * n;
  double* d_r = ctx->stage_out[0].as<double>(); double* d_J = d_r + n; double* d_s = d_J + 3 * n;
  PCD_CUDA(ctx, cudaMemcpyAsync(d_abcd, abcd, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_p, p, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_w, w, 8 * (size_t)n, cudaMemcpyHostToDevice, st));
  int rc = pcd_residual_jacobian_batch_dev(ctx, d_p, d_abcd, d_w, n, d_r, J ? d_J : nullptr, sum_sq ? d_s : nullptr, st);
  if (rc) return rc;
  PCD_CUDA(ctx, cudaMemcpyAsync(r, d_r, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (J) PCD_CUDA(ctx, cudaMemcpyAsync(J, d_J, 24 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (sum_sq) PCD_CUDA(ctx, cudaMemcpyAsync(sum_sq, d_s, 8, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

int pcd_residual_pose_jacobian_batch(pcd_ctx* ctx, const double* p_cam, const double* qvec, const double* tvec,
                                     const double* abcd, const double* w, int64_t n, double* r, double* J_point, double* J_pose) {
  if (!ctx || n < 0 || (n > 0 && (!p_cam || !qvec || !tvec || !abcd || !w || !r))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_residual_pose_jacobian_batch: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(8 * 15 * (size_t)n));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure(8 * 11 * (size_t)n));
  double* d_abcd = ctx->stage_in[0].as<double>();
  double* d_q = d_abcd + 4 * n; double* d_p = d_q + 4 * n; double* d_t = d_p + 3 * n; double* d_w = d_t + 3 * n;
  double* d_r = ctx->stage_out[0].as<double>(); double* d_Jp = d_r + n; double* d_Jq = d_Jp + 3 * n;
  PCD_CUDA(ctx, cudaMemcpyAsync(d_abcd, abcd, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_q, qvec, 32 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_p, p_cam, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_t, tvec, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_w, w, 8 * (size_t)n, cudaMemcpyHostToDevice, st));
  pcd::k_pose_jac<<<pcd_blocks(n, 128), 128, 0, st>>>(d_p, d_q, d_t, (const double4*)d_abcd, d_w, n, d_r, J_point ? d_Jp : nullptr, J_pose ? d_Jq : nullptr);
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaMemcpyAsync(r, d_r, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (J_point) PCD_CUDA(ctx, cudaMemcpyAsync(J_point, d_Jp, 24 * (size_t)n, cudaMemcpyDeviceToHost, st));
  if (J_pose) PCD_CUDA(ctx, cudaMemcpyAsync(J_pose, d_Jq, 56 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

int pcd_weights_from_type_dev(pcd_ctx* ctx, const uint8_t* d_type, int64_t n, double w_proj, double w_icp, double w_ground,
                              double* d_w_out, void* stream) {
  if (!ctx || n < 0 || (n > 0 && (!d_type || !d_w_out))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_weights_from_type_dev: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  pcd::k_weights<<<pcd_blocks(n, 256), 256, 0, (cudaStream_t)stream>>>(d_type, n, w_proj, w_icp, w_ground, d_w_out);
  PCD_LAUNCH_CHECK(ctx);
  return PCD_OK;
}

int pcd_weights_from_type(pcd_ctx* ctx, const uint8_t* type, int64_t n, double w_proj, double w_icp, double w_ground, double* w_out) {
  if (!ctx || n < 0 || (n > 0 && (!type || !w_out))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_weights_from_type: bad argument");
  if (n == 0) return PCD_OK;
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure((size_t)n));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure(8 * (size_t)n));
  PCD_CUDA(ctx, cudaMemcpyAsync(ctx->stage_in[0].p, type, (size_t)n, cudaMemcpyHostToDevice, st));
  pcd::k_weights<<<pcd_blocks(n, 256), 256, 0, st>>>(ctx->stage_in[0].as<uint8_t>(), n, w_proj, w_icp, w_ground, ctx->stage_out[0].as<double>());
  PCD_LAUNCH_CHECK(ctx);
  PCD_CUDA(ctx, cudaMemcpyAsync(w_out, ctx->stage_out[0].p, 8 * (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  return PCD_OK;
}

int pcd_filter_outliers_batch(pcd_ctx* ctx, const double* p, const double* lidar_xyz, const uint8_t* type, int64_t n,
                              double max_proj, double max_icp, uint8_t* erase, int64_t* n_erased) {
  if (!ctx || n < 0 || (n > 0 && (!p || !lidar_xyz || !type || !erase))) PCD_FAIL(ctx, PCD_E_INVALID, "pcd_filter_outliers_batch: bad argument");
  if (n == 0) { if (n_erased) *n_erased = 0; return PCD_OK; }
  DeviceGuard g(ctx->device);
  cudaStream_t st = ctx->stream;
  PCD_CUDA(ctx, ctx->stage_in[0].ensure(49 * (size_t)n + 64));
  PCD_CUDA(ctx, ctx->stage_out[0].ensure((size_t)n + 64));
  double* d_p = ctx->stage_in[0].as<double>(); double* d_l = d_p + 3 * n; uint8_t* d_t = (uint8_t*)(d_l + 3 * n);
  unsigned long long* d_cnt = (unsigned long long*)ctx->stage_out[0].p; uint8_t* d_e = (uint8_t*)(d_cnt + 2);
  PCD_CUDA(ctx, cudaMemcpyAsync(d_p, p, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_l, lidar_xyz, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(d_t, type, (size_t)n, cudaMemcpyHostToDevice, st));
  PCD_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, 8, st));
  pcd::k_filter<<<pcd_blocks(n, 256), 256, 0, st>>>(d_p, d_l, d_t, n, max_proj, max_icp, d_e, d_cnt);
  PCD_LAUNCH_CHECK(ctx);
  unsigned long long cnt = 0;
  PCD_CUDA(ctx, cudaMemcpyAsync(erase, d_e, (size_t)n, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaMemcpyAsync(&cnt, d_cnt, 8, cudaMemcpyDeviceToHost, st));
  PCD_CUDA(ctx, cudaStreamSynchronize(st));
  if (n_erased) *n_erased = (int64_t)cnt;
  return PCD_OK;
}

// ---- allocation-free variant of pcd_residual_jacobian_batch_dev --------------------------------------------------
// For callers that may not allocate device memory (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  The per-block partial sums of sum r^2 -- the one piece of scratch k_resjac needs -- live in a
// file-scope device array instead of the context's growable buffer: 65 536 doubles = 512 KiB, held by every GPU that
// loads the library until the process exits.  That bounds one call at 65 536 * 256 = 16 777 216 constraints.
// All calls on a device share the array and it keeps the last call's contents.  Nothing has to start at zero: block b
// of a call writes slot b before the one-block sum reads slots [0, blocks), in stream order, so stale slots beyond a
// call's block count are never read.  Calls that OVERLAP on one GPU (different streams) race on it: serialise them.
}  // extern "C"

namespace pcd {
constexpr int64_t NOALLOC_MAX_BLOCKS = 65536;
constexpr int64_t NOALLOC_MAX_N = NOALLOC_MAX_BLOCKS * RJ_THREADS;
static __device__ __align__(256) double g_noalloc_partial[NOALLOC_MAX_BLOCKS];

// The array's name is a host-side stand-in in host code, not a device address: its device address is looked up once
// per device, with that device current, and cached.
constexpr int NOALLOC_MAX_DEVICES = 64;
static double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_noalloc_mu;

static cudaError_t noalloc_scratch(int device, double** out) {       // caller has made `device` current
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    double* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_partial);
    if (e != cudaSuccess) return e;
    if (!ptr) return cudaErrorInvalidSymbol;
    g_noalloc_ptr[device] = ptr;
  }
  *out = g_noalloc_ptr[device];
  return cudaSuccess;
}
}  // namespace pcd

extern "C" {

int64_t pcd_residual_jacobian_noalloc_max(void) { return pcd::NOALLOC_MAX_N; }

int pcd_residual_jacobian_batch_dev_noalloc(pcd_ctx* ctx, const double* d_p, const double* d_abcd, const double* d_w, int64_t n,
                                            double* d_r, double* d_J, double* d_sum_sq, void* stream) {
  if (!ctx || n < 0 || (n > 0 && (!d_p || !d_abcd || !d_w || !d_r)))
    PCD_FAIL(ctx, PCD_E_INVALID, "pcd_residual_jacobian_batch_dev_noalloc: bad argument");
  if (n > pcd::NOALLOC_MAX_N)
    PCD_FAIL(ctx, PCD_E_INVALID, "pcd_residual_jacobian_batch_dev_noalloc: %lld constraints exceed the fixed scratch (limit %lld per call); "
             "split the batch or use pcd_residual_jacobian_batch_dev", (long long)n, (long long)pcd::NOALLOC_MAX_N);
  DeviceGuard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) { if (d_sum_sq) PCD_CUDA(ctx, cudaMemsetAsync(d_sum_sq, 0, 8, st)); return PCD_OK; }
  const int nb = pcd_blocks(n, pcd::RJ_THREADS);
  double* partial = nullptr;
  if (d_sum_sq) PCD_CUDA(ctx, pcd::noalloc_scratch(ctx->device, &partial));
  pcd::k_resjac<<<nb, pcd::RJ_THREADS, 0, st>>>(d_p, (const double4*)d_abcd, d_w, n, d_r, d_J, partial);
  PCD_LAUNCH_CHECK(ctx);
  if (d_sum_sq) { pcd::k_sum_partials<<<1, 256, 0, st>>>(partial, nb, d_sum_sq); PCD_LAUNCH_CHECK(ctx); }
  return PCD_OK;
}

}  // extern "C"
