// pcd_b200_dropin.h -- header-only C++ mirror of Colmap-PCD's src/lidar interfaces over the C ABI
// (include/pcd_b200.h).  Same class names, method names, argument meaning and bool / abort-free error
// behaviour as the reference, so the call sites listed in INTEGRATION.md compile unchanged:
//
//   colmap::lidar::LidarPoint / LidarPointcloud(Ptr)   src/lidar/pt_type.h:14-45
//   colmap::lidar::Kdtree                               src/lidar/kdtree.h:13-22, kdtree.cc:5-21
//   colmap::lidar::PcdProjectionOptions / PcdProj       src/lidar/pcd_projection.h:31-66
//   colmap::lidar::PointCloudProcess                    src/lidar/ply.h:13-40, ply.cc:9-107
//   colmap::LidarPoint (the match record, full interface) src/lidar/lidar_point.h:9-52, lidar_point.cc:5-50
//   colmap::BundleAdjustmentLidarCostFunction::Create   src/base/cost_functions.h:180-232
//   colmap::LidarEvaluationBatch / LidarEvaluationCallback (new: the ceres::EvaluationCallback of SURVEY 8b, one batch per problem)
//   colmap::MatchClosestLidarPoints                     (new: the three serial association loops as one batched call)
//   colmap::Project2Image                               (new: src/optim/bundle_adjustment.cc:241-281 for all variable points at once)
// Several GPUs: PCD_B200_DEVICES="0,1,..." or DropinRuntime::SetDevices() shards the batched calls over pcd_group_*.
//
// Inside the reference tree define PCD_DROPIN_WITH_EIGEN / PCD_DROPIN_WITH_CERES before including
// this header: the real Eigen and Ceres types are then used.  Without them (this repository's own
// build, which has neither library) minimal same-shaped stand-ins are declared so that the shim can be
// compiled and tested.  PCL types are always replaced by the small containers below: only `points`,
// `width`, `height` and `size()` of pcl::PointCloud are used on this path.
//
// There is no CPU fallback anywhere in this file: every numeric result comes from libpcd_b200.so.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <array>
#include <cmath>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>
#include "../../include/pcd_b200.h"

#ifdef PCD_DROPIN_WITH_EIGEN
#include <Eigen/Core>
#include <Eigen/Geometry>
#else
namespace Eigen {   // stand-ins with the storage order and accessors the call sites use
template <typename T, int N> struct PcdVec {
  T v[N];
  PcdVec() : v{} {}
  PcdVec(std::initializer_list<T> l) : v{} { int i = 0; for (T x : l) if (i < N) v[i++] = x; }
  T& operator()(int i) { return v[i]; }
  const T& operator()(int i) const { return v[i]; }
  T& operator[](int i) { return v[i]; }
  const T& operator[](int i) const { return v[i]; }
  T* data() { return v; }
  const T* data() const { return v; }
  // the few Eigen operations the reference's association loops spell out (block(i,0,3,1), a.dot(b), v << a,b,c,d)
  PcdVec<T, 3> block(int i, int, int, int) const { PcdVec<T, 3> r; for (int k = 0; k < 3; k++) r.v[k] = v[i + k]; return r; }
  // three-element reductions in Eigen's order, a0 + (a1 + a2) (Redux.h splits at Length / 2)
  T dot(const PcdVec& o) const { if (N == 3) return v[0] * o.v[0] + (v[1] * o.v[1] + v[2] * o.v[2]); T s = T(0); for (int k = 0; k < N; k++) s += v[k] * o.v[k]; return s; }
  T norm() const { return std::sqrt(dot(*this)); }
  PcdVec operator-(const PcdVec& o) const { PcdVec r; for (int k = 0; k < N; k++) r.v[k] = v[k] - o.v[k]; return r; }
  struct Comma {
    PcdVec* m; int i;
    Comma operator,(T x) { if (i < N) m->v[i] = x; return Comma{m, i + 1}; }
  };
  Comma operator<<(T x) { v[0] = x; return Comma{this, 1}; }
};
using Vector2d = PcdVec<double, 2>;
using Vector3d = PcdVec<double, 3>;
using Vector4d = PcdVec<double, 4>;
using Vector6d = PcdVec<double, 6>;
using Vector3ub = PcdVec<uint8_t, 3>;
template <typename T, int R, int C> using Matrix = PcdVec<T, R * C>;
}  // namespace Eigen
#endif

#ifdef PCD_DROPIN_WITH_CERES
#include <ceres/ceres.h>
#else
namespace ceres {   // the two Ceres interfaces this path touches (ceres/cost_function.h, evaluation_callback.h)
class CostFunction {
 public:
  virtual ~CostFunction() {}
  virtual bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const = 0;
  const std::vector<int32_t>& parameter_block_sizes() const { return sizes_; }
  int num_residuals() const { return num_residuals_; }
 protected:
  std::vector<int32_t>* mutable_parameter_block_sizes() { return &sizes_; }
  void set_num_residuals(int n) { num_residuals_ = n; }
 private:
  std::vector<int32_t> sizes_;
  int num_residuals_ = 0;
};
class EvaluationCallback {
 public:
  virtual ~EvaluationCallback() {}
  virtual void PrepareForEvaluation(bool evaluate_jacobians, bool new_evaluation_point) = 0;
};
}  // namespace ceres
#endif

namespace lidarpt {
// src/lidar/pt_type.h:14-20: PCL_ADD_POINT4D + PCL_ADD_NORMAL4D = 2 x 16 bytes.
struct alignas(16) Point {
  float x = 0, y = 0, z = 0, pad0 = 1.0f;
  float normal_x = 0, normal_y = 0, normal_z = 0, pad1 = 0;
};
static_assert(sizeof(Point) == 32, "reference point layout is 32 bytes");
}  // namespace lidarpt

namespace colmap {

typedef uint64_t pcd_point3D_t;   // colmap::point3D_t (src/util/types.h)

namespace lidar {

typedef lidarpt::Point LidarPoint;
struct LidarPointcloud {            // the part of pcl::PointCloud<LidarPoint> used by src/lidar
  std::vector<LidarPoint> points;
  uint32_t width = 0, height = 1;
  size_t size() const { return points.size(); }
  void push_back(const LidarPoint& p) { points.push_back(p); }
};
typedef std::shared_ptr<LidarPointcloud> LidarPointcloudPtr;

// One CUDA context per process and device, shared by every shim object (the reference's objects are
// all owned by one controller thread; batch calls on a context are serialised by `mu`).
struct DropinRuntime {
  pcd_ctx* ctx = nullptr;            // the first device's context (all serial calls)
  pcd_group* group = nullptr;        // more than one device: the batched calls are sharded over the group
  std::vector<int> devices;          // SetDevices(), or the environment variable PCD_B200_DEVICES="0,1,2,3"; default: device 0
  std::mutex mu;
  std::string error;
  static DropinRuntime& Get() { static DropinRuntime r; return r; }
  void SetDevices(const std::vector<int>& d) { std::lock_guard<std::mutex> g(mu); if (!ctx) devices = d; }
  bool Ensure() {
    std::lock_guard<std::mutex> g(mu);
    if (ctx) return true;
    if (devices.empty()) {
      if (const char* e = getenv("PCD_B200_DEVICES")) { for (const char* p = e; *p;) { devices.push_back(atoi(p)); while (*p && *p != ',') p++; if (*p) p++; } }
      if (devices.empty()) devices.push_back(0);
    }
    if (devices.size() > 1) {
      if (pcd_group_create(devices.data(), (int)devices.size(), &group) != PCD_OK) { error = pcd_last_error(nullptr); group = nullptr; return false; }
      ctx = pcd_group_ctx(group, 0);
      return true;
    }
    if (pcd_ctx_create(devices[0], &ctx) != PCD_OK) { error = pcd_last_error(nullptr); ctx = nullptr; return false; }
    return true;
  }
  ~DropinRuntime() { if (group) pcd_group_destroy(group); else if (ctx) pcd_ctx_destroy(ctx); }
  // A context of the calling thread's own (streams and scratch are per context), for the small per-record calls a
  // serial call site or a Ceres worker thread makes: no process-wide lock is taken on that path.
  static pcd_ctx* Thread() {
    struct Holder { pcd_ctx* c = nullptr; ~Holder() { if (c) pcd_ctx_destroy(c); } };
    static thread_local Holder h;
    if (!h.c && pcd_ctx_create(Get().devices.empty() ? 0 : Get().devices[0], &h.c) != PCD_OK) h.c = nullptr;
    return h.c;
  }
};

// Device-resident cloud shared between Kdtree and PcdProj (the reference keeps three host copies).
struct DeviceCloud {
  pcd_cloud* h = nullptr;            // on the first device
  pcd_group_cloud* gc = nullptr;     // replicas on every device of the group (owns h once it exists)
  // after the index and the sub-map are built: copy them to the other devices of the group
  bool Replicate() {
    DropinRuntime& rt = DropinRuntime::Get();
    if (!rt.group || gc || !h) return true;
    std::lock_guard<std::mutex> g(rt.mu);
    if (pcd_group_replicate(rt.group, h, &gc) != PCD_OK) { rt.error = pcd_group_last_error(rt.group); gc = nullptr; return false; }
    return true;
  }
  ~DeviceCloud() { if (gc) pcd_group_cloud_destroy(gc); else if (h) pcd_cloud_destroy(h); }
};

struct PcdProjectionOptions {        // src/lidar/pcd_projection.h:31-47, same fields and defaults
  double depth_image_scale = 0.2;
  bool if_save_depth_image = false;
  std::string depth_image_folder;
  std::string original_image_folder;
  bool if_save_lidar_frame = false;
  std::string lidar_frame_folder;
  int max_proj_scale = 10;
  int min_proj_scale = 2;
  double min_proj_dist = 2;
  float submap_length = 1.0;
  float submap_width = 1.0;
  float submap_height = 1.0;
  float choose_meter = 40.0;
  double min_lidar_proj_dist = 0.5;
  pcd_proj_options ToC() const {
    pcd_proj_options o;
    o.depth_image_scale = depth_image_scale; o.max_proj_scale = max_proj_scale; o.min_proj_scale = min_proj_scale;
    o.min_proj_dist = min_proj_dist; o.submap_length = submap_length; o.submap_width = submap_width;
    o.submap_height = submap_height; o.choose_meter = choose_meter; o.min_lidar_proj_dist = min_lidar_proj_dist;
    return o;
  }
};

inline std::shared_ptr<DeviceCloud> UploadCloud(const LidarPointcloudPtr& ptr, bool direction_trans) {
  auto dc = std::make_shared<DeviceCloud>();
  DropinRuntime& rt = DropinRuntime::Get();
  if (!ptr || !rt.Ensure()) return dc;
  std::lock_guard<std::mutex> g(rt.mu);
  if (pcd_cloud_from_host(rt.ctx, reinterpret_cast<const float*>(ptr->points.data()), (int64_t)ptr->points.size(),
                          direction_trans ? 1 : 0, &dc->h) != PCD_OK) {
    rt.error = pcd_last_error(rt.ctx); dc->h = nullptr;
  }
  return dc;
}

// ---------------------------------------------------------------------------------------------
class Kdtree {                       // src/lidar/kdtree.h:13-22
 public:
  Kdtree() {}
  ~Kdtree() {}
  // kdtree.cc:5-8.  Uploads the cloud (index space = rows of *ptr) and builds the GPU grid index.
  void BuildMap(LidarPointcloudPtr& ptr) { global_pointcloud_ptr_ = ptr; Attach(UploadCloud(ptr, false)); }
  // Shares an already-uploaded cloud (used by PointCloudProcess so the cloud is resident once).
  void Attach(std::shared_ptr<DeviceCloud> dc) {
    dev_ = std::move(dc);
    if (!dev_ || !dev_->h) return;
    DropinRuntime& rt = DropinRuntime::Get();
    std::lock_guard<std::mutex> g(rt.mu);
    ready_ = pcd_index_build(dev_->h, 0.0f) == PCD_OK;
    if (!ready_) rt.error = pcd_last_error(rt.ctx);
  }
  // kdtree.cc:10-21: exact 1-NN; false when the search finds nothing (empty cloud / no device).
  bool GetClosestPoint(const LidarPoint& pt, LidarPoint& closest_point) {
    double q[3] = {pt.x, pt.y, pt.z};
    int32_t idx = -1;
    if (!GetClosestPoints(q, 1, &idx) || idx < 0) return false;
    closest_point = global_pointcloud_ptr_->points[(size_t)idx];
    return true;
  }
  // Batched form for the hoisted loops: idx[i] = -1 when none.
  bool GetClosestPoints(const double* q_xyz, int64_t n, int32_t* idx, float* d2 = nullptr) {
    if (!ready_) return false;
    DropinRuntime& rt = DropinRuntime::Get();
    std::lock_guard<std::mutex> g(rt.mu);
    if (pcd_nn_batch(dev_->h, q_xyz, nullptr, n, idx, d2) != PCD_OK) { rt.error = pcd_last_error(rt.ctx); return false; }
    return true;
  }
  void SetHostCloud(const LidarPointcloudPtr& p) { global_pointcloud_ptr_ = p; }
  pcd_cloud* handle() const { return dev_ ? dev_->h : nullptr; }
  pcd_group_cloud* group_handle() const { return dev_ ? dev_->gc : nullptr; }
 private:
  LidarPointcloudPtr global_pointcloud_ptr_;
  std::shared_ptr<DeviceCloud> dev_;
  bool ready_ = false;
};

// ---------------------------------------------------------------------------------------------
class PcdProj {                      // src/lidar/pcd_projection.h:49-66
 public:
  using PointType = LidarPoint;
  using MapType = LidarPointcloudPtr;
  explicit PcdProj(PcdProjectionOptions options) : options_(options) {}
  ~PcdProj() {}
  // pcd_projection.cc:223-255
  void BuildSubMap(const MapType& ptr) { global_map_ptr_ = ptr; Attach(UploadCloud(ptr, false)); }
  void Attach(std::shared_ptr<DeviceCloud> dc) {
    dev_ = std::move(dc);
    if (!dev_ || !dev_->h) return;
    DropinRuntime& rt = DropinRuntime::Get();
    std::lock_guard<std::mutex> g(rt.mu);
    pcd_proj_options o = options_.ToC();
    ready_ = pcd_submap_build(dev_->h, &o) == PCD_OK;
    if (!ready_) rt.error = pcd_last_error(rt.ctx);
  }
  void SetHostCloud(const MapType& p) { global_map_ptr_ = p; }

  // pcd_projection.cc:13-89.  ImageT / CameraT are colmap::Image / colmap::Camera in the reference
  // tree (Qvec(), Tvec(), Points2D() with HasPoint3D()/XY()/GetPoint3DId(); Params(), Width(), Height()).
  template <typename ImageT, typename CameraT, typename MapT>
  void SetNewImage(const ImageT& image, const CameraT& camera, MapT& map) {
    if (!ready_) return;
    std::vector<double> feat; std::vector<uint64_t> ids;
    for (const auto& p2 : image.Points2D()) {
      if (!p2.HasPoint3D()) continue;
      const auto xy = p2.XY();
      feat.push_back(xy(0)); feat.push_back(xy(1)); ids.push_back((uint64_t)p2.GetPoint3DId());
    }
    std::vector<int32_t> win(ids.size()); std::vector<float> rng(ids.size());
    if (!Project(image, camera, feat, win, rng)) return;
    for (size_t k = 0; k < ids.size(); k++) {
      if (win[k] < 0) continue;
      const LidarPoint& l = global_map_ptr_->points[(size_t)win[k]];
      typename MapT::mapped_type v;
      v(0) = l.x; v(1) = l.y; v(2) = l.z; v(3) = l.normal_x; v(4) = l.normal_y; v(5) = l.normal_z;
      map.insert({(typename MapT::key_type)ids[k], v});       // map.insert keeps the first, as :75
    }
  }
  // Batched form of the loop in BundleAdjustmentConfig::Project2Image (src/optim/bundle_adjustment.cc:241-281): the
  // caller collects the images that are not cached yet (its own decision logic, :247-258, is unchanged) and projects
  // them in ONE call (pcd_project_images); cache[image_id] receives what SetNewImage(image, camera, map) would fill.
  template <typename ImageT, typename CameraT, typename CacheT>
  void SetNewImages(const std::vector<const ImageT*>& images, const std::vector<const CameraT*>& cameras,
                    const std::vector<uint64_t>& image_ids, CacheT& cache) {
    if (!ready_ || images.empty()) return;
    std::vector<pcd_image> ims; std::vector<double> feat; std::vector<uint64_t> ids; std::vector<int64_t> start(1, 0);
    for (size_t k = 0; k < images.size(); k++) {
      ims.push_back(MakeImage(*images[k], *cameras[k]));
      for (const auto& p2 : images[k]->Points2D()) {
        if (!p2.HasPoint3D()) continue;
        const auto xy = p2.XY();
        feat.push_back(xy(0)); feat.push_back(xy(1)); ids.push_back((uint64_t)p2.GetPoint3DId());
      }
      start.push_back((int64_t)ids.size());
    }
    std::vector<int32_t> win(ids.size()); std::vector<float> rng(ids.size());
    DropinRuntime& rt = DropinRuntime::Get();
    {
      std::lock_guard<std::mutex> g(rt.mu);
      pcd_proj_options o = options_.ToC();
      double* coef = SplatCoef(ims[0], o);
      const int rc = dev_->gc ? pcd_group_project_images(dev_->gc, ims.data(), (int64_t)ims.size(), coef, feat.data(), start.data(), win.data(), rng.data(), nullptr)
                              : pcd_project_images(dev_->h, ims.data(), (int64_t)ims.size(), coef, feat.data(), start.data(), win.data(), rng.data(), nullptr);
      if (rc != PCD_OK) { rt.error = pcd_last_error(rt.ctx); return; }
    }
    for (size_t k = 0; k < images.size(); k++) {
      auto& map = cache[(typename CacheT::key_type)image_ids[k]];
      for (int64_t f = start[k]; f < start[k + 1]; f++) {
        if (win[f] < 0) continue;
        const LidarPoint& l = global_map_ptr_->points[(size_t)win[f]];
        typename CacheT::mapped_type::mapped_type v;
        v(0) = l.x; v(1) = l.y; v(2) = l.z; v(3) = l.normal_x; v(4) = l.normal_y; v(5) = l.normal_z;
        map.insert({(typename CacheT::mapped_type::key_type)ids[f], v});
      }
    }
  }

  // pcd_projection.cc:102-220 (initial image pair): pt_xys[k].second = matched, pt_xyzs[k] = camera-frame point.
  template <typename ImageT, typename CameraT, typename XYs, typename XYZs>
  void SetNewImage(const ImageT& image, const CameraT& camera, XYs& pt_xys, XYZs& pt_xyzs) {
    if (!ready_) return;
    std::vector<double> feat;
    for (const auto& p : pt_xys) { feat.push_back(p.first(0)); feat.push_back(p.first(1)); }
    const size_t n = pt_xys.size();
    std::vector<int32_t> win(n); std::vector<float> rng(n);
    if (!Project(image, camera, feat, win, rng)) return;
    std::vector<uint8_t> matched(n); std::vector<double> xyz(3 * n);
    pcd_image im = MakeImage(image, camera);
    DropinRuntime& rt = DropinRuntime::Get();
    {
      std::lock_guard<std::mutex> g(rt.mu);
      if (pcd_ray_plane_batch(dev_->h, &im, feat.data(), win.data(), (int64_t)n, matched.data(), xyz.data()) != PCD_OK) {
        rt.error = pcd_last_error(rt.ctx); return;
      }
    }
    for (size_t k = 0; k < n; k++) {                       // one push_back per keypoint, zero vector when unmatched (:147-205)
      pt_xys[k].second = matched[k] != 0;
      typename XYZs::value_type v;
      v(0) = xyz[3 * k]; v(1) = xyz[3 * k + 1]; v(2) = xyz[3 * k + 2];
      pt_xyzs.push_back(v);
    }
  }
  pcd_cloud* handle() const { return dev_ ? dev_->h : nullptr; }

 private:
  template <typename ImageT, typename CameraT>
  static pcd_image MakeImage(const ImageT& image, const CameraT& camera) {
    pcd_image im;
    const auto q = image.Qvec(); const auto t = image.Tvec();
    for (int i = 0; i < 4; i++) im.qvec[i] = q(i);
    for (int i = 0; i < 3; i++) im.tvec[i] = t(i);
    const std::vector<double> params = camera.Params();
    for (int i = 0; i < 8; i++) im.cam[i] = i < (int)params.size() ? params[i] : 0.0;
    im.width = (int32_t)camera.Width(); im.height = (int32_t)camera.Height();
    return im;
  }
  // the function-local statics a_x b_x a_y b_y of pcd_projection.cc:391-397: fixed by the first camera of the process
  static double* SplatCoef(const pcd_image& im, const pcd_proj_options& o) {
    static double coef[4]; static bool have_coef = false;
    if (!have_coef) { pcd_splat_coef(im.cam, &o, coef); have_coef = true; }
    return coef;
  }
  template <typename ImageT, typename CameraT>
  bool Project(const ImageT& image, const CameraT& camera, const std::vector<double>& feat, std::vector<int32_t>& win,
               std::vector<float>& rng) {
    pcd_image im = MakeImage(image, camera);
    DropinRuntime& rt = DropinRuntime::Get();
    std::lock_guard<std::mutex> g(rt.mu);
    pcd_proj_options o = options_.ToC();
    double* coef = SplatCoef(im, o);
    if (pcd_project_image(dev_->h, &im, coef, feat.data(), (int64_t)win.size(), win.data(), rng.data(), nullptr, nullptr) != PCD_OK) {
      rt.error = pcd_last_error(rt.ctx); return false;
    }
    return true;
  }
  MapType global_map_ptr_;
  PcdProjectionOptions options_;
  std::shared_ptr<DeviceCloud> dev_;
  bool ready_ = false;
};

// ---------------------------------------------------------------------------------------------
class PointCloudProcess {            // src/lidar/ply.h:13-40
 public:
  PointCloudProcess(const std::string& path) : path_(path) {}
  ~PointCloudProcess() {}
  // ply.cc:9-31: false when the file cannot be read (ply.cc:14-17) or no device is usable.
  const bool Initialize(const PcdProjectionOptions& pp_options) {
    DropinRuntime& rt = DropinRuntime::Get();
    if (!rt.Ensure()) return false;
    float* raw = nullptr; int64_t n = 0;
    if (pcd_ply_read(path_.c_str(), &raw, &n) != PCD_OK) { rt.error = pcd_last_error(nullptr); return false; }
    auto dc = std::make_shared<DeviceCloud>();
    int rc;
    {
      std::lock_guard<std::mutex> g(rt.mu);
      rc = pcd_cloud_from_host(rt.ctx, raw, n, 1, &dc->h);       // PointCloudDirectionTrans on the device
    }
    pcd_free(raw);
    if (rc != PCD_OK) { rt.error = pcd_last_error(rt.ctx); return false; }
    // host mirror of the post-swap cloud: GetClosestPoint returns full points by value, as the reference
    global_pcd_ptr_ = std::make_shared<LidarPointcloud>();
    const int64_t m = pcd_cloud_size(dc->h);
    global_pcd_ptr_->points.resize((size_t)m);
    global_pcd_ptr_->width = (uint32_t)m;
    {
      std::lock_guard<std::mutex> g(rt.mu);
      if (pcd_cloud_download(dc->h, reinterpret_cast<float*>(global_pcd_ptr_->points.data())) != PCD_OK) return false;
    }
    pcd_proj_ = std::make_shared<PcdProj>(pp_options);
    pcd_proj_->SetHostCloud(global_pcd_ptr_);
    pcd_proj_->Attach(dc);
    kdtree_ptr_ = std::make_shared<Kdtree>();
    kdtree_ptr_->SetHostCloud(global_pcd_ptr_);
    kdtree_ptr_->Attach(dc);
    if (!dc->Replicate()) return false;                       // several devices: index and sub-map copied, not rebuilt
    return pcd_proj_->handle() != nullptr && kdtree_ptr_->handle() != nullptr;
  }
  // ply.cc:90-107
  bool SearchNearestNeiborByKdtree(const Eigen::Vector3d& point_3d, Eigen::Vector6d& l_pt) {
    double q[3] = {point_3d(0), point_3d(1), point_3d(2)}, l[6]; uint8_t ok = 0;
    if (!SearchNearestNeiborBatch(q, 1, l, &ok) || !ok) return false;
    for (int i = 0; i < 6; i++) l_pt(i) = l[i];
    return true;
  }
  // Batched form: l_pt = 6 doubles per query, ok = the per-query bool.
  bool SearchNearestNeiborBatch(const double* q_xyz, int64_t n, double* l_pt, uint8_t* ok) {
    if (!kdtree_ptr_ || !kdtree_ptr_->handle()) return false;
    DropinRuntime& rt = DropinRuntime::Get();
    std::lock_guard<std::mutex> g(rt.mu);
    if (pcd_search_nn_batch(kdtree_ptr_->handle(), q_xyz, n, l_pt, ok) != PCD_OK) { rt.error = pcd_last_error(rt.ctx); return false; }
    return true;
  }
  LidarPointcloudPtr GetCloud() const { return global_pcd_ptr_; }

  std::shared_ptr<PcdProj> pcd_proj_;
  std::shared_ptr<Kdtree> kdtree_ptr_;

 private:
  std::string path_;
  LidarPointcloudPtr global_pcd_ptr_;
};

}  // namespace lidar

// ---------------------------------------------------------------------------------------------
enum class LidarPointType { Proj, Icp, IcpGround };       // src/lidar/lidar_point.h:9

// The match record (src/lidar/lidar_point.h:10-52), full interface.  Records produced by the batched association
// (pcd_associate_batch) are installed with FromBatch() and carry every field already.  The constructors and the
// Compute* methods exist for call sites that stay serial (src/optim/bundle_adjustment.cc:368-408,
// src/sfm/incremental_mapper.cc:1432-1466, src/controllers/bundle_adjustment.cc:142-183, src/base/reconstruction.cc:787):
// like the reference both constructors Normalize(); the arithmetic runs in pcd_lidar_point_ops (one record per call, on
// the calling thread's own context), the same device code the batched record kernel uses, so serial and batched
// results are bit-identical.  (Inside the reference tree the class of src/lidar/lidar_point.{h,cc} -- no PCL in it -- may
// simply be kept: define PCD_DROPIN_KEEP_LIDAR_POINT and include it before this header.)
#ifndef PCD_DROPIN_KEEP_LIDAR_POINT
class LidarPoint {
 public:
  LidarPoint() {}
  LidarPoint(LidarPointType type, Eigen::Vector3d& xyz, Eigen::Vector4d& abcd) : type_(type), xyz_(xyz), abcd_(abcd) { Normalize(); }
  LidarPoint(Eigen::Vector3d& xyz, Eigen::Vector4d& abcd) : xyz_(xyz), abcd_(abcd) { Normalize(); }
  ~LidarPoint() {}
  static LidarPoint FromBatch(const pcd_assoc_out& o, int64_t i) {
    LidarPoint p;
    p.type_ = o.type[i] == PCD_TYPE_ICP_GROUND ? LidarPointType::IcpGround
              : (o.type[i] == PCD_TYPE_ICP ? LidarPointType::Icp : LidarPointType::Proj);
    for (int k = 0; k < 3; k++) p.xyz_(k) = o.xyz[3 * i + k];
    for (int k = 0; k < 4; k++) p.abcd_(k) = o.abcd[4 * i + k];
    if (o.rgb) for (int k = 0; k < 3; k++) p.color_(k) = o.rgb[3 * i + k];
    if (o.dist) p.dist_ = o.dist[i];
    if (o.angle) p.angle_ = o.angle[i];
    return p;
  }
  static LidarPoint FromFields(LidarPointType type, const double xyz[3], const double abcd[4], double dist, double angle, uint8_t r, uint8_t g, uint8_t b) {
    LidarPoint p;
    p.type_ = type;
    for (int k = 0; k < 3; k++) p.xyz_(k) = xyz[k];
    for (int k = 0; k < 4; k++) p.abcd_(k) = abcd[k];
    p.dist_ = dist; p.angle_ = angle;
    p.color_(0) = r; p.color_(1) = g; p.color_(2) = b;
    return p;
  }
  void SetType(LidarPointType type) { type_ = type; }
  void SetColor(Eigen::Vector3ub& color) { color_ = color; }
  void SetDist(const double& dist) { dist_ = dist; }
  void SetAngle(const double& angle) { angle_ = angle; }
  Eigen::Vector3d LidarXYZ() { return xyz_; }
  const Eigen::Vector3d LidarXYZ() const { return xyz_; }
  Eigen::Vector4d LidarABCD() { return abcd_; }
  const Eigen::Vector4d LidarABCD() const { return abcd_; }
  LidarPointType Type() { return type_; }
  LidarPointType Type() const { return type_; }
  const double Dist() { return dist_; }
  const double Angle() { return angle_; }
  const Eigen::Vector3ub& Color() const { return color_; }
  Eigen::Vector3ub& Color() { return color_; }
  const double ComputeDist(Eigen::Vector3d& pt) { double d = NaN(); Ops(&pt, 0, nullptr, &d, nullptr, nullptr); return d; }
  const double ComputePointToPointDist(Eigen::Vector3d& pt) { double d = NaN(); Ops(&pt, 0, nullptr, nullptr, &d, nullptr); return d; }
  const double ComputeAngle(Eigen::Vector3d& pt) { double a = NaN(); Ops(&pt, 0, nullptr, nullptr, nullptr, &a); return a; }
  void Normalize() { double out[4]; if (Ops(nullptr, 1, out, nullptr, nullptr, nullptr)) for (int k = 0; k < 4; k++) abcd_(k) = out[k]; }
 private:
  static double NaN() { return std::nan(""); }
  // without a usable device the record keeps its fields / the methods return NaN: nothing is computed on the host
  bool Ops(const Eigen::Vector3d* pt, int normalize, double* abcd_out, double* d_plane, double* d_point, double* angle) const {
    pcd_ctx* c = lidar::DropinRuntime::Thread();
    if (!c) return false;
    const double xyz[3] = {xyz_(0), xyz_(1), xyz_(2)}, pl[4] = {abcd_(0), abcd_(1), abcd_(2), abcd_(3)};
    double p[3] = {0, 0, 0};
    if (pt) for (int k = 0; k < 3; k++) p[k] = (*pt)(k);
    return pcd_lidar_point_ops(c, xyz, pl, pt ? p : nullptr, 1, normalize, abcd_out, d_plane, d_point, angle) == PCD_OK;
  }
  LidarPointType type_ = LidarPointType::Icp;
  Eigen::Vector3d xyz_;
  Eigen::Vector4d abcd_;
  Eigen::Vector3ub color_;
  double dist_ = 0, angle_ = 0;
};
#endif

// BundleAdjustmentConfig::Project2Image (src/optim/bundle_adjustment.cc:241-281) for a whole set of variable points:
// the reference's decision, element by element (pcd_project2image_select: skip an image whose pair with the new image
// has <= match_features_threshold correspondences, :247-253; skip what lidar_searched_image_ids_ already holds), then
// ONE batched projection of everything selected instead of one SetNewImage per image.  ReconstructionT is
// colmap::Reconstruction (Point3D(id).Track().Elements(), ExistsImagePair, ImagePair().num_total_corrs, Image(), Camera()).
template <typename ReconstructionT, typename PointIds, typename CacheT>
inline size_t Project2Image(lidar::PointCloudProcess& pcp, ReconstructionT* reconstruction, const PointIds& point3D_ids,
                            uint32_t image_id, int match_features_threshold, CacheT& lidar_searched_image_ids_) {
  std::vector<uint32_t> todo;
  for (const auto point3D_id : point3D_ids) {
    const auto& track_els = reconstruction->Point3D(point3D_id).Track().Elements();
    const int64_t n = (int64_t)track_els.size();
    std::vector<uint32_t> img((size_t)n); std::vector<int64_t> corrs((size_t)n); std::vector<uint8_t> cached((size_t)n), sel((size_t)n);
    for (int64_t k = 0; k < n; k++) {
      const uint32_t id2 = (uint32_t)track_els[(size_t)k].image_id;
      img[(size_t)k] = id2;
      corrs[(size_t)k] = (id2 != image_id && reconstruction->ExistsImagePair(image_id, id2)) ? (int64_t)reconstruction->ImagePair(image_id, id2).num_total_corrs : -1;
      bool have = lidar_searched_image_ids_.find(id2) != lidar_searched_image_ids_.end();
      for (uint32_t t : todo) have = have || t == id2;
      cached[(size_t)k] = have ? 1 : 0;
    }
    if (pcd_project2image_select(img.data(), n, image_id, corrs.data(), match_features_threshold, cached.data(), sel.data()) < 0) return 0;
    for (int64_t k = 0; k < n; k++) if (sel[(size_t)k]) todo.push_back(img[(size_t)k]);
  }
  if (todo.empty() || !pcp.pcd_proj_) return 0;
  typedef typename std::remove_reference<decltype(reconstruction->Image(todo[0]))>::type ImageT;
  typedef typename std::remove_reference<decltype(reconstruction->Camera(reconstruction->Image(todo[0]).CameraId()))>::type CameraT;
  std::vector<const ImageT*> images; std::vector<const CameraT*> cameras; std::vector<uint64_t> ids;
  for (uint32_t id : todo) {
    const ImageT& image = reconstruction->Image(id);
    images.push_back(&image); cameras.push_back(&reconstruction->Camera(image.CameraId())); ids.push_back(id);
  }
  pcp.pcd_proj_->template SetNewImages<ImageT, CameraT>(images, cameras, ids, lidar_searched_image_ids_);
  return todo.size();
}

// BundleAdjustmentConfig::MatchVariablePoint2LidarPoint (src/optim/bundle_adjustment.cc:288-350) for many points at once:
// per point the candidates are the projected lidar points its track's cached images hold for it, in track order; the
// kernel keeps the one with the smallest |v . n| / (|n| |v|) (strict <, start value 360) and builds the Proj record
// (Normalize, dist = ComputeDist, angle = ComputeAngle; colour 255,0,0 as :343-345).
template <typename ReconstructionT, typename PointIds, typename CacheT>
inline bool MatchVariablePoints2LidarPoints(ReconstructionT* reconstruction, const PointIds& point3D_ids,
                                            const CacheT& lidar_searched_image_ids_, std::map<uint64_t, LidarPoint>* out) {
  std::vector<double> p, cand; std::vector<int64_t> start(1, 0); std::vector<uint64_t> ids;
  for (const auto point3D_id : point3D_ids) {
    auto& point3D = reconstruction->Point3D(point3D_id);
    const auto xyz = point3D.XYZ();
    for (int k = 0; k < 3; k++) p.push_back(xyz(k));
    for (const auto& track_el : point3D.Track().Elements()) {
      auto iter = lidar_searched_image_ids_.find(track_el.image_id);
      if (iter == lidar_searched_image_ids_.end()) continue;
      auto lidar_pt_ptr = iter->second.find(point3D_id);
      if (lidar_pt_ptr == iter->second.end()) continue;
      for (int k = 0; k < 6; k++) cand.push_back(lidar_pt_ptr->second(k));
    }
    start.push_back((int64_t)cand.size() / 6); ids.push_back((uint64_t)point3D_id);
  }
  const int64_t n = (int64_t)ids.size();
  if (n == 0) return true;
  std::vector<uint8_t> found((size_t)n); std::vector<double> xyz(3 * (size_t)n), abcd(4 * (size_t)n), dist((size_t)n), angle((size_t)n);
  lidar::DropinRuntime& rt = lidar::DropinRuntime::Get();
  if (!rt.Ensure()) return false;
  {
    std::lock_guard<std::mutex> g(rt.mu);
    if (pcd_match_proj_batch(rt.ctx, p.data(), cand.data(), start.data(), n, found.data(), xyz.data(), abcd.data(), dist.data(), angle.data()) != PCD_OK) {
      rt.error = pcd_last_error(rt.ctx); return false;
    }
  }
  for (int64_t i = 0; i < n; i++)
    if (found[(size_t)i])
      out->insert({ids[(size_t)i], LidarPoint::FromFields(LidarPointType::Proj, &xyz[3 * (size_t)i], &abcd[4 * (size_t)i], dist[(size_t)i], angle[(size_t)i], 255, 0, 0)});
  return true;
}

// Batched replacement of the three serial association loops (src/sfm/incremental_mapper.cc:1138-1166,
// :1411-1470; src/controllers/bundle_adjustment.cc:126-185).  ids / xyz: the variable points in the
// caller's iteration order; returns the accepted matches keyed by point id.
inline bool MatchClosestLidarPoints(lidar::PointCloudProcess& pcp, const std::vector<uint64_t>& ids, const std::vector<double>& xyz,
                                    const std::vector<int32_t>& global_opt_num, double max_r, double min_r, double drop,
                                    int variant, std::map<uint64_t, LidarPoint>* out) {
  const int64_t n = (int64_t)ids.size();
  if (!pcp.kdtree_ptr_ || !pcp.kdtree_ptr_->handle() || (int64_t)xyz.size() != 3 * n) return false;
  std::vector<int32_t> idx(n); std::vector<uint8_t> acc(n), type(n), rgb(3 * n);
  std::vector<double> lx(3 * n), abcd(4 * n), dist(n), angle(n);
  pcd_assoc_out o = {idx.data(), acc.data(), type.data(), rgb.data(), lx.data(), abcd.data(), dist.data(), angle.data()};
  pcd_range_sweep sw = {max_r, min_r, drop};
  lidar::DropinRuntime& rt = lidar::DropinRuntime::Get();
  {
    std::lock_guard<std::mutex> g(rt.mu);
    const bool sweep = variant != PCD_ASSOC_WHOLE_MAP;
    pcd_group_cloud* gc = pcp.kdtree_ptr_->group_handle();
    const int rc = gc ? pcd_group_associate_evaluate_batch(gc, xyz.data(), nullptr, sweep ? global_opt_num.data() : nullptr, sweep ? &sw : nullptr,
                                                           n, variant, 0.0, 0.0, 0.0, &o, nullptr, nullptr, nullptr)
                      : pcd_associate_batch(pcp.kdtree_ptr_->handle(), xyz.data(), nullptr, sweep ? global_opt_num.data() : nullptr,
                                            sweep ? &sw : nullptr, n, variant, &o);
    if (rc != PCD_OK) { rt.error = gc ? pcd_group_last_error(rt.group) : pcd_last_error(rt.ctx); return false; }
  }
  for (int64_t i = 0; i < n; i++)
    if (acc[i]) out->insert({ids[i], LidarPoint::FromBatch(o, i)});
  return true;
}

// ---------------------------------------------------------------------------------------------
// Cost function + evaluation callback (src/base/cost_functions.h:180-232).
// A LidarEvaluationBatch belongs to ONE ceres::Problem.  The bundle adjuster opens it before it adds residual blocks
// (LidarEvaluationBatch::Begin(), which makes it the calling thread's current batch), every cost function created by
// Create() takes a slot in it, the caller binds the slot to the point's own storage (SetPointPtr -- the same pointer it
// hands to AddResidualBlock, src/optim/bundle_adjustment.cc:1037) and installs a LidarEvaluationCallback(batch) in
// Problem::Options.  PrepareForEvaluation gathers the current points from those pointers (Ceres has copied its state
// into them), runs pcd_residual_jacobian_batch once and fills r / J; Evaluate() is then a read-only copy of its slot,
// safe from any number of Ceres threads.  The pointer Ceres passes to Evaluate() is its own state vector, not the
// block's storage, so Evaluate() never records it: it only compares VALUES with the prepared point.  An evaluation
// at another point (no callback installed, a probe between Prepare calls) is a one-element batch on the calling
// thread's own context -- slower, lock-free, never a host formula.
class LidarEvaluationBatch {
 public:
  int64_t Add(const double abcd[4], double weight, const double* point_ptr) {
    for (int k = 0; k < 4; k++) abcd_.push_back(abcd[k]);
    w_.push_back(weight); ptr_.push_back(point_ptr);
    r_.push_back(0.0); J_.insert(J_.end(), 3, 0.0);
    p_.insert(p_.end(), 3, 0.0);
    return (int64_t)w_.size() - 1;
  }
  void SetPointPtr(int64_t slot, const double* p) { ptr_[(size_t)slot] = p; }
  bool Prepare(bool jacobians) {
    const int64_t n = (int64_t)w_.size();
    prepared_ = false;
    for (int64_t i = 0; i < n; i++) {
      if (!ptr_[i]) return false;                           // an unbound slot: the caller forgot SetPointPtr
      for (int k = 0; k < 3; k++) p_[3 * i + k] = ptr_[i][k];
    }
    lidar::DropinRuntime& rt = lidar::DropinRuntime::Get();
    if (!rt.Ensure()) return false;
    std::lock_guard<std::mutex> g(rt.mu);
    prepared_ = pcd_residual_jacobian_batch(rt.ctx, p_.data(), abcd_.data(), w_.data(), n, r_.data(),
                                            jacobians ? J_.data() : nullptr, &sum_sq_) == PCD_OK;
    has_J_ = prepared_ && jacobians;
    return prepared_;
  }
  bool Matches(int64_t slot, const double* p) const {
    return prepared_ && p_[3 * slot] == p[0] && p_[3 * slot + 1] == p[1] && p_[3 * slot + 2] == p[2];
  }
  bool has_jacobians() const { return has_J_; }
  double r(int64_t s) const { return r_[(size_t)s]; }
  const double* J(int64_t s) const { return &J_[3 * (size_t)s]; }
  const double* abcd(int64_t s) const { return &abcd_[4 * (size_t)s]; }
  double w(int64_t s) const { return w_[(size_t)s]; }
  double sum_sq() const { return sum_sq_; }
  int64_t size() const { return (int64_t)w_.size(); }
  // The batch new cost functions attach to (per thread: problems are built by one thread, src/optim/bundle_adjustment.cc:452-475).
  static std::shared_ptr<LidarEvaluationBatch>& Current() { static thread_local std::shared_ptr<LidarEvaluationBatch> b; if (!b) b = std::make_shared<LidarEvaluationBatch>(); return b; }
  // Opens the batch of the next problem; the previous one lives on for as long as its cost functions do.
  static std::shared_ptr<LidarEvaluationBatch> Begin() { Current() = std::make_shared<LidarEvaluationBatch>(); return Current(); }
  static void Reset() { Begin(); }
 private:
  std::vector<double> abcd_, w_, p_, r_, J_;
  std::vector<const double*> ptr_;
  double sum_sq_ = 0;
  bool prepared_ = false, has_J_ = false;
};

class LidarEvaluationCallback : public ceres::EvaluationCallback {
 public:
  explicit LidarEvaluationCallback(std::shared_ptr<LidarEvaluationBatch> b = LidarEvaluationBatch::Current()) : batch_(std::move(b)) {}
  void PrepareForEvaluation(bool evaluate_jacobians, bool /*new_evaluation_point*/) override { batch_->Prepare(evaluate_jacobians); }
 private:
  std::shared_ptr<LidarEvaluationBatch> batch_;
};

class BundleAdjustmentLidarCostFunction : public ceres::CostFunction {
 public:
  // cost_functions.h:180: same signature; ownership of the returned object passes to ceres::Problem.  Bind the
  // block's storage afterwards: batch->SetPointPtr(fn->slot(), point3D.XYZ().data()), or use the three-argument form.
  static ceres::CostFunction* Create(const Eigen::Matrix<double, 4, 1>& abcd, const double& weight) {
    return new BundleAdjustmentLidarCostFunction(abcd, weight, nullptr, LidarEvaluationBatch::Current());
  }
  static ceres::CostFunction* Create(const Eigen::Matrix<double, 4, 1>& abcd, const double& weight, const double* point_xyz) {
    return new BundleAdjustmentLidarCostFunction(abcd, weight, point_xyz, LidarEvaluationBatch::Current());
  }
  bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const override {
    const double* p = parameters[0];
    const bool want_J = jacobians && jacobians[0];
    if (batch_->Matches(slot_, p) && (!want_J || batch_->has_jacobians())) {       // read-only
      residuals[0] = batch_->r(slot_);
      if (want_J) for (int k = 0; k < 3; k++) jacobians[0][k] = batch_->J(slot_)[k];
      return true;
    }
    // not prepared for this point: one-element batch on this thread's own context
    pcd_ctx* c = lidar::DropinRuntime::Thread();
    if (!c) return false;
    double w = batch_->w(slot_), J[3];
    if (pcd_residual_jacobian_batch(c, p, batch_->abcd(slot_), &w, 1, residuals, want_J ? J : nullptr, nullptr) != PCD_OK) return false;
    if (want_J) for (int k = 0; k < 3; k++) jacobians[0][k] = J[k];
    return true;
  }
  int64_t slot() const { return slot_; }
  const std::shared_ptr<LidarEvaluationBatch>& batch() const { return batch_; }
 private:
  BundleAdjustmentLidarCostFunction(const Eigen::Matrix<double, 4, 1>& abcd, double weight, const double* point_xyz,
                                    std::shared_ptr<LidarEvaluationBatch> b)
      : batch_(std::move(b)) {
    const double a[4] = {abcd(0), abcd(1), abcd(2), abcd(3)};
    slot_ = batch_->Add(a, weight, point_xyz);
    set_num_residuals(1);
    mutable_parameter_block_sizes()->push_back(3);
  }
  std::shared_ptr<LidarEvaluationBatch> batch_;
  int64_t slot_ = 0;
};

}  // namespace colmap
