// dropin_check.cpp -- drives the header-only drop-in shim exactly the way the reference's call sites
// do (src/sfm/incremental_mapper.cc:201,607,1146-1164; src/optim/bundle_adjustment.cc:993-1040) and
// dumps what it got, so tests/test_gpu_parity.py can compare it with the oracle.  Compiled by
// __graft_entry__.build() (needs no GPU to compile; needs one to run).
//
//   dropin_check cloud.ply queries.f64 out.f64
//     queries.f64: n x [x y z opt_num] doubles;  out.f64: see the writes below.
#include <stdio.h>
#include <stdlib.h>
#include "pcd_b200_dropin.h"

using namespace colmap;

// minimal stand-ins for colmap::Image / Camera / Point2D (src/base/image.h, camera.h, point2d.h)
struct MockPoint2D {
  Eigen::Vector2d xy; uint64_t id; bool has;
  bool HasPoint3D() const { return has; }
  Eigen::Vector2d XY() const { return xy; }
  uint64_t GetPoint3DId() const { return id; }
};
struct MockImage {
  Eigen::Vector4d q; Eigen::Vector3d t; std::vector<MockPoint2D> pts;
  Eigen::Vector4d Qvec() const { return q; }
  Eigen::Vector3d Tvec() const { return t; }
  const std::vector<MockPoint2D>& Points2D() const { return pts; }
};
struct MockCamera {
  std::vector<double> params; size_t w, h;
  std::vector<double> Params() const { return params; }
  size_t Width() const { return w; }
  size_t Height() const { return h; }
};

static std::vector<double> read_f64(const char* path) {
  std::vector<double> v;
  FILE* f = fopen(path, "rb");
  if (!f) return v;
  fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
  v.resize((size_t)sz / 8);
  if (fread(v.data(), 8, v.size(), f) != v.size()) v.clear();
  fclose(f);
  return v;
}

// ---- the reference's three serial association loop bodies, transcribed: only the surrounding containers
// (Reconstruction, BundleAdjustmentConfig) are replaced by a std::map; every statement that touches
// PointCloudProcess, Eigen or colmap::LidarPoint is the reference's own text.
// src/optim/bundle_adjustment.cc:358-410 (MatchClosestLidarPoint, local BA)
static void LoopLocal(lidar::PointCloudProcess* point_cloud_process_, uint64_t point3D_id, Eigen::Vector3d pt_xyz,
                      double& max_search_range, std::map<uint64_t, LidarPoint>& lidar_maps_) {
  Eigen::Vector6d lidar_pt;
  if (point_cloud_process_->SearchNearestNeiborByKdtree(pt_xyz,lidar_pt)){
    Eigen::Vector3d norm = lidar_pt.block(3,0,3,1);
    Eigen::Vector3d l_pt = lidar_pt.block(0,0,3,1);
    double d = 0 - l_pt.dot(norm);
    Eigen::Vector4d plane;
    plane << norm(0),norm(1),norm(2),d;

    LidarPoint lidar_point(l_pt,plane);
    if (std::abs(norm(1)/norm(0))>10 && std::abs(norm(1)/norm(2))>10) {
      lidar_point.SetType(LidarPointType::IcpGround);
      Eigen::Vector3ub color;
      color << 255,255,0;
      lidar_point.SetColor(color);
    } else {
      lidar_point.SetType(LidarPointType::Icp);
      Eigen::Vector3ub color;
      color << 0,255,0;
      lidar_point.SetColor(color);
    }

    double dist = lidar_point.ComputePointToPointDist(pt_xyz);
    if (dist > max_search_range) return;

    lidar_point.SetDist(dist);
    lidar_point.SetAngle( lidar_point.ComputeAngle(pt_xyz));

    lidar_maps_.insert({point3D_id,lidar_point});
  }
}
// src/sfm/incremental_mapper.cc:1432-1466 (global BA by lidar)
static void LoopGlobal(lidar::PointCloudProcess* lidar_pointcloud_process_, uint64_t point3D_id, Eigen::Vector3d pt_xyz,
                       double max_search_range, std::map<uint64_t, LidarPoint>& lidar_maps_) {
      Eigen::Vector6d lidar_pt;
      if (lidar_pointcloud_process_->SearchNearestNeiborByKdtree(pt_xyz,lidar_pt)) {
        Eigen::Vector3d norm = lidar_pt.block(3,0,3,1);
        Eigen::Vector3d l_pt = lidar_pt.block(0,0,3,1);

        double d = 0 - l_pt.dot(norm);
        Eigen::Vector4d plane;
        plane << norm(0),norm(1),norm(2),d;

        LidarPoint lidar_point(l_pt,plane);

        if (std::abs(norm(1)/norm(0))>10 && std::abs(norm(1)/norm(2))>10) {
          lidar_point.SetType(LidarPointType::IcpGround);
          Eigen::Vector3ub color;
          color << 255,255,0;
          lidar_point.SetColor(color);
        } else {
          lidar_point.SetType(LidarPointType::Icp);
          Eigen::Vector3ub color;
          color << 0,0,255;
          lidar_point.SetColor(color);
        }

        double dist = lidar_point.ComputePointToPointDist(pt_xyz);
        if (dist > max_search_range) return;

        lidar_maps_.insert({point3D_id,lidar_point});
      }
}
// src/controllers/bundle_adjustment.cc:142-183 (whole-map refine)
static void LoopWholeMap(lidar::PointCloudProcess* lidar_pointcloud_process_, uint64_t point3d_id, Eigen::Vector3d pt_xyz,
                         std::map<uint64_t, LidarPoint>& lidar_maps_) {
      Eigen::Vector6d lidar_pt;
      if (lidar_pointcloud_process_->SearchNearestNeiborByKdtree(pt_xyz,lidar_pt)){
        Eigen::Vector3d norm = lidar_pt.block(3,0,3,1);
        Eigen::Vector3d l_pt = lidar_pt.block(0,0,3,1);

        double d = 0 - l_pt.dot(norm);
        Eigen::Vector4d plane;
        plane << norm(0),norm(1),norm(2),d;

        LidarPoint lidar_point(l_pt,plane);

        const double dist2plane = lidar_point.ComputeDist(pt_xyz);
        const double dist2point = lidar_point.ComputePointToPointDist(pt_xyz);

        if (dist2plane > 1 || dist2point > 2) return;

        if (std::abs(norm(1)/norm(0))>10 && std::abs(norm(1)/norm(2))>10) {
          lidar_point.SetType(LidarPointType::IcpGround);
          Eigen::Vector3ub color;
          color << 255,255,0;
          lidar_point.SetColor(color);
        } else {
          lidar_point.SetType(LidarPointType::Icp);
          Eigen::Vector3ub color;
          color << 0,0,255;
          lidar_point.SetColor(color);
        }

        lidar_maps_.insert({point3d_id,lidar_point});
      }
}

// serial loop vs batched MatchClosestLidarPoints on the same points: same ids, bit-identical records
static bool SameRecords(std::map<uint64_t, LidarPoint>& a, std::map<uint64_t, LidarPoint>& b, bool with_dist_angle, const char* what) {
  if (a.size() != b.size()) { fprintf(stderr, "%s: %zu serial vs %zu batched matches\n", what, a.size(), b.size()); return false; }
  for (auto& kv : a) {
    auto it = b.find(kv.first);
    if (it == b.end()) { fprintf(stderr, "%s: id %llu missing\n", what, (unsigned long long)kv.first); return false; }
    LidarPoint &x = kv.second, &y = it->second;
    bool ok = x.Type() == y.Type();
    for (int c = 0; c < 3; c++) ok = ok && x.LidarXYZ()(c) == y.LidarXYZ()(c) && x.Color()(c) == y.Color()(c);
    for (int c = 0; c < 4; c++) ok = ok && (x.LidarABCD()(c) == y.LidarABCD()(c) || (x.LidarABCD()(c) != x.LidarABCD()(c) && y.LidarABCD()(c) != y.LidarABCD()(c)));
    if (with_dist_angle) ok = ok && x.Dist() == y.Dist() && (x.Angle() == y.Angle() || (x.Angle() != x.Angle() && y.Angle() != y.Angle()));
    if (!ok) { fprintf(stderr, "%s: record of id %llu differs\n", what, (unsigned long long)kv.first); return false; }
  }
  return true;
}

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: dropin_check cloud.ply queries.f64 out.f64 [image.f64]\n"); return 2; }
  lidar::PcdProjectionOptions opt;
  opt.min_lidar_proj_dist = 0.5;
  // error behaviour: a missing file makes Initialize return false (ply.cc:14-17), nothing aborts
  { lidar::PointCloudProcess bad("/nonexistent/cloud.ply"); if (bad.Initialize(opt)) { fprintf(stderr, "missing file accepted\n"); return 1; } }
  lidar::PointCloudProcess pcp(argv[1]);
  if (!pcp.Initialize(opt)) { fprintf(stderr, "Initialize failed: %s\n", lidar::DropinRuntime::Get().error.c_str()); return 1; }
  std::vector<double> qin = read_f64(argv[2]);
  const int64_t n = (int64_t)qin.size() / 4;
  std::vector<double> out;
  out.push_back((double)pcp.GetCloud()->size());
  // (1) the serial call of the reference: one query per call
  const int64_t n_serial = n < 64 ? n : 64;
  for (int64_t i = 0; i < n_serial; i++) {
    Eigen::Vector3d p{qin[4 * i], qin[4 * i + 1], qin[4 * i + 2]};
    Eigen::Vector6d l;
    bool ok = pcp.SearchNearestNeiborByKdtree(p, l);
    out.push_back(ok ? 1.0 : 0.0);
    for (int k = 0; k < 6; k++) out.push_back(ok ? l(k) : 0.0);
  }
  // Kdtree::GetClosestPoint on the same queries
  for (int64_t i = 0; i < n_serial; i++) {
    lidar::LidarPoint a, b; a.x = (float)qin[4 * i]; a.y = (float)qin[4 * i + 1]; a.z = (float)qin[4 * i + 2];
    bool ok = pcp.kdtree_ptr_->GetClosestPoint(a, b);
    out.push_back(ok ? 1.0 : 0.0); out.push_back(b.x); out.push_back(b.y); out.push_back(b.z);
  }
  // (2) the hoisted loop: one batched association (global variant) + cost functions + callback
  std::vector<uint64_t> ids(n); std::vector<double> xyz(3 * n); std::vector<int32_t> on(n);
  for (int64_t i = 0; i < n; i++) {
    ids[i] = 1000 + (uint64_t)i; on[i] = (int32_t)qin[4 * i + 3];
    for (int k = 0; k < 3; k++) xyz[3 * i + k] = qin[4 * i + k];
  }
  std::map<uint64_t, LidarPoint> matches;
  if (!MatchClosestLidarPoints(pcp, ids, xyz, on, 1.5, 0.2, 0.1, PCD_ASSOC_GLOBAL, &matches)) { fprintf(stderr, "batch association failed\n"); return 1; }
  out.push_back((double)matches.size());
  {
    // (2b) the reference's own serial loop bodies over the shim == the batched calls, for all three variants
    const int64_t m = n < 400 ? n : 400;
    std::vector<uint64_t> ids_m(ids.begin(), ids.begin() + m); std::vector<double> xyz_m(xyz.begin(), xyz.begin() + 3 * m);
    std::vector<int32_t> on_m(on.begin(), on.begin() + m);
    std::map<uint64_t, LidarPoint> sl, sg, sw, bl, bg, bw;
    for (int64_t i = 0; i < m; i++) {
      Eigen::Vector3d p{xyz_m[3 * i], xyz_m[3 * i + 1], xyz_m[3 * i + 2]};
      double range = 1.5 - on_m[i] * 0.1;                                  // src/sfm/incremental_mapper.cc:1159-1163
      if (range <= 0.2) range = 0.2;
      LoopLocal(&pcp, ids_m[i], p, range, sl);
      LoopGlobal(&pcp, ids_m[i], p, range, sg);
      LoopWholeMap(&pcp, ids_m[i], p, sw);
    }
    if (!MatchClosestLidarPoints(pcp, ids_m, xyz_m, on_m, 1.5, 0.2, 0.1, PCD_ASSOC_LOCAL, &bl) ||
        !MatchClosestLidarPoints(pcp, ids_m, xyz_m, on_m, 1.5, 0.2, 0.1, PCD_ASSOC_GLOBAL, &bg) ||
        !MatchClosestLidarPoints(pcp, ids_m, xyz_m, on_m, 1.5, 0.2, 0.1, PCD_ASSOC_WHOLE_MAP, &bw)) { fprintf(stderr, "batched variants failed\n"); return 1; }
    if (!SameRecords(sl, bl, true, "local loop") || !SameRecords(sg, bg, false, "global loop") || !SameRecords(sw, bw, false, "whole-map loop")) return 1;
    if (sl.size() < 100) { fprintf(stderr, "too few matches to mean anything (%zu)\n", sl.size()); return 1; }
    // Reconstruction::FilterLidarOutlier's test (src/base/reconstruction.cc:783-800) on the stored records, moved points
    size_t num_filtered = 0, expect = 0;
    for (auto& kv : sg) {
      LidarPoint l_pt = kv.second;
      const int64_t i = (int64_t)(kv.first - 1000);
      Eigen::Vector3d i_pt_xyz{xyz_m[3 * i] + 0.05 * (double)(i % 7), xyz_m[3 * i + 1], xyz_m[3 * i + 2]};
      double dist = l_pt.ComputePointToPointDist(i_pt_xyz);
      if (l_pt.Type() == LidarPointType::Proj){
        if (dist > 0.2) num_filtered += 1;
      } else {
        if (dist > 0.25) num_filtered += 1;
      }
      double lx[3] = {l_pt.LidarXYZ()(0), l_pt.LidarXYZ()(1), l_pt.LidarXYZ()(2)}; uint8_t ty = PCD_TYPE_ICP, er = 0; int64_t cnt = 0;
      if (pcd_filter_outliers_batch(lidar::DropinRuntime::Get().ctx, i_pt_xyz.data(), lx, &ty, 1, 0.2, 0.25, &er, &cnt) != PCD_OK) return 1;
      expect += er;
    }
    if (num_filtered != expect || num_filtered == 0) { fprintf(stderr, "FilterLidarOutlier: serial %zu vs batched %zu\n", num_filtered, expect); return 1; }
  }
  std::shared_ptr<LidarEvaluationBatch> batch = LidarEvaluationBatch::Begin();     // one batch per ceres::Problem
  std::vector<ceres::CostFunction*> fns; std::vector<const double*> blocks;
  for (auto& kv : matches) {
    LidarPoint& lp = kv.second;
    const double w = lp.Type() == LidarPointType::IcpGround ? 10000.0 : (lp.Type() == LidarPointType::Icp ? 1000.0 : 10.0);
    ceres::CostFunction* f = BundleAdjustmentLidarCostFunction::Create(lp.LidarABCD(), w);     // bundle_adjustment.cc:1032
    const double* block = &xyz[3 * (kv.first - 1000)];
    batch->SetPointPtr(static_cast<BundleAdjustmentLidarCostFunction*>(f)->slot(), block);     // the pointer given to AddResidualBlock (:1037)
    fns.push_back(f); blocks.push_back(block);
  }
  {
    // a batch with an unbound slot refuses to prepare instead of reading through a stale pointer
    std::shared_ptr<LidarEvaluationBatch> other = LidarEvaluationBatch::Begin();
    Eigen::Vector4d pl{0.0, 0.0, 1.0, -1.0};
    ceres::CostFunction* g = BundleAdjustmentLidarCostFunction::Create(pl, 10.0);
    if (other->Prepare(true)) { fprintf(stderr, "unbound slot was prepared\n"); return 1; }
    delete g;
  }
  LidarEvaluationCallback cb(batch);
  cb.PrepareForEvaluation(true, true);
  size_t k = 0;
  for (auto& kv : matches) {
    // Ceres hands Evaluate its own state vector, never the block's storage: a COPY of the values must be served from
    // the prepared batch and must not rebind anything
    double r = 0, J[3] = {0, 0, 0}; double* jac[1] = {J};
    const double state[3] = {blocks[k][0], blocks[k][1], blocks[k][2]}; const double* par[1] = {state};
    if (!fns[k]->Evaluate(par, &r, jac)) { fprintf(stderr, "Evaluate failed\n"); return 1; }
    LidarPoint& lp = kv.second;
    out.push_back((double)kv.first); out.push_back((double)(int)lp.Type());
    for (int c = 0; c < 3; c++) out.push_back(lp.LidarXYZ()(c));
    for (int c = 0; c < 4; c++) out.push_back(lp.LidarABCD()(c));
    out.push_back(lp.Dist()); out.push_back(r);
    for (int c = 0; c < 3; c++) out.push_back(J[c]);
    k++;
  }
  // an unprepared evaluation (moved point) must still come from the GPU and agree in form
  if (!fns.empty()) {
    double moved[3] = {blocks[0][0] + 0.5, blocks[0][1], blocks[0][2]}; const double* par[1] = {moved};
    double r = -1, J[3]; double* jac[1] = {J};
    if (!fns[0]->Evaluate(par, &r, jac)) { fprintf(stderr, "unprepared Evaluate failed\n"); return 1; }
    out.push_back(r);
    // ... and the probe left the batch bound to the blocks: the next iteration still prepares from the user's storage
    xyz[3 * (matches.begin()->first - 1000)] += 0.25;
    cb.PrepareForEvaluation(true, true);
    const double* par2[1] = {blocks[0]};
    double r2 = -1;
    if (!fns[0]->Evaluate(par2, &r2, jac) || !batch->Matches(0, blocks[0])) { fprintf(stderr, "batch lost its binding\n"); return 1; }
    out.push_back(r2);
    xyz[3 * (matches.begin()->first - 1000)] -= 0.25;
  }
  for (auto* f : fns) delete f;
  // (3) depth projection through PcdProj::SetNewImage (map overload), optional
  if (argc > 4) {
    std::vector<double> im = read_f64(argv[4]);       // qvec[4] tvec[3] cam[8] w h nfeat then xy pairs
    MockImage image; MockCamera cam;
    for (int i = 0; i < 4; i++) image.q(i) = im[i];
    for (int i = 0; i < 3; i++) image.t(i) = im[4 + i];
    cam.params.assign(im.begin() + 7, im.begin() + 15); cam.w = (size_t)im[15]; cam.h = (size_t)im[16];
    const int64_t nf = (int64_t)im[17];
    for (int64_t i = 0; i < nf; i++) image.pts.push_back({Eigen::Vector2d{im[18 + 2 * i], im[19 + 2 * i]}, (uint64_t)i, true});
    std::map<uint64_t, Eigen::Vector6d> m;
    pcp.pcd_proj_->SetNewImage(image, cam, m);
    out.push_back((double)m.size());
    for (auto& kv : m) { out.push_back((double)kv.first); for (int c = 0; c < 6; c++) out.push_back(kv.second(c)); }
    // the batched Project2Image form must fill the cache with exactly the same maps
    MockImage image2 = image; image2.t(0) += 1.0;
    std::map<uint64_t, std::map<uint64_t, Eigen::Vector6d>> cache;
    pcp.pcd_proj_->SetNewImages<MockImage, MockCamera>({&image, &image2}, {&cam, &cam}, {7, 9}, cache);
    std::map<uint64_t, Eigen::Vector6d> m2;
    pcp.pcd_proj_->SetNewImage(image2, cam, m2);
    auto same = [](const std::map<uint64_t, Eigen::Vector6d>& a, const std::map<uint64_t, Eigen::Vector6d>& b) {
      if (a.size() != b.size()) return false;
      for (auto& kv : a) { auto it = b.find(kv.first); if (it == b.end()) return false; for (int c = 0; c < 6; c++) if (it->second(c) != kv.second(c)) return false; }
      return true;
    };
    if (cache.size() != 2 || !same(cache[7], m) || !same(cache[9], m2)) { fprintf(stderr, "batched SetNewImages differs from SetNewImage\n"); return 1; }
    // (5) Project2Image over a mock reconstruction: image 9 (the shifted pose) is reached through a pair above the
    // threshold, image 11 through a pair at the threshold (skipped), image 7 is cached already
    {
      struct El { uint32_t image_id; };
      struct TrackT { std::vector<El> e; const std::vector<El>& Elements() const { return e; } };
      struct P3 { TrackT t; const TrackT& Track() const { return t; } };
      struct Pair { size_t num_total_corrs; };
      struct Recon {
        std::map<uint64_t, P3> pts; MockCamera cam; std::map<uint32_t, size_t> pairs;
        P3& Point3D(uint64_t id) { return pts[id]; }
        bool ExistsImagePair(uint32_t, uint32_t b) { return pairs.count(b) != 0; }
        Pair ImagePair(uint32_t, uint32_t b) { return Pair{pairs[b]}; }
        struct Img : MockImage { uint32_t CameraId() const { return 0; } };
        std::map<uint32_t, Img> ims;
        Img& Image(uint32_t id) { return ims[id]; }
        MockCamera& Camera(uint32_t) { return cam; }
      } rec;
      rec.cam = cam;
      static_cast<MockImage&>(rec.ims[7]) = image; static_cast<MockImage&>(rec.ims[9]) = image2; static_cast<MockImage&>(rec.ims[11]) = image2;
      rec.pairs[9] = 31; rec.pairs[11] = 30;
      rec.pts[1].t.e = {{7}, {9}, {11}}; rec.pts[2].t.e = {{9}, {7}};
      std::map<uint64_t, std::map<uint64_t, Eigen::Vector6d>> cache2;
      cache2[7] = m;
      std::vector<uint64_t> pids = {1, 2};
      const size_t np = Project2Image(pcp, &rec, pids, 7u, 30, cache2);
      if (np != 1 || cache2.size() != 2 || !same(cache2[9], m2)) { fprintf(stderr, "Project2Image selected %zu images\n", np); return 1; }
    }
    // (6) MatchVariablePoint2LidarPoint (src/optim/bundle_adjustment.cc:288-350): the reference body, transcribed, per
    // point over the shim's LidarPoint -- against the batched MatchVariablePoints2LidarPoints
    {
      struct El { uint32_t image_id; };
      struct TrackT { std::vector<El> e; const std::vector<El>& Elements() const { return e; } };
      struct P3 { TrackT t; Eigen::Vector3d xyz; const TrackT& Track() const { return t; } Eigen::Vector3d XYZ() const { return xyz; } };
      struct Recon { std::map<uint64_t, P3> pts; P3& Point3D(uint64_t id) { return pts[id]; } } rec;
      std::map<uint64_t, std::map<uint64_t, Eigen::Vector6d>> lidar_searched_image_ids_;
      lidar_searched_image_ids_[7] = m; lidar_searched_image_ids_[9] = m2;
      std::vector<uint64_t> pids;
      for (auto& kv : m) {                     // feature ids seen in image 7; most are in image 9 as well
        P3 p3; p3.t.e = {{9}, {5}, {7}};
        for (int c = 0; c < 3; c++) p3.xyz(c) = kv.second(c) + 0.03 * (double)((kv.first + (uint64_t)c) % 5) - 0.05;
        rec.pts[kv.first] = p3; pids.push_back(kv.first);
        if (pids.size() >= 300) break;
      }
      std::map<uint64_t, LidarPoint> serial, batched;
      for (uint64_t point3D_id : pids) {
        P3& point3D = rec.Point3D(point3D_id);
        Eigen::Vector3d pt_xyz = point3D.XYZ();
        double angle = 360;
        Eigen::Vector6d lidar_pt;
        const auto& track_els = point3D.Track().Elements();
        for (auto& track_el : track_els){
          auto iter = lidar_searched_image_ids_.find(track_el.image_id);
          if (iter == lidar_searched_image_ids_.end()) continue;
          auto lidar_pt_ptr = iter->second.find(point3D_id);
          if (lidar_pt_ptr != iter->second.end()){
            Eigen::Vector6d lidar_pt_temp = lidar_pt_ptr->second;
            Eigen::Vector3d norm = lidar_pt_temp.block(3,0,3,1);
            Eigen::Vector3d vec = pt_xyz - lidar_pt_temp.block(0,0,3,1);
            double angle_temp = std::abs(vec.dot(norm)/(norm.norm()*vec.norm()));
            if (angle_temp < angle){
              angle = angle_temp;
              lidar_pt = lidar_pt_temp;
            }
          }
        }
        if (angle != 360){
          Eigen::Vector3d norm = lidar_pt.block(3,0,3,1);
          Eigen::Vector3d l_pt = lidar_pt.block(0,0,3,1);
          double d = 0 - l_pt.dot(norm);
          Eigen::Vector4d plane;
          plane << norm(0),norm(1),norm(2),d;

          LidarPoint lidar_point(LidarPointType::Proj,l_pt,plane);
          lidar_point.SetDist( lidar_point.ComputeDist(pt_xyz));
          lidar_point.SetAngle( lidar_point.ComputeAngle(pt_xyz));
          Eigen::Vector3ub color;
          color << 255,0,0;
          lidar_point.SetColor(color);
          serial.insert({point3D_id,lidar_point});
        }
      }
      if (!MatchVariablePoints2LidarPoints(&rec, pids, lidar_searched_image_ids_, &batched)) { fprintf(stderr, "batched a17 failed\n"); return 1; }
      if (serial.size() < 50 || !SameRecords(serial, batched, true, "MatchVariablePoint2LidarPoint")) return 1;
      for (auto& kv : batched) if (kv.second.Type() != LidarPointType::Proj) { fprintf(stderr, "a17 type\n"); return 1; }
    }
    // (4) the initial-pair overload (src/lidar/pcd_projection.cc:102-220): one push_back per keypoint after what the vector held
    std::vector<std::pair<Eigen::Vector2d, bool>> pt_xys;
    for (int64_t i = 0; i < nf; i++) pt_xys.push_back({Eigen::Vector2d{im[18 + 2 * i], im[19 + 2 * i]}, false});
    std::vector<Eigen::Vector3d> pt_xyzs(1, Eigen::Vector3d{7.0, 7.0, 7.0});
    pcp.pcd_proj_->SetNewImage(image, cam, pt_xys, pt_xyzs);
    if (pt_xyzs.size() != (size_t)nf + 1 || pt_xyzs[0](0) != 7.0) { fprintf(stderr, "init SetNewImage must append one entry per keypoint\n"); return 1; }
    out.push_back((double)nf);
    for (int64_t i = 0; i < nf; i++) { out.push_back(pt_xys[i].second ? 1.0 : 0.0); for (int c = 0; c < 3; c++) out.push_back(pt_xyzs[i + 1](c)); }
  }
  FILE* f = fopen(argv[3], "wb");
  if (!f || fwrite(out.data(), 8, out.size(), f) != out.size()) return 1;
  fclose(f);
  printf("dropin_check ok: %lld points, %lld queries, %zu matches\n", (long long)pcp.GetCloud()->size(), (long long)n, matches.size());
  return 0;
}
