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
  LidarEvaluationBatch::Reset();
  std::vector<ceres::CostFunction*> fns; std::vector<const double*> blocks;
  for (auto& kv : matches) {
    LidarPoint& lp = kv.second;
    const double w = lp.Type() == LidarPointType::IcpGround ? 10000.0 : (lp.Type() == LidarPointType::Icp ? 1000.0 : 10.0);
    ceres::CostFunction* f = BundleAdjustmentLidarCostFunction::Create(lp.LidarABCD(), w);     // bundle_adjustment.cc:1032
    const double* block = &xyz[3 * (kv.first - 1000)];
    LidarEvaluationBatch::Current()->SetPointPtr(static_cast<BundleAdjustmentLidarCostFunction*>(f)->slot(), block);
    fns.push_back(f); blocks.push_back(block);
  }
  LidarEvaluationCallback cb;
  cb.PrepareForEvaluation(true, true);
  size_t k = 0;
  for (auto& kv : matches) {
    double r = 0, J[3] = {0, 0, 0}; double* jac[1] = {J}; const double* par[1] = {blocks[k]};
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
  }
  FILE* f = fopen(argv[3], "wb");
  if (!f || fwrite(out.data(), 8, out.size(), f) != out.size()) return 1;
  fclose(f);
  printf("dropin_check ok: %lld points, %lld queries, %zu matches\n", (long long)pcp.GetCloud()->size(), (long long)n, matches.size());
  return 0;
}
