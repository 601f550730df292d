/*
 * pcd_b200.h -- C ABI of the B200-native image-to-point-cloud association path.
 *
 * Drop-in boundary for Colmap-PCD's src/lidar (Wangshihu12/colmap-pcd @ 0eecc21).  Each entry
 * point names the reference interface it replaces.  Plain pointers and sizes only; every
 * function returns 0 on success or a PCD_E_* code, and pcd_last_error() gives the message.
 * No exception or abort crosses this boundary.  There is NO CPU fallback: every compute entry
 * point fails with PCD_E_CUDA when no sm_100 device/driver is usable.
 *
 * Cloud layout: the reference's 32-byte AoS point (src/lidar/pt_type.h:14-20),
 *   float x, y, z, pad, normal_x, normal_y, normal_z, pad.
 * Index space: row numbers of the cloud AFTER PointCloudDirectionTrans' NaN-row drop
 * (src/lidar/ply.cc:33-57).
 *
 * "_dev" variants take device pointers and a cudaStream_t (as void*), launch asynchronously
 * and never synchronise; the plain variants take HOST pointers, stage through pinned
 * buffers and return when the results are in the caller's memory.
 */
#ifndef PCD_B200_H_
#define PCD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PCD_API __attribute__((visibility("default")))
#else
#define PCD_API
#endif

enum {
  PCD_OK = 0,
  PCD_E_INVALID = 1,   /* bad argument / missing prerequisite (e.g. index not built)   */
  PCD_E_CUDA = 2,      /* CUDA runtime / driver error (sticky per context)              */
  PCD_E_IO = 3,        /* file could not be read or parsed                              */
  PCD_E_NOMEM = 4
};

/* LidarPointType, src/lidar/lidar_point.h:9 */
enum { PCD_TYPE_PROJ = 0, PCD_TYPE_ICP = 1, PCD_TYPE_ICP_GROUND = 2, PCD_TYPE_NONE = 255 };

/* Which of the three twin association loops is being served (they differ in accept test and colour). */
enum {
  PCD_ASSOC_LOCAL = 0,     /* BundleAdjustmentConfig::MatchClosestLidarPoint  src/optim/bundle_adjustment.cc:358-410 */
  PCD_ASSOC_GLOBAL = 1,    /* AdjustGlobalBundleByLidar body                  src/sfm/incremental_mapper.cc:1430-1468 */
  PCD_ASSOC_WHOLE_MAP = 2  /* BundleAdjustmentController::Run body            src/controllers/bundle_adjustment.cc:141-183 */
};

typedef struct pcd_ctx pcd_ctx;       /* one CUDA device + scratch memory + sticky error */
typedef struct pcd_cloud pcd_cloud;   /* device-resident cloud, NN grid index, 1 m sub-map */

/* PcdProjectionOptions, src/lidar/pcd_projection.h:31-47 (numeric fields, same order/meaning). */
typedef struct pcd_proj_options {
  double depth_image_scale;     /* 0.2  */
  int32_t max_proj_scale;       /* 10   */
  int32_t min_proj_scale;       /* 2    */
  double min_proj_dist;         /* 2    */
  float submap_length;          /* 1.0  */
  float submap_width;           /* 1.0  */
  float submap_height;          /* 1.0  */
  float choose_meter;           /* 40.0 */
  double min_lidar_proj_dist;   /* 0.5 (src/controllers/incremental_mapper.h:92) */
} pcd_proj_options;

/* ---------------------------------------------------------------- context ---------------- */
PCD_API int pcd_ctx_create(int device, pcd_ctx** out);
PCD_API void pcd_ctx_destroy(pcd_ctx* ctx);
PCD_API const char* pcd_last_error(const pcd_ctx* ctx);   /* ctx may be NULL: last global error */
PCD_API int pcd_ctx_device(const pcd_ctx* ctx);
/* Counters since creation: [0] queries answered, [1] accepted, [2] ground, [3] kernels launched. */
PCD_API int pcd_ctx_counters(const pcd_ctx* ctx, int64_t out[4]);
PCD_API int64_t pcd_ctx_launches(const pcd_ctx* ctx);      /* number of kernels this ctx launched */
/* Per-kernel device timing for bench.py's roofline: when enabled, every launch of the kernel
 * classes below is bracketed by CUDA events on its launching stream.  pcd_ctx_kernel_time
 * synchronises the pending events and returns the accumulated milliseconds and launch count
 * since the last pcd_ctx_profile(ctx, 1). */
enum {
  PCD_K_NN_KEYSORT = 0,   /* query cell keys + radix sort                       */
  PCD_K_NN_SEARCH = 1,    /* k_nn_search (the dominant kernel)                  */
  PCD_K_ASSOC_RECORD = 2, /* k_assoc_record                                     */
  PCD_K_RESJAC = 3,       /* k_resjac (+ partial-sum reduction)                 */
  PCD_K_PROJECT = 4,      /* k_project (z-buffer splat)                         */
  PCD_K_NN_FAR = 5,       /* k_nn_far (far queries; timed apart from k_nn_search) */
  PCD_K_COUNT = 8
};
PCD_API int pcd_ctx_profile(pcd_ctx* ctx, int enable);
PCD_API int pcd_ctx_kernel_time(pcd_ctx* ctx, int which, double* total_ms, int64_t* launches);
/* Pinned host memory for callers that want truly asynchronous staging. */
PCD_API int pcd_host_alloc(void** out, int64_t bytes);
PCD_API void pcd_host_free(void* p);

/* ---------------------------------------------------------------- loader (a1, a2) -------- */
/* PLY parse on the host (ASCII / binary little-endian; x y z + normal_x|nx ...), raw lidar
 * frame, into a malloc'd AoS8 buffer released with pcd_free().  Replaces
 * pcl::io::loadPLYFile<LidarPoint> at src/lidar/ply.cc:14. */
PCD_API int pcd_ply_read(const char* path, float** aos8, int64_t* n);
PCD_API int pcd_ply_write(const char* path, const float* aos8, int64_t n, int binary);
PCD_API void pcd_free(void* p);

/* Pose-prior file "pose.ply" (src/controllers/incremental_mapper.cc:922-997, LoadPose): lidar-frame x y z roll pitch
 * yaw per image -> rows [tx ty tz qw qx qy qz] of the camera-frame pose, image_ids counting lines from 1 (lines with
 * "nan" are skipped).  *n = rows available; at most `capacity` rows are stored.  The writer emits the same format
 * (the reference's writer lives in GUI code, src/ui/main_window.cc:1096-1160). */
PCD_API int pcd_pose_ply_read(const char* path, int64_t* image_ids, double* tq7, int64_t capacity, int64_t* n);
PCD_API int pcd_pose_ply_write(const char* path, const double* xyzrpy, int64_t n);

/* Upload a cloud.  direction_trans != 0 applies PointCloudDirectionTrans on the device
 * ((x,y,z)<-(-y,-z,x), same for normals, NaN rows dropped, order kept: src/lidar/ply.cc:33-57).
 * Replaces the data half of PointCloudProcess::Initialize (src/lidar/ply.cc:9-31). */
PCD_API int pcd_cloud_from_host(pcd_ctx* ctx, const float* aos8, int64_t n, int direction_trans, pcd_cloud** out);
PCD_API int pcd_cloud_from_device(pcd_ctx* ctx, const float* d_aos8, int64_t n, int direction_trans, pcd_cloud** out);
/* pcd_ply_read + pcd_cloud_from_host(direction_trans = 1) + pcd_index_build + pcd_submap_build. */
PCD_API int pcd_cloud_load_ply(pcd_ctx* ctx, const char* path, const pcd_proj_options* opt, pcd_cloud** out);
/* N3, one-time load (the reference re-parses and re-indexes the map on every refine,
 * src/controllers/bundle_adjustment.cc:123-124): as pcd_cloud_load_ply, but the parsed, axis-swapped, NaN-compacted cloud
 * is kept in a binary sidecar path + ".pcdb200" keyed by the source file (pcd_file_key); *cache_hit = 1 when it was used.
 * A stale, truncated or foreign sidecar is ignored and rewritten.  The GPU indices are rebuilt either way. */
PCD_API int pcd_cloud_load_ply_cached(pcd_ctx* ctx, const char* path, const pcd_proj_options* opt, pcd_cloud** out, int* cache_hit);
PCD_API int pcd_file_key(const char* path, uint64_t* key);      /* FNV-1a of size, mtime, first and last 64 KiB */
PCD_API int pcd_sidecar_write(const char* path, uint64_t key, int direction_trans, const float* aos8, int64_t n);
PCD_API int pcd_sidecar_read(const char* path, uint64_t key, int direction_trans, float** aos8, int64_t* n);   /* pcd_free() */

/* a16, the decision half of BundleAdjustmentConfig::Project2Image (src/optim/bundle_adjustment.cc:241-281): for one
 * point's track, which elements' images go to SetNewImage now.  track_image[k]: the element's image; pair_corrs[k]:
 * num_total_corrs of the pair (image_id, track_image[k]), -1 when the pair does not exist; cached[k] != 0 (or cached ==
 * NULL: none): already in lidar_searched_image_ids_.  An element is skipped iff its image differs from image_id, the pair
 * exists and corrs <= match_features_threshold (:247-253); otherwise it is projected unless cached or already selected
 * by an earlier element of the same track.  Returns the number selected (>= 0) or a negative status. */
PCD_API int64_t pcd_project2image_select(const uint32_t* track_image, int64_t n, uint32_t image_id, const int64_t* pair_corrs,
                                         int match_features_threshold, const uint8_t* cached, uint8_t* project);
PCD_API void pcd_cloud_destroy(pcd_cloud* cloud);
PCD_API int64_t pcd_cloud_size(const pcd_cloud* cloud);
PCD_API int pcd_cloud_download(const pcd_cloud* cloud, float* aos8);   /* the post-swap cloud */
PCD_API const float* pcd_cloud_device_ptr(const pcd_cloud* cloud);     /* AoS8 on the device    */

/* ---------------------------------------------------------------- kd-tree wrapper (a3-a5) - */
/* Builds the GPU spatial index (uniform grid, power-of-two cell, cell-sorted float4 points).
 * cell_size <= 0 picks the cell from the cloud density.  Replaces Kdtree::BuildMap
 * (src/lidar/kdtree.cc:5-8). */
PCD_API int pcd_index_build(pcd_cloud* cloud, float cell_size);
PCD_API int pcd_index_info(const pcd_cloud* cloud, float* cell_size, int32_t dims[3], float origin[3], int64_t* n_cells);

/* Batched exact 1-NN: float32 L2 on the float-cast query, ties to the lowest index
 * (Kdtree::GetClosestPoint, src/lidar/kdtree.cc:10-21).  radius == NULL: unbounded, as the
 * reference.  radius != NULL: per-query bound; a query with no cloud point within
 * radius[i] reports idx = -1 (the reference finds a farther point and then rejects it at
 * src/optim/bundle_adjustment.cc:400, so the accepted set is identical).
 * idx: int32[n]; d2: float[n] squared float distance (inf when idx < 0); either may be NULL. */
PCD_API int pcd_nn_batch(pcd_cloud* cloud, const double* q_xyz, const double* radius, int64_t n,
                         int32_t* idx, float* d2);
PCD_API int pcd_nn_batch_dev(pcd_cloud* cloud, const double* d_q_xyz, const double* d_radius, int64_t n,
                             int32_t* d_idx, float* d_d2, void* stream);

/* PointCloudProcess::SearchNearestNeiborByKdtree (src/lidar/ply.cc:90-107) for a batch:
 * l_pt: double[6n] = [x y z nx ny nz]; ok: uint8[n] is the reference's bool return. */
PCD_API int pcd_search_nn_batch(pcd_cloud* cloud, const double* q_xyz, int64_t n, double* l_pt, uint8_t* ok);

/* ---------------------------------------------------------------- association (a6-a8) ---- */
/* range = max(min_r, max_r - opt_num*drop): src/sfm/incremental_mapper.cc:1159-1163,:1423-1427 */
typedef struct pcd_range_sweep { double max_r, min_r, drop; } pcd_range_sweep;

/* Output record arrays (SoA).  Any pointer may be NULL (not produced / not copied back). */
typedef struct pcd_assoc_out {
  int32_t* idx;        /* [n]   NN index, -1 when none                                    */
  uint8_t* accepted;   /* [n]   1 iff the reference would have stored the LidarPoint       */
  uint8_t* type;       /* [n]   PCD_TYPE_*                                                  */
  uint8_t* rgb;        /* [3n]  colour (src/optim/bundle_adjustment.cc:385-394)             */
  double* xyz;         /* [3n]  LidarXYZ()                                                  */
  double* abcd;        /* [4n]  LidarABCD() after Normalize (src/lidar/lidar_point.cc:39-50) */
  double* dist;        /* [n]   Dist()  = |l - p|                                           */
  double* angle;       /* [n]   Angle() (src/lidar/lidar_point.cc:32-37)                    */
} pcd_assoc_out;

/* Fused NN + record + classification + accept test for one of the three reference loops.
 * range: double[n] per-query search range, or NULL together with opt_num/sweep to derive it
 * from GlobalOptNum, or all NULL (whole-map variant uses its fixed thresholds). */
PCD_API int pcd_associate_batch(pcd_cloud* cloud, const double* p_xyz, const double* range,
                                const int32_t* opt_num, const pcd_range_sweep* sweep,
                                int64_t n, int variant, const pcd_assoc_out* out);
PCD_API int pcd_associate_batch_dev(pcd_cloud* cloud, const double* d_p_xyz, const double* d_range,
                                    const int32_t* d_opt_num, const pcd_range_sweep* sweep,
                                    int64_t n, int variant, const pcd_assoc_out* d_out, void* stream);

/* One pipelined pass for the end-to-end flow: pcd_associate_batch, then for every ACCEPTED match the weight by
 * type (src/optim/bundle_adjustment.cc:1013-1028) and the cost function with its 1x3 Jacobian evaluated at the
 * SfM point itself (src/base/cost_functions.h:204-232) -- what AddLidarToProblem + the first Ceres evaluation
 * produce.  The SfM points cross PCIe once and the records never come back to be re-uploaded.
 * r[n], J[3n] (may be NULL): 0 for queries without an accepted match; sum_sq (may be NULL) = sum r^2. */
PCD_API int pcd_associate_evaluate_batch(pcd_cloud* cloud, const double* p_xyz, const double* range,
                                         const int32_t* opt_num, const pcd_range_sweep* sweep, int64_t n, int variant,
                                         double w_proj, double w_icp, double w_ground, const pcd_assoc_out* out,
                                         double* r, double* J, double* sum_sq);

/* Device-pointer form of the fused pass (one record kernel writes the record fields, r, J and the partial sums). */
PCD_API int pcd_associate_evaluate_batch_dev(pcd_cloud* cloud, const double* d_p_xyz, const double* d_range,
                                             const int32_t* d_opt_num, const pcd_range_sweep* sweep, int64_t n, int variant,
                                             double w_proj, double w_icp, double w_ground, const pcd_assoc_out* d_out,
                                             double* d_r, double* d_J, double* d_sum_sq, void* stream);

/* The arithmetic of colmap::LidarPoint's own methods (src/lidar/lidar_point.h:12-13,37-41, lidar_point.cc:21-50) for n
 * records: normalize != 0 applies Normalize() to (xyz, abcd) first (what both constructors do) and abcd_out receives the
 * plane; with pt (3 doubles per record) given, dist_plane = ComputeDist(pt) = |pt . abc + d|, dist_point =
 * ComputePointToPointDist(pt) = |xyz - pt|, angle = ComputeAngle(pt) = |abc . (pt - xyz)| / |pt - xyz| on that plane.
 * Any output may be NULL.  The shim's colmap::LidarPoint calls this with n = 1 where a call site stays serial. */
PCD_API int pcd_lidar_point_ops(pcd_ctx* ctx, const double* xyz, const double* abcd, const double* pt, int64_t n, int normalize,
                                double* abcd_out, double* dist_plane, double* dist_point, double* angle);

/* MatchVariablePoint2LidarPoint (src/optim/bundle_adjustment.cc:288-350): per point, among its
 * candidates cand[cand_start[i] .. cand_start[i+1]) (6 doubles each, track order) choose min
 * |v.n|/(|n||v|); outputs a Proj record (dist = point-to-plane).  found: uint8[n]. */
PCD_API int pcd_match_proj_batch(pcd_ctx* ctx, const double* p_xyz, const double* cand, const int64_t* cand_start,
                                 int64_t n, uint8_t* found, double* xyz, double* abcd, double* dist, double* angle);

/* ---------------------------------------------------------------- cost function (a18-a19) - */
/* Batched BundleAdjustmentLidarCostFunction (src/base/cost_functions.h:204-232) with the
 * Jet<double,3> derivative: r[n], J[3n] row-major 1x3 blocks (NaN at u == 0 as autodiff gives),
 * sum_sq = sum r^2 (deterministic reduction).  J / sum_sq may be NULL. */
PCD_API int pcd_residual_jacobian_batch(pcd_ctx* ctx, const double* p_xyz, const double* abcd, const double* w,
                                        int64_t n, double* r, double* J, double* sum_sq);
PCD_API int pcd_residual_jacobian_batch_dev(pcd_ctx* ctx, const double* d_p_xyz, const double* d_abcd, const double* d_w,
                                            int64_t n, double* d_r, double* d_J, double* d_sum_sq, void* stream);
/* Allocation-free variant of pcd_residual_jacobian_batch_dev.  No counterpart in the reference: added
 * for callers that may not allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same arguments, same kernels, bit-identical r / J / sum_sq; the block partial sums
 * live in a fixed 512 KiB device array inside the library (held by every GPU that loads it until exit), so the call
 * makes no cudaMalloc / cudaFree and n is limited to pcd_residual_jacobian_noalloc_max() = 16 777 216 constraints per
 * call -- larger n returns PCD_E_INVALID with a message.  All calls on one device share that array (it keeps the last
 * call's contents; nothing in it needs clearing): calls that overlap on the GPU race, so issue them in one stream or
 * order them with events.  Not timed by pcd_ctx_profile. */
PCD_API int pcd_residual_jacobian_batch_dev_noalloc(pcd_ctx* ctx, const double* d_p_xyz, const double* d_abcd, const double* d_w,
                                                    int64_t n, double* d_r, double* d_J, double* d_sum_sq, void* stream);
PCD_API int64_t pcd_residual_jacobian_noalloc_max(void);
/* fp32 wire format of the residual blocks for the multi-GPU exchange (north_star: "fp32 paths stated separately";
 * no reference counterpart).  d_out[4n] = {r, J0, J1, J2} per constraint, each rounded to nearest float; d_out is
 * 16-byte aligned.  The fp64 blocks are unchanged; this is a transport format only. */
PCD_API int pcd_pack_blocks_f32_dev(pcd_ctx* ctx, const double* d_r, const double* d_J, int64_t n, float* d_out, void* stream);
/* north_star extension (no reference counterpart): the point expressed through
 * a camera, p_w = R(q)^T (p_c - t).  J_pose[7n] = d r / d (qw qx qy qz tx ty tz). */
PCD_API int pcd_residual_pose_jacobian_batch(pcd_ctx* ctx, const double* p_cam, const double* qvec, const double* tvec,
                                             const double* abcd, const double* w, int64_t n,
                                             double* r, double* J_point, double* J_pose);
/* Weight by type, src/optim/bundle_adjustment.cc:1013-1028. w_out[n]; type NONE -> 0. */
PCD_API int pcd_weights_from_type(pcd_ctx* ctx, const uint8_t* type, int64_t n, double w_proj, double w_icp,
                                  double w_ground, double* w_out);
PCD_API int pcd_weights_from_type_dev(pcd_ctx* ctx, const uint8_t* d_type, int64_t n, double w_proj, double w_icp,
                                      double w_ground, double* d_w_out, void* stream);
/* Reconstruction::FilterLidarOutlier distance test (src/base/reconstruction.cc:771-804). */
PCD_API int pcd_filter_outliers_batch(pcd_ctx* ctx, const double* p_xyz, const double* lidar_xyz, const uint8_t* type,
                                      int64_t n, double max_proj_dist, double max_icp_dist, uint8_t* erase, int64_t* n_erased);

/* ---------------------------------------------------------------- depth projection (a9-a15) */
/* PcdProj::BuildSubMap (src/lidar/pcd_projection.cc:223-255): cells keyed by round(coord/len). */
PCD_API int pcd_submap_build(pcd_cloud* cloud, const pcd_proj_options* opt);
PCD_API int64_t pcd_submap_cells(const pcd_cloud* cloud);

/* The function-local statics a_x b_x a_y b_y of src/lidar/pcd_projection.cc:391-397, computed
 * from the first camera projected in the process (quirk F7). */
PCD_API int pcd_splat_coef(const double cam[8], const pcd_proj_options* opt, double coef[4]);

/* Measurement aid for the z-buffer's roofline (north_star: "atomic throughput for the z-buffer"): the rate of
 * contention-free, warp-coalesced 64-bit atomicMin updates on a buffer of `words` words (one z-buffer), `sweeps`
 * passes, timed with CUDA events; best of three. */
PCD_API int pcd_atomic_peak(pcd_ctx* ctx, int64_t words, int sweeps, double* atomics_per_s);

typedef struct pcd_image {
  double qvec[4];      /* Image::Qvec()  (w x y z)                                   */
  double tvec[3];      /* Image::Tvec()                                              */
  double cam[8];       /* OPENCV params fx fy cx cy k1 k2 p1 p2 (Camera::Params())   */
  int32_t width, height;   /* Camera::Width()/Height() (unscaled)                    */
} pcd_image;

/* One image through SearchSubMap + ImageMapProj (src/lidar/pcd_projection.cc:258-468).
 * feat_xy: double[2*n_feat] keypoints (unscaled pixels) or NULL.
 * winner_idx[n_feat] (-1 = none) / winner_range[n_feat]: min-range lidar point at each
 * keypoint's pixel (ties -> lowest point index).  dense_zbuf: optional uint64[img_w*img_h]
 * (range bits << 32 | index, UINT64_MAX empty) = the if_save_depth_image map.
 * stats (optional) int64[4]: cells selected, points projected, pixel updates, features hit. */
PCD_API int pcd_project_image(pcd_cloud* cloud, const pcd_image* img, const double coef[4],
                              const double* feat_xy, int64_t n_feat,
                              int32_t* winner_idx, float* winner_range,
                              uint64_t* dense_zbuf, int64_t* stats);
/* Batch of images (each with its own feature list: feat_start[n_img+1]). */
PCD_API int pcd_project_images(pcd_cloud* cloud, const pcd_image* imgs, int64_t n_img, const double coef[4],
                               const double* feat_xy, const int64_t* feat_start,
                               int32_t* winner_idx, float* winner_range, int64_t* stats);
/* SetNewImage init overload's ray/plane step (src/lidar/pcd_projection.cc:183-205). */
PCD_API int pcd_ray_plane_batch(pcd_cloud* cloud, const pcd_image* img, const double* feat_xy, const int32_t* winner_idx,
                                int64_t n_feat, uint8_t* matched, double* xyz);

/* ---------------------------------------------------------------- several GPUs of one node */
/* A group is one context per device.  The cloud and both indices are built once on the group's first device and
 * replicated device-to-device (peer copies over NVLink); a batch is cut into contiguous query shards (association) or
 * image runs (projection), one per device, each an ordinary single-device call on a host thread bound to the CPUs next
 * to its GPU.  Results land in the caller's host arrays at the shards' offsets: the consumer is one host process
 * (Ceres, src/optim/bundle_adjustment.cc:452-530), so the gather of the residual / Jacobian blocks is the shards' own
 * D2H copies over each GPU's PCIe link, and the one reduction (sum r^2, counters) is added on the host in rank order.
 * devices == NULL: devices 0 .. n-1.  A device may be named more than once (several shards on one GPU). */
typedef struct pcd_group pcd_group;
typedef struct pcd_group_cloud pcd_group_cloud;
PCD_API int pcd_group_create(const int* devices, int n, pcd_group** out);
PCD_API void pcd_group_destroy(pcd_group* g);
PCD_API int pcd_group_size(const pcd_group* g);
PCD_API pcd_ctx* pcd_group_ctx(pcd_group* g, int rank);
PCD_API const char* pcd_group_last_error(const pcd_group* g);
/* cloud + NN index (+ sub-map when opt != NULL) on every device of the group */
PCD_API int pcd_group_cloud_from_host(pcd_group* g, const float* aos8, int64_t n, int direction_trans, float cell_size,
                                      const pcd_proj_options* opt, pcd_group_cloud** out);
PCD_API int pcd_group_cloud_load_ply(pcd_group* g, const char* path, const pcd_proj_options* opt, pcd_group_cloud** out);
/* replicates a cloud of the group's first context (ownership passes to the group cloud on success) */
PCD_API int pcd_group_replicate(pcd_group* g, pcd_cloud* first, pcd_group_cloud** out);
PCD_API pcd_cloud* pcd_group_cloud_rank(pcd_group_cloud* gc, int rank);
PCD_API void pcd_group_cloud_destroy(pcd_group_cloud* gc);
/* pcd_associate_evaluate_batch over the group (r == NULL: pcd_associate_batch); same arguments and results */
PCD_API int pcd_group_associate_evaluate_batch(pcd_group_cloud* gc, const double* p_xyz, const double* range,
                                               const int32_t* opt_num, const pcd_range_sweep* sweep, int64_t n, int variant,
                                               double w_proj, double w_icp, double w_ground, const pcd_assoc_out* out,
                                               double* r, double* J, double* sum_sq);
/* pcd_project_images over the group (contiguous runs of images per device) */
PCD_API int pcd_group_project_images(pcd_group_cloud* gc, const pcd_image* imgs, int64_t n_img, const double coef[4],
                                     const double* feat_xy, const int64_t* feat_start, int32_t* winner_idx,
                                     float* winner_range, int64_t* stats);
/* Binds the calling thread to the CPUs next to `device` (sysfs local_cpulist): pinned buffers allocated afterwards and
 * the copies issued from this thread stay on the GPU's NUMA node.  Returns the CPU count bound to (negative: unchanged). */
PCD_API int pcd_host_bind_near(int device);

#ifdef __cplusplus
}
#endif
#endif /* PCD_B200_H_ */
