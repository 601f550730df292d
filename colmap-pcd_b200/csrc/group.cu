// group.cu -- the path on several GPUs of one node, behind the C ABI.
//
// Partitioning is the natural one of the path: the cloud and both indices are REPLICATED (built once on the first
// device, copied to the others device-to-device over NVLink -- no re-indexing per rank), the queries of a batch are
// cut into contiguous shards, the images of a projection batch into contiguous runs.  Every shard is one ordinary
// single-device call (pcd_associate_evaluate_batch, pcd_project_images) issued by a host thread of its own, pinned
// to the CPUs next to its GPU, on that device's context: H2D, kernels and D2H of all devices run concurrently, each
// over its own PCIe link.  The consumer of the results is ONE host process (Ceres), so the "gather" of the residual
// and Jacobian blocks is the shards' D2H copies landing in the caller's arrays at their offsets -- no block crosses
// between GPUs; the only reduction, sum r^2 and the counters, is added on the host in rank order (deterministic).
#include "common.cuh"
#include <sched.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <thread>
#include <vector>

struct pcd_group {
  std::vector<pcd_ctx*> ctx;
  std::vector<int> dev;
  std::string err;
};

struct pcd_group_cloud {
  pcd_group* g = nullptr;
  std::vector<pcd_cloud*> c;
};

namespace pcd {

// CPUs local to a device (sysfs local_cpulist of its PCI function), parsed into a cpu_set; false when unknown.
static bool device_cpus(int device, cpu_set_t* set) {
  char bus[32] = {0};
  if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) return false;
  for (char* p = bus; *p; p++) if (*p >= 'A' && *p <= 'Z') *p = (char)(*p - 'A' + 'a');
  std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
  FILE* f = fopen(path.c_str(), "r");
  if (!f) return false;
  char line[4096] = {0};
  const bool got = fgets(line, sizeof line, f) != nullptr;
  fclose(f);
  if (!got) return false;
  CPU_ZERO(set);
  int n = 0;
  for (char* tok = strtok(line, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
    int a = 0, b = 0;
    if (sscanf(tok, "%d-%d", &a, &b) == 2) { for (int k = a; k <= b && k < CPU_SETSIZE; k++) { CPU_SET(k, set); n++; } }
    else if (sscanf(tok, "%d", &a) == 1 && a < CPU_SETSIZE) { CPU_SET(a, set); n++; }
  }
  return n > 0;
}

// Restricts the calling thread to the CPUs next to `device`, intersected with what the process is allowed.
static void bind_thread_near(int device) {
  cpu_set_t near, allowed, both;
  if (!device_cpus(device, &near)) return;
  if (sched_getaffinity(0, sizeof allowed, &allowed) != 0) return;
  CPU_AND(&both, &near, &allowed);
  if (CPU_COUNT(&both) > 0) sched_setaffinity(0, sizeof both, &both);
}

template <typename T>
static cudaError_t copy_array(T** dst, const T* src, size_t count, int dst_dev, int src_dev, cudaStream_t st) {
  *dst = nullptr;
  cudaError_t e = cudaMalloc((void**)dst, sizeof(T) * (count > 0 ? count : 1));
  if (e != cudaSuccess || count == 0 || !src) return e;
  if (dst_dev == src_dev) return cudaMemcpyAsync(*dst, src, sizeof(T) * count, cudaMemcpyDeviceToDevice, st);
  return cudaMemcpyPeerAsync(*dst, dst_dev, src, src_dev, sizeof(T) * count, st);
}

// A replica of `src` (cloud + NN grid + sub-map) on the device of `dctx`.
static int replicate_cloud(const pcd_cloud* src, pcd_ctx* dctx, pcd_cloud** out) {
  DeviceGuard g(dctx->device);
  cudaStream_t st = dctx->stream;
  const int sd = src->ctx->device, dd = dctx->device;
  pcd_cloud* c = new pcd_cloud(*src);            // metadata (sizes, grid geometry, options); device pointers replaced below
  c->ctx = dctx;
  c->aos = nullptr; c->grid_pts = nullptr; c->cell_start = nullptr; c->occ = nullptr;
  c->sub_keys = nullptr; c->sub_start = nullptr; c->sub_pts = nullptr; c->sub_seg = nullptr;
  cudaError_t e = copy_array(&c->aos, src->aos, 8 * (size_t)src->n, dd, sd, st);
  if (e == cudaSuccess && src->has_index) {
    const size_t words = (size_t)src->grid.nbx * src->grid.nby * src->grid.nbz;
    e = copy_array(&c->grid_pts, src->grid_pts, (size_t)src->n, dd, sd, st);
    if (e == cudaSuccess) e = copy_array(&c->cell_start, src->cell_start, (size_t)(src->n_cells + 1), dd, sd, st);
    if (e == cudaSuccess) e = copy_array(&c->occ, src->occ, words > 0 ? words : 1, dd, sd, st);
    c->grid.pts = c->grid_pts; c->grid.cell_start = c->cell_start; c->grid.occ = c->occ;
  }
  if (e == cudaSuccess && src->has_submap) {
    e = copy_array(&c->sub_pts, src->sub_pts, (size_t)src->n, dd, sd, st);
    if (e == cudaSuccess) e = copy_array(&c->sub_keys, src->sub_keys, 3 * (size_t)src->sub_cells, dd, sd, st);
    if (e == cudaSuccess) e = copy_array(&c->sub_start, src->sub_start, (size_t)(src->sub_cells + 1), dd, sd, st);
    if (e == cudaSuccess && src->sub_seg) e = copy_array(&c->sub_seg, src->sub_seg, (size_t)src->sub_nseg, dd, sd, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) {
    pcd_cloud_destroy(c);
    PCD_FAIL(dctx, PCD_E_CUDA, "replicating the cloud to device %d failed: %s", dd, cudaGetErrorString(e));
  }
  *out = c;
  return PCD_OK;
}

static void shard_bounds(int64_t n, int parts, int r, int64_t* lo, int64_t* hi) {
  *lo = n * r / parts; *hi = n * (r + 1) / parts;
}

}  // namespace pcd

extern "C" {

int pcd_host_bind_near(int device) {
  cpu_set_t before, after;
  if (sched_getaffinity(0, sizeof before, &before) != 0) return 0;
  pcd::bind_thread_near(device);
  if (sched_getaffinity(0, sizeof after, &after) != 0) return 0;
  return CPU_EQUAL(&before, &after) ? -CPU_COUNT(&after) : CPU_COUNT(&after);
}

int pcd_group_create(const int* devices, int n, pcd_group** out) {
  if (!out || n <= 0 || n > 64) PCD_FAIL((pcd_ctx*)nullptr, PCD_E_INVALID, "pcd_group_create: bad argument");
  *out = nullptr;
  pcd_group* g = new pcd_group();
  for (int r = 0; r < n; r++) {
    const int d = devices ? devices[r] : r;
    pcd_ctx* c = nullptr;
    int rc = pcd_ctx_create(d, &c);
    if (rc) { for (pcd_ctx* x : g->ctx) pcd_ctx_destroy(x); delete g; return rc; }
    g->ctx.push_back(c); g->dev.push_back(d);
  }
  // peer access between every pair of distinct devices (NVLink / NVSwitch); not fatal when unavailable: the
  // replication then goes through host memory
  for (int a = 0; a < n; a++)
    for (int b = 0; b < n; b++) {
      if (g->dev[a] == g->dev[b]) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, g->dev[a], g->dev[b]) == cudaSuccess && can) {
        DeviceGuard dg(g->dev[a]);
        cudaError_t e = cudaDeviceEnablePeerAccess(g->dev[b], 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      }
    }
  *out = g;
  return PCD_OK;
}

void pcd_group_destroy(pcd_group* g) {
  if (!g) return;
  for (pcd_ctx* c : g->ctx) pcd_ctx_destroy(c);
  delete g;
}

int pcd_group_size(const pcd_group* g) { return g ? (int)g->ctx.size() : 0; }
pcd_ctx* pcd_group_ctx(pcd_group* g, int rank) { return (g && rank >= 0 && rank < (int)g->ctx.size()) ? g->ctx[(size_t)rank] : nullptr; }
const char* pcd_group_last_error(const pcd_group* g) { return g ? g->err.c_str() : ""; }

void pcd_group_cloud_destroy(pcd_group_cloud* gc) {
  if (!gc) return;
  for (pcd_cloud* c : gc->c) pcd_cloud_destroy(c);
  delete gc;
}

pcd_cloud* pcd_group_cloud_rank(pcd_group_cloud* gc, int rank) {
  return (gc && rank >= 0 && rank < (int)gc->c.size()) ? gc->c[(size_t)rank] : nullptr;
}

// Takes ownership of `first` (a cloud of the group's first context, index and / or sub-map already built) and
// replicates it to every other rank.
int pcd_group_replicate(pcd_group* g, pcd_cloud* first, pcd_group_cloud** out) {
  if (!g || !first || !out || first->ctx != g->ctx[0]) PCD_FAIL(g ? g->ctx[0] : (pcd_ctx*)nullptr, PCD_E_INVALID, "pcd_group_replicate: the cloud must belong to the group's first context");
  *out = nullptr;
  pcd_group_cloud* gc = new pcd_group_cloud();
  gc->g = g;
  gc->c.push_back(first);
  {
    DeviceGuard dg(first->ctx->device);
    cudaStreamSynchronize(first->ctx->stream);
  }
  for (size_t r = 1; r < g->ctx.size(); r++) {
    pcd_cloud* c = nullptr;
    int rc = pcd::replicate_cloud(first, g->ctx[r], &c);
    if (rc) {
      g->err = pcd_last_error(g->ctx[r]);
      gc->c.erase(gc->c.begin());                      // the caller keeps `first` on failure
      pcd_group_cloud_destroy(gc);
      return rc;
    }
    gc->c.push_back(c);
  }
  *out = gc;
  return PCD_OK;
}

int pcd_group_cloud_from_host(pcd_group* g, const float* aos8, int64_t n, int direction_trans, float cell_size,
                              const pcd_proj_options* opt, pcd_group_cloud** out) {
  if (!g || !out) return PCD_E_INVALID;
  pcd_cloud* first = nullptr;
  int rc = pcd_cloud_from_host(g->ctx[0], aos8, n, direction_trans, &first);
  if (!rc) rc = pcd_index_build(first, cell_size);
  if (!rc && opt) rc = pcd_submap_build(first, opt);
  if (!rc) rc = pcd_group_replicate(g, first, out);
  if (rc) { g->err = pcd_last_error(g->ctx[0]); if (first) pcd_cloud_destroy(first); }
  return rc;
}

int pcd_group_cloud_load_ply(pcd_group* g, const char* path, const pcd_proj_options* opt, pcd_group_cloud** out) {
  if (!g || !out || !path) return PCD_E_INVALID;
  pcd_cloud* first = nullptr;
  int rc = pcd_cloud_load_ply(g->ctx[0], path, opt, &first);
  if (!rc) rc = pcd_group_replicate(g, first, out);
  if (rc) { g->err = pcd_last_error(g->ctx[0]); if (first) pcd_cloud_destroy(first); }
  return rc;
}

// The fused pass of pcd_associate_evaluate_batch on every device of the group at once.  r == NULL: association only
// (pcd_associate_batch).  Outputs land in the caller's arrays at the shards' offsets.
int pcd_group_associate_evaluate_batch(pcd_group_cloud* gc, const double* p_xyz, const double* range, const int32_t* opt_num,
                                       const pcd_range_sweep* sweep, int64_t n, int variant, double w_proj, double w_icp,
                                       double w_ground, const pcd_assoc_out* out, double* r, double* J, double* sum_sq) {
  if (!gc || !out || n < 0 || (n > 0 && !p_xyz)) return PCD_E_INVALID;
  const int G = (int)gc->c.size();
  std::vector<int> rc((size_t)G, PCD_OK);
  std::vector<double> sums((size_t)G, 0.0);
  std::vector<std::thread> th;
  bool spawn_failed = false;
  for (int k = 0; k < G && !spawn_failed; k++) {
    try {
    th.emplace_back([&, k]() {
      int64_t lo, hi;
      pcd::shard_bounds(n, G, k, &lo, &hi);
      const int64_t m = hi - lo;
      if (m == 0) return;
      pcd::bind_thread_near(gc->c[(size_t)k]->ctx->device);
      pcd_assoc_out o = *out;
      if (o.idx) o.idx += lo;
      if (o.accepted) o.accepted += lo;
      if (o.type) o.type += lo;
      if (o.rgb) o.rgb += 3 * lo;
      if (o.xyz) o.xyz += 3 * lo;
      if (o.abcd) o.abcd += 4 * lo;
      if (o.dist) o.dist += lo;
      if (o.angle) o.angle += lo;
      if (r)
        rc[(size_t)k] = pcd_associate_evaluate_batch(gc->c[(size_t)k], p_xyz + 3 * lo, range ? range + lo : nullptr, opt_num ? opt_num + lo : nullptr,
                                                     sweep, m, variant, w_proj, w_icp, w_ground, &o, r + lo, J ? J + 3 * lo : nullptr,
                                                     sum_sq ? &sums[(size_t)k] : nullptr);
      else
        rc[(size_t)k] = pcd_associate_batch(gc->c[(size_t)k], p_xyz + 3 * lo, range ? range + lo : nullptr, opt_num ? opt_num + lo : nullptr,
                                            sweep, m, variant, &o);
    });
    } catch (...) { spawn_failed = true; }             // no exception crosses the C ABI
  }
  for (auto& t : th) t.join();
  if (spawn_failed) { gc->g->err = "could not start a host thread per device"; return PCD_E_NOMEM; }
  for (int k = 0; k < G; k++)
    if (rc[(size_t)k]) { gc->g->err = pcd_last_error(gc->c[(size_t)k]->ctx); return rc[(size_t)k]; }
  if (sum_sq) { double t = 0.0; for (double s : sums) t += s; *sum_sq = t; }      // rank order: deterministic
  return PCD_OK;
}

// pcd_project_images with the images cut into contiguous runs, one per device.
int pcd_group_project_images(pcd_group_cloud* gc, const pcd_image* imgs, int64_t n_img, const double coef[4], const double* feat_xy,
                             const int64_t* feat_start, int32_t* winner_idx, float* winner_range, int64_t* stats) {
  if (!gc || n_img < 0 || (n_img > 0 && (!imgs || !feat_start)) || !coef) return PCD_E_INVALID;
  const int G = (int)gc->c.size();
  std::vector<int> rc((size_t)G, PCD_OK);
  std::vector<std::vector<int64_t>> st((size_t)G, std::vector<int64_t>(4, 0));
  std::vector<std::thread> th;
  bool spawn_failed = false;
  for (int k = 0; k < G && !spawn_failed; k++) {
    try {
    th.emplace_back([&, k]() {
      int64_t lo, hi;
      pcd::shard_bounds(n_img, G, k, &lo, &hi);
      if (hi == lo) return;
      pcd::bind_thread_near(gc->c[(size_t)k]->ctx->device);
      std::vector<int64_t> fs((size_t)(hi - lo + 1));
      const int64_t f0 = feat_start[lo];
      for (int64_t i = lo; i <= hi; i++) fs[(size_t)(i - lo)] = feat_start[i] - f0;
      rc[(size_t)k] = pcd_project_images(gc->c[(size_t)k], imgs + lo, hi - lo, coef, feat_xy ? feat_xy + 2 * f0 : nullptr, fs.data(),
                                         winner_idx ? winner_idx + f0 : nullptr, winner_range ? winner_range + f0 : nullptr, st[(size_t)k].data());
    });
    } catch (...) { spawn_failed = true; }
  }
  for (auto& t : th) t.join();
  if (spawn_failed) { gc->g->err = "could not start a host thread per device"; return PCD_E_NOMEM; }
  for (int k = 0; k < G; k++)
    if (rc[(size_t)k]) { gc->g->err = pcd_last_error(gc->c[(size_t)k]->ctx); return rc[(size_t)k]; }
  if (stats) for (int i = 0; i < 4; i++) { stats[i] = 0; for (int k = 0; k < G; k++) stats[i] += st[(size_t)k][(size_t)i]; }
  return PCD_OK;
}

}  // extern "C"
