// ctx.cu -- context lifetime, error plumbing, pinned host memory.
#include "common.cuh"
#include <stdlib.h>

namespace pcd {
static thread_local std::string g_err;
void set_global_error(const std::string& s) { g_err = s; }
}  // namespace pcd

extern "C" {

void pcd_ctx_destroy(pcd_ctx* c);

int pcd_ctx_create(int device, pcd_ctx** out) {
  if (!out) PCD_FAIL((pcd_ctx*)nullptr, PCD_E_INVALID, "pcd_ctx_create: out is NULL");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    PCD_FAIL((pcd_ctx*)nullptr, PCD_E_CUDA, "no usable CUDA device (%s); this library has no CPU fallback",
             e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (device < 0 || device >= count)
    PCD_FAIL((pcd_ctx*)nullptr, PCD_E_INVALID, "device %d out of range (have %d)", device, count);
  cudaDeviceProp prop;
  PCD_CUDA((pcd_ctx*)nullptr, cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    PCD_FAIL((pcd_ctx*)nullptr, PCD_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device,
             prop.major, prop.minor);
  DeviceGuard g(device);
  pcd_ctx* c = new pcd_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  cudaError_t ce = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->copy_in, cudaStreamNonBlocking);
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->copy_out, cudaStreamNonBlocking);
  for (int i = 0; i < 8 && ce == cudaSuccess; i++) ce = cudaEventCreateWithFlags(&c->ev[i], cudaEventDisableTiming);
  if (ce != cudaSuccess) {
    pcd_ctx_destroy(c);                                      // whatever was created so far
    PCD_FAIL((pcd_ctx*)nullptr, PCD_E_CUDA, "creating the context's streams / events failed: %s", cudaGetErrorString(ce));
  }
  *out = c;
  return PCD_OK;
}

void pcd_ctx_destroy(pcd_ctx* c) {
  if (!c) return;
  DeviceGuard g(c->device);
  cudaDeviceSynchronize();
  pcd::DevBuf* bufs[] = {&c->keys, &c->keys_alt, &c->vals, &c->vals_alt, &c->cub_tmp, &c->misc, &c->partial,
                         &c->stage_in[0], &c->stage_in[1], &c->stage_out[0], &c->stage_out[1],
                         &c->stage_aux[0], &c->stage_aux[1], &c->img_buf, &c->feat_buf, &c->zbuf, &c->sel_buf, &c->far_q, &c->qrec, &c->eval_partial};
  for (auto* b : bufs) b->release();
  for (int i = 0; i < 2; i++) { c->pin_in[i].release(); c->pin_out[i].release(); }
  for (int i = 0; i < 8; i++) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
  for (auto& s : c->prof_pending) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  for (auto e : c->prof_free) cudaEventDestroy(e);
  if (c->stream) cudaStreamDestroy(c->stream);
  if (c->copy_in) cudaStreamDestroy(c->copy_in);
  if (c->copy_out) cudaStreamDestroy(c->copy_out);
  delete c;
}

const char* pcd_last_error(const pcd_ctx* c) {
  if (c) return c->err.c_str();
  return pcd::g_err.c_str();
}

int pcd_ctx_device(const pcd_ctx* c) { return c ? c->device : -1; }

int pcd_ctx_counters(const pcd_ctx* c, int64_t out[4]) {
  if (!c || !out) return PCD_E_INVALID;
  for (int i = 0; i < 3; i++) out[i] = c->counters[i];
  out[3] = c->launches;
  return PCD_OK;
}

int64_t pcd_ctx_launches(const pcd_ctx* c) { return c ? c->launches : 0; }

int pcd_ctx_profile(pcd_ctx* c, int enable) {
  if (!c) return PCD_E_INVALID;
  DeviceGuard g(c->device);
  for (auto& s : c->prof_pending) { cudaEventSynchronize(s.b); c->prof_free.push_back(s.a); c->prof_free.push_back(s.b); }
  c->prof_pending.clear();
  for (int i = 0; i < PCD_K_COUNT; i++) { c->prof_ms[i] = 0.0; c->prof_n[i] = 0; }
  c->prof = enable != 0;
  return PCD_OK;
}

int pcd_ctx_kernel_time(pcd_ctx* c, int which, double* total_ms, int64_t* launches) {
  if (!c || which < 0 || which >= PCD_K_COUNT) return PCD_E_INVALID;
  DeviceGuard g(c->device);
  for (auto& s : c->prof_pending) {
    PCD_CUDA(c, cudaEventSynchronize(s.b));
    float ms = 0.0f;
    PCD_CUDA(c, cudaEventElapsedTime(&ms, s.a, s.b));
    c->prof_ms[s.which] += ms; c->prof_n[s.which]++;
    c->prof_free.push_back(s.a); c->prof_free.push_back(s.b);
  }
  c->prof_pending.clear();
  if (total_ms) *total_ms = c->prof_ms[which];
  if (launches) *launches = c->prof_n[which];
  return PCD_OK;
}

int pcd_host_alloc(void** out, int64_t bytes) {
  if (!out || bytes < 0) return PCD_E_INVALID;
  cudaError_t e = cudaMallocHost(out, (size_t)(bytes > 0 ? bytes : 1));
  if (e != cudaSuccess) PCD_FAIL((pcd_ctx*)nullptr, PCD_E_CUDA, "cudaMallocHost(%lld) failed: %s", (long long)bytes, cudaGetErrorString(e));
  return PCD_OK;
}

void pcd_host_free(void* p) { if (p) cudaFreeHost(p); }

void pcd_free(void* p) { free(p); }

}  // extern "C"
