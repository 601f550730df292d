// nn.cuh -- internal interface of the batched grid search (shared by nn.cu and assoc.cu).
#pragma once
#include "common.cuh"

namespace pcd {

// Runs key -> sort -> search on `stream`.  d_q: double[3n] device; d_radius: double[n] or NULL;
// fixed_radius used when d_radius is NULL (INFINITY = unbounded).  When d_opt_num != NULL the
// radius is max(min_r, max_r - opt_num*drop) (src/sfm/incremental_mapper.cc:1159-1163).
// Writes idx (original order) and, if non-NULL, d2.  The kernels search the float ball of the radius rounded
// up; exact_radius = true adds the pass that applies the reference's fp64 test |l - p| <= R and reports -1
// beyond it (callers that apply the test themselves, like k_assoc_record, pass false).
int nn_search_dev(pcd_cloud* cloud, const double* d_q, const double* d_radius, double fixed_radius,
                  const int32_t* d_opt_num, const pcd_range_sweep* sweep, int64_t n, int32_t* d_idx, float* d_d2,
                  bool exact_radius, cudaStream_t stream);

__host__ __device__ inline double sweep_range(int opt_num, double max_r, double min_r, double drop) {
  double r = max_r - opt_num * drop;
  if (r <= min_r) r = min_r;
  return r;
}

}  // namespace pcd
