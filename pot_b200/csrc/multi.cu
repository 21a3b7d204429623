// C ABI of the scaling iteration for several target histograms on one kernel matrix (see include/b2ot.h,
// multi_hist_kernels.cuh): the loop body of ot.bregman.sinkhorn_stabilized with a 2-D b.
#include <algorithm>
#include <cstdint>

#include "../../include/b2ot.h"
#include "host_util.h"
#include "multi_hist_kernels.cuh"

using namespace b2ot;

extern "C" size_t b2ot_multi_scratch_bytes(int64_t, int64_t m, int64_t H) {
  return (size_t)MH_SPLIT * (size_t)H * (size_t)m * sizeof(double);
}

template <typename T>
static int multi_step(const T* K, int64_t n, int64_t m, int64_t ld, int H, const T* a, const T* B, const T* Up,
                      const T* Vp, T* U, T* V, double* stats, double* scratch, cudaStream_t s) {
  const int S = (int)std::min<long long>(n, MH_SPLIT);
  const unsigned chunks = (unsigned)((H + MH_HC - 1) / MH_HC);
  const unsigned cblocks = (unsigned)((m + 255) / 256);
  mh_col_partial_kernel<T><<<dim3(cblocks, (unsigned)S, chunks), 256, 0, s>>>(K, n, m, ld, H, Up, scratch);
  B2OT_CHECK_LAUNCH();
  mh_col_finish_kernel<T><<<dim3(cblocks, (unsigned)H), 256, 0, s>>>(scratch, m, H, S, B, V);
  B2OT_CHECK_LAUNCH();
  mh_row_kernel<T><<<dim3((unsigned)((n + 7) / 8), 1, chunks), 256, 0, s>>>(K, n, m, ld, H, V, a, U);
  B2OT_CHECK_LAUNCH();
  if (stats != nullptr) {
    mh_stats_kernel<T><<<1, 1024, 0, s>>>(U, Up, (long long)n * H, V, Vp, (long long)m * H, stats);
    B2OT_CHECK_LAUNCH();
  }
  return 0;
}

extern "C" int b2ot_multi_scaling_step(int dtype, const void* K, int64_t n, int64_t m, int64_t ld, int64_t H,
                                       const void* a, const void* B, const void* u_prev, const void* v_prev,
                                       void* u_new, void* v_new, double* stats_dev, void* scratch_dev, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_multi_scaling_step: bad dtype %d", dtype);
  B2OT_REQUIRE(n >= 1 && m >= 1 && ld >= m && H >= 1 && H <= 65535,
               "b2ot_multi_scaling_step: bad shape n=%lld m=%lld ld=%lld H=%lld",
               (long long)n, (long long)m, (long long)ld, (long long)H);
  B2OT_REQUIRE(K && a && B && u_prev && u_new && v_new && scratch_dev, "b2ot_multi_scaling_step: null device pointer");
  B2OT_REQUIRE(stats_dev == nullptr || v_prev != nullptr, "b2ot_multi_scaling_step: the statistics need v_prev");
  B2OT_REQUIRE(u_new != u_prev && v_new != v_prev, "b2ot_multi_scaling_step: the update is out of place");
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B2OT_F32)
    return multi_step<float>((const float*)K, n, m, ld, (int)H, (const float*)a, (const float*)B,
                             (const float*)u_prev, (const float*)v_prev, (float*)u_new, (float*)v_new, stats_dev,
                             (double*)scratch_dev, s);
  return multi_step<double>((const double*)K, n, m, ld, (int)H, (const double*)a, (const double*)B,
                            (const double*)u_prev, (const double*)v_prev, (double*)u_new, (double*)v_new, stats_dev,
                            (double*)scratch_dev, s);
}

extern "C" int b2ot_multi_absorb(int dtype, int64_t n, int64_t m, int64_t H, const void* u, const void* v,
                                 double* alpha_r, double* beta_r, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_multi_absorb: bad dtype %d", dtype);
  B2OT_REQUIRE(n >= 1 && m >= 1 && H >= 1 && u && v && alpha_r && beta_r, "b2ot_multi_absorb: bad arguments");
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B2OT_F32)
    mh_absorb_kernel<float><<<1, 1024, 0, s>>>((const float*)u, n, (const float*)v, m, (int)H, alpha_r, beta_r);
  else
    mh_absorb_kernel<double><<<1, 1024, 0, s>>>((const double*)u, n, (const double*)v, m, (int)H, alpha_r, beta_r);
  B2OT_CHECK_LAUNCH();
  return 0;
}
