// C ABI of the scaling iteration for several target histograms on one kernel matrix (see include/b2ot.h,
// multi_hist_kernels.cuh): the loop body of ot.bregman.sinkhorn_stabilized with a 2-D b.
#include <algorithm>
#include <cstdint>

#include "../../include/b2ot.h"
#include "host_util.h"
#include "multi_hist_kernels.cuh"

using namespace b2ot;

static_assert(MH_STATS == B2OT_MULTI_STATS && MH_KNOPP == B2OT_MULTI_KNOPP && MH_UNBAL == B2OT_MULTI_UNBALANCED,
              "mode constants of include/b2ot.h and multi_hist_kernels.cuh");

extern "C" size_t b2ot_multi_scratch_bytes(int64_t, int64_t m, int64_t H) {
  const size_t cblocks = (size_t)((m + 255) / 256);
  return ((size_t)MH_SPLIT * (size_t)H * (size_t)m + 2 * cblocks * (size_t)H) * sizeof(double);
}

// the two passes for the histograms [h_base, h_base + HC z): exact HC, no guards in the kernels
template <typename T, int HC>
static int launch_col_partial(const T* K, int64_t n, int64_t m, int64_t ld, int H, int h_base, unsigned z, const T* U,
                              double* scratch, const MhCtl& ctl, cudaStream_t s) {
  const int S = (int)std::min<long long>(n, MH_SPLIT);
  mh_col_partial_kernel<T, HC><<<dim3((unsigned)((m + 255) / 256), (unsigned)S, z), 256, 0, s>>>(K, n, m, ld, H, h_base,
                                                                                                U, scratch, ctl);
  B2OT_CHECK_LAUNCH();
  return 0;
}
template <typename T, int HC>
static int launch_row(const T* K, int64_t n, int64_t m, int64_t ld, int H, int h_base, unsigned z, const T* V,
                      const T* a, T* U, const MhCtl& ctl, cudaStream_t s) {
  mh_row_kernel<T, HC><<<dim3((unsigned)((n + MH_ROWS - 1) / MH_ROWS), 1, z), MH_ROWS * 32, 0, s>>>(K, n, m, ld, H, h_base,
                                                                                                   V, a, U, ctl);
  B2OT_CHECK_LAUNCH();
  return 0;
}
#define B2OT_MH_REST(FN, rest, ...)                  \
  switch (rest) {                                    \
    case 1: rc = FN<T, 1>(__VA_ARGS__); break;       \
    case 2: rc = FN<T, 2>(__VA_ARGS__); break;       \
    case 3: rc = FN<T, 3>(__VA_ARGS__); break;       \
    case 4: rc = FN<T, 4>(__VA_ARGS__); break;       \
    case 5: rc = FN<T, 5>(__VA_ARGS__); break;       \
    case 6: rc = FN<T, 6>(__VA_ARGS__); break;       \
    case 7: rc = FN<T, 7>(__VA_ARGS__); break;       \
    default: break;                                  \
  }

template <typename T>
static int col_pass(const T* K, int64_t n, int64_t m, int64_t ld, int H, const T* U, double* scratch, const MhCtl& ctl,
                    cudaStream_t s) {
  const int full = H / 8, rest = H % 8;
  int rc = 0;
  if (full) rc = launch_col_partial<T, 8>(K, n, m, ld, H, 0, (unsigned)full, U, scratch, ctl, s);
  if (rc) return rc;
  B2OT_MH_REST(launch_col_partial, rest, K, n, m, ld, H, full * 8, 1u, U, scratch, ctl, s)
  return rc;
}
template <typename T>
static int row_pass(const T* K, int64_t n, int64_t m, int64_t ld, int H, const T* V, const T* a, T* U, const MhCtl& ctl,
                    cudaStream_t s) {
  const int full = H / 8, rest = H % 8;
  int rc = 0;
  if (full) rc = launch_row<T, 8>(K, n, m, ld, H, 0, (unsigned)full, V, a, U, ctl, s);
  if (rc) return rc;
  B2OT_MH_REST(launch_row, rest, K, n, m, ld, H, full * 8, 1u, V, a, U, ctl, s)
  return rc;
}

template <typename T>
static int multi_step(const T* K, int64_t n, int64_t m, int64_t ld, int H, const T* a, const T* B, const T* Up,
                      const T* Vp, T* U, T* V, const MhCtl& ctl, double* scratch, cudaStream_t s) {
  const int S = (int)std::min<long long>(n, MH_SPLIT);
  const unsigned cblocks = (unsigned)((m + 255) / 256);
  double* blk = ctl.stats != nullptr ? scratch + (size_t)MH_SPLIT * H * m : nullptr;
  const dim3 gfin(cblocks, (unsigned)H);
  int rc;
  if (ctl.mode == MH_UNBAL) {  // u-then-v (ot/unbalanced/_sinkhorn.py:756-759)
    if ((rc = row_pass<T>(K, n, m, ld, H, Vp, a, U, ctl, s))) return rc;
    if ((rc = col_pass<T>(K, n, m, ld, H, U, scratch, ctl, s))) return rc;
    mh_col_finish_kernel<T><<<gfin, 256, 0, s>>>(scratch, m, H, S, B, (const T*)nullptr, V, blk, ctl);
    B2OT_CHECK_LAUNCH();
  } else {  // v-then-u (ot/bregman/_sinkhorn.py:629-631, :1270-1271)
    if ((rc = col_pass<T>(K, n, m, ld, H, Up, scratch, ctl, s))) return rc;
    mh_col_finish_kernel<T><<<gfin, 256, 0, s>>>(scratch, m, H, S, B, ctl.mode == MH_KNOPP ? Vp : (const T*)nullptr, V,
                                                 blk, ctl);
    B2OT_CHECK_LAUNCH();
    if ((rc = row_pass<T>(K, n, m, ld, H, V, a, U, ctl, s))) return rc;
  }
  if (ctl.stats != nullptr) {
    mh_stats_kernel<T><<<1, 1024, 0, s>>>(U, Up, (long long)n * H, V, Vp, (long long)m * H, blk,
                                          (long long)cblocks * H, ctl);
    B2OT_CHECK_LAUNCH();
  }
  return 0;
}

extern "C" int b2ot_multi_scaling_step(int dtype, int mode, const void* K, int64_t n, int64_t m, int64_t ld, int64_t H,
                                       const void* a, const void* B, const void* u_prev, const void* v_prev,
                                       void* u_new, void* v_new, int iter, int check_every, double stop_thr,
                                       double fi_1, double fi_2, double* stats_dev, double* err_hist_dev,
                                       int err_capacity, void* scratch_dev, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_multi_scaling_step: bad dtype %d", dtype);
  B2OT_REQUIRE(mode == B2OT_MULTI_STATS || mode == B2OT_MULTI_KNOPP || mode == B2OT_MULTI_UNBALANCED,
               "b2ot_multi_scaling_step: bad mode %d", mode);
  B2OT_REQUIRE(n >= 1 && m >= 1 && ld >= m && H >= 1 && H <= 65535,
               "b2ot_multi_scaling_step: bad shape n=%lld m=%lld ld=%lld H=%lld", (long long)n, (long long)m,
               (long long)ld, (long long)H);
  B2OT_REQUIRE(K && a && B && u_prev && v_prev && u_new && v_new && scratch_dev,
               "b2ot_multi_scaling_step: null device pointer");
  B2OT_REQUIRE(u_new != u_prev && v_new != v_prev, "b2ot_multi_scaling_step: the update is out of place");
  B2OT_REQUIRE(fi_1 > 0.0 && fi_1 <= 1.0 && fi_2 > 0.0 && fi_2 <= 1.0,
               "b2ot_multi_scaling_step: exponents must be in (0, 1], got %g %g", fi_1, fi_2);
  if (mode != B2OT_MULTI_STATS) {
    B2OT_REQUIRE(stats_dev != nullptr && err_hist_dev != nullptr && err_capacity >= 1 && iter >= 0,
                 "b2ot_multi_scaling_step: the solver modes need the state block, the error history and iter >= 0");
    B2OT_REQUIRE(mode != B2OT_MULTI_KNOPP || check_every >= 1, "b2ot_multi_scaling_step: check_every must be >= 1");
  }
  MhCtl ctl;
  ctl.mode = mode;
  ctl.iter = iter;
  ctl.check_every = check_every;
  ctl.stop_thr = stop_thr;
  ctl.fi_u = fi_1;
  ctl.fi_v = fi_2;
  ctl.stats = stats_dev;
  ctl.err_hist = err_hist_dev;
  ctl.err_cap = err_capacity;
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B2OT_F32)
    return multi_step<float>((const float*)K, n, m, ld, (int)H, (const float*)a, (const float*)B,
                             (const float*)u_prev, (const float*)v_prev, (float*)u_new, (float*)v_new, ctl,
                             (double*)scratch_dev, s);
  return multi_step<double>((const double*)K, n, m, ld, (int)H, (const double*)a, (const double*)B,
                            (const double*)u_prev, (const double*)v_prev, (double*)u_new, (double*)v_new, ctl,
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
