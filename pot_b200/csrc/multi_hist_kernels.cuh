// Scaling iteration for SEVERAL target histograms that share one kernel matrix (sm_100a).
//
// ot.bregman.sinkhorn_stabilized with a 2-D b (ot/bregman/_sinkhorn.py:1230-1234) keeps u (n x H) and v (m x H) on ONE
// K = exp(-(M - alpha - beta) / reg): the H problems are coupled through the joint absorption (:1275-1279) and the joint
// error (:1297-1301), so they cannot run as H independent solves.  One iteration (:1270-1271) is
//     V = B / (K^T U)        U = a / (K V)
// i.e. two passes over K with H right-hand sides each - K is read twice per iteration whatever H is (HBM / L2 bound, the
// arithmetic is H FMAs per loaded element).  Kernels:
//   mh_col_partial_kernel  partial[s][h][j] = sum over rows i = s, s + S, ... of K_ij U_ih   (coalesced along j, the H
//                          values of row i are warp-uniform loads); deterministic split over S row blocks;
//   mh_col_finish_kernel   V_jh = B_jh / sum_s partial[s][h][j];
//   mh_row_kernel          one warp per row i: U_ih = a_i / sum_j K_ij V_jh (lane-strided, fixed shuffle tree);
//   mh_stats_kernel        one block: max|U|, max|V|, max|U - Uprev|, max|Uprev|, max|V - Vprev|, max|Vprev| and the
//                          "contains nan" flags of U and V - everything the host's absorption / error / stop tests need,
//                          one 64-byte read per iteration (the reference syncs every iteration at :1274 as well);
//   mh_absorb_kernel       one block: alpha/reg += max_h log U_ih, beta/reg += max over ALL of log V (a scalar - that is
//                          what :1278 computes), in double; U, V are kept (:1275-1279 has no reset in this branch).
// Sums are accumulated in double for both dtypes and rounded once.  Histograms are processed MH_HC at a time
// (blockIdx.z selects the chunk).
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int MH_HC = 4;       // histograms per thread
constexpr int MH_SPLIT = 32;   // row blocks of the column pass

template <typename T>
__global__ void __launch_bounds__(256) mh_col_partial_kernel(const T* __restrict__ K, long long n, long long m,
                                                             long long ld, int H, const T* __restrict__ U,
                                                             double* __restrict__ partial) {
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const int h0 = blockIdx.z * MH_HC;
  if (j >= m) return;
  double acc[MH_HC];
#pragma unroll
  for (int c = 0; c < MH_HC; ++c) acc[c] = 0.0;
  for (long long i = blockIdx.y; i < n; i += gridDim.y) {
    const double k = (double)K[i * ld + j];
#pragma unroll
    for (int c = 0; c < MH_HC; ++c)
      if (h0 + c < H) acc[c] += k * (double)U[i * H + h0 + c];
  }
#pragma unroll
  for (int c = 0; c < MH_HC; ++c)
    if (h0 + c < H) partial[((long long)blockIdx.y * H + h0 + c) * m + j] = acc[c];
}

// grid = (ceil(m / 256), H)
template <typename T>
__global__ void __launch_bounds__(256) mh_col_finish_kernel(const double* __restrict__ partial, long long m, int H,
                                                            int S, const T* __restrict__ B, T* __restrict__ V) {
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const int h = blockIdx.y;
  if (j >= m) return;
  double s = 0.0;
  for (int q = 0; q < S; ++q) s += partial[((long long)q * H + h) * m + j];
  V[j * H + h] = B[j * H + h] / (T)s;
}

// grid = (ceil(n / 8), 1, ceil(H / MH_HC)), 8 warps per block, one row per warp
template <typename T>
__global__ void __launch_bounds__(256) mh_row_kernel(const T* __restrict__ K, long long n, long long m, long long ld,
                                                     int H, const T* __restrict__ V, const T* __restrict__ a,
                                                     T* __restrict__ U) {
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int h0 = blockIdx.z * MH_HC;
  if (i >= n) return;
  double acc[MH_HC];
#pragma unroll
  for (int c = 0; c < MH_HC; ++c) acc[c] = 0.0;
  const T* row = K + i * ld;
  for (long long j = lane; j < m; j += 32) {
    const double k = (double)row[j];
#pragma unroll
    for (int c = 0; c < MH_HC; ++c)
      if (h0 + c < H) acc[c] += k * (double)V[j * H + h0 + c];
  }
#pragma unroll
  for (int c = 0; c < MH_HC; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_down_sync(0xffffffffu, acc[c], o);
  }
  if (lane == 0) {
    const T ai = a[i];
#pragma unroll
    for (int c = 0; c < MH_HC; ++c)
      if (h0 + c < H) U[i * H + h0 + c] = ai / (T)acc[c];
  }
}

// block-wide max of non-negative values (nan-free: callers pass fmax-filtered values); result valid in every thread
__device__ __forceinline__ double mh_block_max(double x, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_xor_sync(0xffffffffu, x, o));
  __syncthreads();  // sh may still be read from the previous call
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = x;
  __syncthreads();
  double r = sh[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = fmax(r, sh[w]);
  return r;
}

// stats[0..7] = max|U|, max|V|, max|U - Uprev|, max|Uprev|, max|V - Vprev|, max|Vprev|, nan in U, nan in V
// (the maxima skip nan entries; the host puts the nan back where numpy's max would propagate it).  One block.
template <typename T>
__global__ void __launch_bounds__(1024) mh_stats_kernel(const T* __restrict__ U, const T* __restrict__ Up,
                                                        long long nu, const T* __restrict__ V,
                                                        const T* __restrict__ Vp, long long nv,
                                                        double* __restrict__ stats) {
  __shared__ double sh[32];
  double mu = 0.0, mdu = 0.0, mup = 0.0, mv = 0.0, mdv = 0.0, mvp = 0.0, nanu = 0.0, nanv = 0.0;
  for (long long q = threadIdx.x; q < nu; q += blockDim.x) {
    const double x = (double)U[q], p = (double)Up[q];
    if (x != x) nanu = 1.0;
    mu = fmax(mu, fabs(x));
    mup = fmax(mup, fabs(p));
    mdu = fmax(mdu, fabs((double)(T)(U[q] - Up[q])));
  }
  for (long long q = threadIdx.x; q < nv; q += blockDim.x) {
    const double x = (double)V[q], p = (double)Vp[q];
    if (x != x) nanv = 1.0;
    mv = fmax(mv, fabs(x));
    mvp = fmax(mvp, fabs(p));
    mdv = fmax(mdv, fabs((double)(T)(V[q] - Vp[q])));
  }
  mu = mh_block_max(mu, sh);
  mv = mh_block_max(mv, sh);
  mdu = mh_block_max(mdu, sh);
  mup = mh_block_max(mup, sh);
  mdv = mh_block_max(mdv, sh);
  mvp = mh_block_max(mvp, sh);
  nanu = mh_block_max(nanu, sh);
  nanv = mh_block_max(nanv, sh);
  if (threadIdx.x == 0) {
    stats[0] = mu;
    stats[1] = mv;
    stats[2] = mdu;
    stats[3] = mup;
    stats[4] = mdv;
    stats[5] = mvp;
    stats[6] = nanu;
    stats[7] = nanv;
  }
}

// nan-propagating max, like numpy's
__device__ __forceinline__ double mh_nanmax(double x, double y) { return (x != x || y != y) ? (x + y) : fmax(x, y); }

// alpha_r[i] += max_h log U_ih;  beta_r[j] += max over all (j, h) of log V_jh.  One block.
template <typename T>
__global__ void __launch_bounds__(1024) mh_absorb_kernel(const T* __restrict__ U, long long n, const T* __restrict__ V,
                                                         long long m, int H, double* __restrict__ alpha_r,
                                                         double* __restrict__ beta_r) {
  __shared__ double sh[32];
  __shared__ double lvmax;
  double lv = -INFINITY;
  for (long long q = threadIdx.x; q < m * H; q += blockDim.x) lv = mh_nanmax(lv, log((double)V[q]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lv = mh_nanmax(lv, __shfl_xor_sync(0xffffffffu, lv, o));
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = lv;
  __syncthreads();
  if (threadIdx.x == 0) {
    double r = sh[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = mh_nanmax(r, sh[w]);
    lvmax = r;
  }
  __syncthreads();
  const double add = lvmax;
  for (long long j = threadIdx.x; j < m; j += blockDim.x) beta_r[j] += add;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) {
    double lu = -INFINITY;
    for (int h = 0; h < H; ++h) lu = mh_nanmax(lu, log((double)U[i * H + h]));
    alpha_r[i] += lu;
  }
}

}  // namespace b2ot
