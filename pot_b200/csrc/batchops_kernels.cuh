// Batched cost matrices and per-problem reductions of ot.batch (sm_100a).
//
//   batch_dist_kernel<T, METRIC>   M[b] = dist(X[b], Y[b])  for B problems in one launch
//       (ot/batch/_linear.py: dist_euclidean_batch :54-85 -- NO clamp at 0, unlike ot.dist; dist_lp_batch :19-51;
//        dist_kl_batch :88-120).  Bound by the HBM write of the (B, n, m) output: a 64 x 64 tile per block, 4 x 4
//       values per thread, the coordinates staged through shared memory in chunks of 16.
//   batch_reduce_kernel<T, OP>     per-problem sums over the (n, m) plane: <M_b, T_b> (loss_linear_batch, :123-147)
//       and sum T log(T + eps) (entropy_batch, ot/batch/_utils.py:4-22).  Bound by the HBM read: 16-byte loads,
//       fp64 accumulators, a fixed-order two-level reduction (deterministic).
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int BD_TILE = 64;
constexpr int BD_KC = 16;
constexpr int BDM_SQEUCLIDEAN = 0, BDM_EUCLIDEAN = 1, BDM_MINKOWSKI = 2, BDM_KL = 3;

struct BatchDistParams {
  const void* X;   // (B, n, d) contiguous
  const void* Y;   // (B, m, d) contiguous
  const void* w;   // optional d weights of the Minkowski metric
  void* out;       // (B, n, m) contiguous
  long long B, n, m;
  int d;
  long long tiles_n, tiles_m;
  double p, eps;
  int vec_ok;      // out rows allow 16-byte stores
};

template <typename T> __device__ __forceinline__ T sqrt_t(T x);
template <> __device__ __forceinline__ float sqrt_t<float>(float x) { return sqrtf(x); }
template <> __device__ __forceinline__ double sqrt_t<double>(double x) { return sqrt(x); }

template <typename T, int METRIC>
__global__ void __launch_bounds__(256) batch_dist_kernel(const BatchDistParams p) {
  __shared__ T xs[BD_KC][BD_TILE + 1];
  __shared__ T ys[BD_KC][BD_TILE + 1];
  const long long per = p.tiles_n * p.tiles_m;
  const long long blk = blockIdx.x;
  const long long b = blk / per;
  const long long t = blk - b * per;
  const long long row0 = (t / p.tiles_m) * BD_TILE, col0 = (t % p.tiles_m) * BD_TILE;
  const T* X = (const T*)p.X + b * p.n * p.d;
  const T* Y = (const T*)p.Y + b * p.m * p.d;
  const T* W = (const T*)p.w;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const T eps = (T)p.eps, pw = (T)p.p;
  T acc[4][4], xa[4], ya[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    xa[i] = ya[i] = (T)0;
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = (T)0;
  }
  for (int k0 = 0; k0 < p.d; k0 += BD_KC) {
    // stage 64 rows x 16 coordinates of each cloud (transposed); KL stores log(x + eps)
    const int k = threadIdx.x & 15;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int row = (threadIdx.x >> 4) + 16 * r;
      const bool kin = k0 + k < p.d;
      T xv = (T)0, yv = (T)0;
      if (kin && row0 + row < p.n) xv = X[(row0 + row) * p.d + k0 + k];
      if (kin && col0 + row < p.m) yv = Y[(col0 + row) * p.d + k0 + k];
      if (METRIC == BDM_KL) xv = kin ? log_t(xv + eps) : (T)0;
      xs[k][row] = xv;
      ys[k][row] = yv;
    }
    __syncthreads();
    const int kn = min(BD_KC, p.d - k0);
    for (int kk = 0; kk < kn; ++kk) {
      T xv[4], yv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        xv[i] = xs[kk][ty * 4 + i];
        yv[i] = ys[kk][tx * 4 + i];
      }
      if (METRIC == BDM_SQEUCLIDEAN || METRIC == BDM_EUCLIDEAN) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          xa[i] += xv[i] * xv[i];
          ya[i] += yv[i] * yv[i];
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] += xv[i] * yv[j];
        }
      } else if (METRIC == BDM_MINKOWSKI) {
        const T wk = W ? W[k0 + kk] : (T)1;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const T dd = abs_t(xv[i] - yv[j]);
            const T pp = pw == (T)1 ? dd : (pw == (T)2 ? dd * dd : pow_t(dd, pw));
            acc[i][j] += wk * pp;
          }
      } else {  // KL(y_j || x_i) = sum_d y log(y + eps) - sum_d y log(x + eps)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          ya[i] += yv[i] * log_t(yv[i] + eps);
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] += yv[j] * xv[i];
        }
      }
    }
    __syncthreads();
  }
  T* O = (T*)p.out + b * p.n * p.m;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long row = row0 + ty * 4 + i;
    if (row >= p.n) continue;
    T r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (METRIC == BDM_SQEUCLIDEAN) r[j] = (xa[i] + ya[j]) - (T)2 * acc[i][j];
      else if (METRIC == BDM_EUCLIDEAN) r[j] = sqrt_t((xa[i] + ya[j]) - (T)2 * acc[i][j]);  // negative round-off -> nan, as the reference
      else if (METRIC == BDM_MINKOWSKI) r[j] = pw == (T)1 ? acc[i][j] : (pw == (T)2 ? sqrt_t(acc[i][j]) : pow_t(acc[i][j], (T)1 / pw));
      else r[j] = ya[j] - acc[i][j];
    }
    const long long col = col0 + tx * 4;
    T* dst = O + row * p.m + col;
    if (p.vec_ok && col + 3 < p.m) {
      if (sizeof(T) == 4) {
        *reinterpret_cast<float4*>(dst) = make_float4((float)r[0], (float)r[1], (float)r[2], (float)r[3]);
      } else {
        reinterpret_cast<double2*>(dst)[0] = make_double2((double)r[0], (double)r[1]);
        reinterpret_cast<double2*>(dst)[1] = make_double2((double)r[2], (double)r[3]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (col + j < p.m) dst[j] = r[j];
    }
  }
}

// ---- per-problem reductions ----
constexpr int BR_DOT = 0;      // sum A * B
constexpr int BR_ENTROPY = 1;  // sum A * log(A + eps)

template <typename T, int OP>
__global__ void __launch_bounds__(256) batch_reduce_kernel(const T* __restrict__ A, const T* __restrict__ Bm, long long per,
                                                           int nsplit, long long chunk, double eps, int vec_ok,
                                                           double* __restrict__ partial) {
  using V = typename Vec<T>::type;
  constexpr int VN = Vec<T>::N;
  const long long b = blockIdx.x / nsplit;
  const int s = blockIdx.x % nsplit;
  const long long lo = (long long)s * chunk, hi = min(per, lo + chunk);
  const T* a = A + b * per;
  const T* bm = OP == BR_DOT ? Bm + b * per : nullptr;
  const T e = (T)eps;
  double acc0 = 0.0, acc1 = 0.0;
  if (vec_ok) {  // chunk is a multiple of VN and the planes are 16-byte aligned
    const long long nv = (hi - lo) / VN;
    const V* av = reinterpret_cast<const V*>(a + lo);
    const V* bv = OP == BR_DOT ? reinterpret_cast<const V*>(bm + lo) : nullptr;
    for (long long i = threadIdx.x; i < nv; i += 256) {
      const V x = __ldcs(av + i);
      T part = (T)0;
      if (OP == BR_DOT) {
        const V y = __ldcs(bv + i);
#pragma unroll
        for (int q = 0; q < VN; ++q) part += elem(x, q) * elem(y, q);
      } else {
#pragma unroll
        for (int q = 0; q < VN; ++q) part += elem(x, q) * log_t(elem(x, q) + e);
      }
      if (i & 256) acc1 += (double)part; else acc0 += (double)part;
    }
    for (long long i = lo + nv * VN + threadIdx.x; i < hi; i += 256)
      acc0 += OP == BR_DOT ? (double)(a[i] * bm[i]) : (double)(a[i] * log_t(a[i] + e));
  } else {
    for (long long i = lo + threadIdx.x; i < hi; i += 256)
      acc0 += OP == BR_DOT ? (double)(a[i] * bm[i]) : (double)(a[i] * log_t(a[i] + e));
  }
  double v = warp_sum(acc0 + acc1);
  __shared__ double red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += red[w];
    partial[blockIdx.x] = tot;
  }
}

// out[b] = sign * sum_s partial[b][s] in a fixed order
template <typename T>
__global__ void batch_reduce_final_kernel(const double* __restrict__ partial, long long B, int nsplit, double sign,
                                          T* __restrict__ out) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double tot = 0.0;
  for (int s = 0; s < nsplit; ++s) tot += partial[b * nsplit + s];
  out[b] = (T)(sign * tot);
}

}  // namespace b2ot
