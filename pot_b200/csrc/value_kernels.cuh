// Plan / objective epilogue of the dense solvers (sm_100a).
//
// plan_value_kernel: ONE pass that (optionally) writes the transport plan and reduces, from the same
// registers, the three sums the reference forms with separate n x m temporaries after a solve:
//     sum M * P                        value_linear            (ot/solvers.py:476, ot/bregman/_sinkhorn.py:435-438)
//     sum P * log(P + 1e-16)           entropy term            (ot/solvers.py:479)
//     sum P * log(P / q + 1e-16)       KL term, q = a_i b_j or a reference matrix c
//                                      (ot/solvers.py:481-483, ot/backend.py:1542-1546,
//                                       ot/unbalanced/_sinkhorn.py:791-797)
// and the row marginal P 1.  The Gibbs kernel can be rebuilt from the cost on the fly (K pointer null):
//     LOG   : K = -M / reg                  (ot/bregman/_sinkhorn.py:860)
//     KNOPP : K = exp(M / (-reg)) [* q]     (ot/bregman/_sinkhorn.py:621, ot/unbalanced/_sinkhorn.py:736-740)
// so that the epilogue reads M once and writes P once (2 s n m bytes) instead of the reference's
// five n x m passes.  A block owns whole rows (grid-stride over rows), 16-byte vector accesses when the
// layout allows it, per-thread double accumulators, deterministic two-level reduction.
//
// colsum_kernel / colsum_reduce_kernel: column marginal P^T 1 (deterministic split over row blocks).
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int PV_LINEAR = 1;
constexpr int PV_ENTROPY = 2;
constexpr int PV_KL = 4;
constexpr int PV_KMULQ = 8;  // K rebuilt from the cost is multiplied by q (unbalanced: K = exp(-M/reg) * c)

struct PlanValueParams {
  const void* K;  // n x m or null (rebuild from M with reg)
  const void* M;  // cost, needed for PV_LINEAR or when K is null
  const void* C;  // optional reference matrix q_ij; null: q_ij = a_i b_j
  const void *u, *v, *a, *b;
  void* out;      // plan or null
  void* rowsum;   // n values of T or null
  long long n, m, ldK, ldM, ldC, ldO;
  double reg;
  int terms;
  int nblk;
  double* blk;    // [3][nblk]
};

template <typename T> __device__ __forceinline__ T eps16();
template <> __device__ __forceinline__ float eps16<float>() { return 1e-16f; }
template <> __device__ __forceinline__ double eps16<double>() { return 1e-16; }

// log / divide of the REDUCED terms only (the plan itself uses the accurate exp): fp32 takes the MUFU forms
// (lg2.approx * ln 2, rcp-based divide; ~1e-7 relative), far inside the 1e-4 budget of a sum of n m terms.
__device__ __forceinline__ float log_red(float x) { return __logf(x); }
__device__ __forceinline__ double log_red(double x) { return log(x); }
__device__ __forceinline__ float div_red(float x, float y) { return __fdividef(x, y); }
__device__ __forceinline__ double div_red(double x, double y) { return x / y; }

template <typename T, int VEC>
__device__ __forceinline__ void load_vec(const T* p, T (&x)[VEC]) {
  if constexpr (VEC == 1) {
    x[0] = *p;
  } else {
    using V = typename Vec<T>::type;
    const V q = *reinterpret_cast<const V*>(p);
#pragma unroll
    for (int e = 0; e < VEC; ++e) x[e] = elem(q, e);
  }
}
template <typename T, int VEC>
__device__ __forceinline__ void load_vec_stream(const T* p, T (&x)[VEC]) {
  if constexpr (VEC == 1) {
    x[0] = __ldcs(p);
  } else {
    using V = typename Vec<T>::type;
    const V q = __ldcs(reinterpret_cast<const V*>(p));
#pragma unroll
    for (int e = 0; e < VEC; ++e) x[e] = elem(q, e);
  }
}
template <typename T, int VEC>
__device__ __forceinline__ void store_vec_stream(T* p, const T (&x)[VEC]) {
  if constexpr (VEC == 1) {
    __stcs(p, x[0]);
  } else {
    using V = typename Vec<T>::type;
    V q;
#pragma unroll
    for (int e = 0; e < VEC; ++e) elem(q, e) = x[e];
    __stcs(reinterpret_cast<V*>(p), q);
  }
}

__device__ __forceinline__ double block_sum_256(double x, double* sred) {
  x = warp_sum(x);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = x;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x < 8) t = sred[threadIdx.x];
  if (threadIdx.x < 32) {
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;  // valid in thread 0
}

// SPEC >= 0: the per-element conditions are compile-time constants (straight-line code per vector, no branches inside
// the element loop): bits 0-3 = terms, 16 = K given, 32 = reference matrix given, 64 = plan written.  SPEC = -1: read
// them from the parameters (any combination).
constexpr int PV_SPEC_K = 16, PV_SPEC_C = 32, PV_SPEC_OUT = 64;
#ifndef B2OT_PV_UNROLL
#define B2OT_PV_UNROLL 2
#endif
constexpr int kPvUnroll = B2OT_PV_UNROLL;  // column vectors per thread and trip in fp32

template <typename T, int MODE, bool VECT, int SPEC = -1>
__global__ void __launch_bounds__(256) plan_value_kernel(const PlanValueParams p) {
  constexpr int VEC = VECT ? Vec<T>::N : 1;
  constexpr bool CT = SPEC >= 0;
  __shared__ double sred[8];
  const T* K = reinterpret_cast<const T*>(p.K);
  const T* M = reinterpret_cast<const T*>(p.M);
  const T* Cq = reinterpret_cast<const T*>(p.C);
  const T* u = reinterpret_cast<const T*>(p.u);
  const T* v = reinterpret_cast<const T*>(p.v);
  const T* a = reinterpret_cast<const T*>(p.a);
  const T* b = reinterpret_cast<const T*>(p.b);
  T* out = reinterpret_cast<T*>(p.out);
  T* rowsum = reinterpret_cast<T*>(p.rowsum);
  const int terms = CT ? (SPEC & 15) : p.terms;
  const bool hasK = CT ? ((SPEC & PV_SPEC_K) != 0) : (K != nullptr);
  const bool hasC = CT ? ((SPEC & PV_SPEC_C) != 0) : (Cq != nullptr);
  const bool hasOut = CT ? ((SPEC & PV_SPEC_OUT) != 0) : (out != nullptr);
  const bool need_m = (terms & PV_LINEAR) || !hasK;
  const bool need_q = (terms & PV_KL) || ((terms & PV_KMULQ) && !hasK);
  const T reg = (T)p.reg;
  // x / reg as x * (1/reg) with one FMA refinement step (Markstein): the correctly rounded quotient except in rare
  // one-ulp cases, without the IEEE division's special-case path (a quarter of this kernel's instructions); one ulp of
  // -M/reg moves a plan entry by < 1e-5 relative in fp32, 1e-14 in fp64
  const T rinv = (T)1 / reg;
  auto div_reg = [&](T x) {
    const T q = x * rinv;
    return fma(fma(-q, reg, x), rinv, q);
  };
  double s_lin = 0.0, s_ent = 0.0, s_kl = 0.0;
  for (long long i = blockIdx.x; i < p.n; i += gridDim.x) {
    const T ui = u[i];
    const T ai = (need_q && !hasC) ? a[i] : (T)0;
    // per-row partial sums in T (at most m / 256 terms per thread), folded into the double accumulators once per row
    T r_sum = (T)0, r_lin = (T)0, r_ent = (T)0, r_kl = (T)0;
#pragma unroll(sizeof(T) == 4 ? kPvUnroll : 1)  // fp64 is FP64-pipe bound: unrolling only costs occupancy there
    for (long long j = (long long)threadIdx.x * VEC; j < p.m; j += 256LL * VEC) {
      T kx[VEC], mx[VEC], vx[VEC], qx[VEC], px[VEC];
      load_vec<T, VEC>(v + j, vx);
      if (hasK) load_vec_stream<T, VEC>(K + i * p.ldK + j, kx);
      if (need_m) load_vec_stream<T, VEC>(M + i * p.ldM + j, mx);
      if (need_q) {
        if (hasC) {
          load_vec_stream<T, VEC>(Cq + i * p.ldC + j, qx);
        } else {
          load_vec<T, VEC>(b + j, qx);
#pragma unroll
          for (int e = 0; e < VEC; ++e) qx[e] = ai * qx[e];
        }
      }
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        T k;
        if (hasK) {
          k = kx[e];
        } else if (MODE == MODE_LOG) {
          k = div_reg(-mx[e]);
        } else {
          k = exp_t(-div_reg(mx[e]));
          if (terms & PV_KMULQ) k = k * qx[e];
        }
        const T g = (MODE == MODE_KNOPP) ? (ui * k) * vx[e] : exp_t((k + ui) + vx[e]);
        px[e] = g;
        r_sum += g;
        if (terms & PV_LINEAR) r_lin = fma(mx[e], g, r_lin);
        if (terms & PV_ENTROPY) r_ent = fma(g, log_red(g + eps16<T>()), r_ent);
        if (terms & PV_KL) r_kl = fma(g, log_red(div_red(g, qx[e]) + eps16<T>()), r_kl);
      }
      if (hasOut) store_vec_stream<T, VEC>(out + i * p.ldO + j, px);
    }
    s_lin += (double)r_lin;
    s_ent += (double)r_ent;
    s_kl += (double)r_kl;
    if (rowsum != nullptr) {  // block-uniform branch
      const double t = block_sum_256((double)r_sum, sred);
      if (threadIdx.x == 0) rowsum[i] = (T)t;
    }
  }
  double t = block_sum_256(s_lin, sred);
  if (threadIdx.x == 0) p.blk[blockIdx.x] = t;
  t = block_sum_256(s_ent, sred);
  if (threadIdx.x == 0) p.blk[p.nblk + blockIdx.x] = t;
  t = block_sum_256(s_kl, sred);
  if (threadIdx.x == 0) p.blk[2 * p.nblk + blockIdx.x] = t;
}

// values[t] = sum_b blk[t][b]  (grid = 3 blocks, fixed summation order)
__global__ void __launch_bounds__(256) plan_value_reduce_kernel(const double* blk, int nblk, double* values) {
  __shared__ double sred[8];
  double x = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 256) x += blk[(long long)blockIdx.x * nblk + i];
  const double t = block_sum_256(x, sred);
  if (threadIdx.x == 0) values[blockIdx.x] = t;
}

// partial[s][j] = sum over rows i = s, s + S, ... of P_ij   (grid = (ceil(m / 256), S))
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ P, long long n, long long m, long long ld,
                                                     double* __restrict__ partial) {
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  if (j >= m) return;
  double s = 0.0;
  for (long long i = blockIdx.y; i < n; i += gridDim.y) s += (double)P[i * ld + j];
  partial[(long long)blockIdx.y * m + j] = s;
}
template <typename T>
__global__ void __launch_bounds__(256) colsum_reduce_kernel(const double* __restrict__ partial, long long m, int S,
                                                            T* __restrict__ out) {
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  if (j >= m) return;
  double s = 0.0;
  for (int q = 0; q < S; ++q) s += partial[(long long)q * m + j];
  out[j] = (T)s;
}

}  // namespace b2ot
