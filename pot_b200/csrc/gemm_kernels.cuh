// Strided, batched matrix product for the Gromov-Wasserstein contraction (sm_100a).
//
//   C[b, f](i, j) = beta * C0[b](i, j) + alpha * sum_{s < Sf} sum_k A[b, f + s](i, k) * B[b, f + s](k, j)
//
// Every operand is addressed by element strides (batch b, feature f, row, column), so transposes and the interleaved
// feature axis of the (B, n, n, d) structure tensors need no copies.  It serves
//   * ot/gromov/_utils.py:244-280   tensor_product:        constC - hC1 T hC2^T                      (two calls)
//   * ot/batch/_quadratic.py:535-572 tensor_product_batch:  const - sum_d hC1[..., d] T hC2[..., d]^T  (two calls, four
//     when the structures are asymmetric)
// which were torch.mm / torch.einsum (cuBLAS) until round 2.
//
// Register-tiled FMA kernel on the FP32 / FP64 pipes: a block of 256 threads owns a (16 TM) x (16 TN) output tile, a
// thread a TM x TN micro-tile (fp32: 8 x 8 on a 128 x 128 tile, fp64: 8 x 4 on 128 x 64; micro-tile rows / columns come
// in groups of four, 64 apart, so the shared-memory reads are conflict-free 16-byte vectors); K advances in chunks of 16
// through shared memory (the next chunk is fetched into registers while the current one is multiplied).  A first
// version with 4 x 4 micro-tiles read 8 shared-memory values per 16 FMA and sat at the shared-memory bound (50 % of the
// FMA peak in fp32, 32 % in fp64: scripts/bench_gemm.py).  The global-load mapping of each operand follows its unit
// stride (coalesced along k or along the row / column, whichever is contiguous).  Fixed summation order:
// deterministic.  fp64 is what the Gromov solvers run in (parity target 1e-10); the tensor pipe has no fp64 path, and
// fp32 on tcgen05 with the operand split of emp_tc_kernels.cuh is the listed next step.
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int GM_K = 16;    // K chunk
constexpr int GM_PAD = 4;

struct GemmParams {
  const void* A;
  const void* B;
  const void* C0;   // may be null (beta ignored)
  void* C;
  long long M, N, K;
  long long Zf, Sf;                    // output feature count, summed feature count
  long long sAb, sAf, sAi, sAk;
  long long sBb, sBf, sBk, sBj;
  long long sC0b, sC0i;                // unit column stride
  long long sCb, sCf, sCi;             // unit column stride
  double alpha, beta;
};

// offset of micro-tile entry r (0..TX-1) of thread t (0..15) inside a tile edge of 16 * TX: groups of four, 64 apart
template <int TX>
__device__ __forceinline__ int gm_off(int t, int r) {
  return TX == 8 ? ((r >> 2) * 64 + t * 4 + (r & 3)) : (t * TX + r);
}

template <typename T, int TM, int TN>
__global__ void __launch_bounds__(256) strided_gemm_kernel(const GemmParams p) {
  constexpr int BM = 16 * TM, BN = 16 * TN;
  __shared__ __align__(16) T As[GM_K][BM + GM_PAD];
  __shared__ __align__(16) T Bs[GM_K][BN + GM_PAD];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const long long z = blockIdx.z, b = z / p.Zf, f = z % p.Zf;
  const long long i0 = (long long)blockIdx.y * BM, j0 = (long long)blockIdx.x * BN;
  // global -> register staging: TM (TN) elements of A (B) per thread and chunk.  Mapping by contiguity:
  //   A: k fastest across threads when A is contiguous along k (sAk == 1), rows fastest otherwise;
  //   B: columns fastest when B is contiguous along j (sBj == 1), k fastest otherwise.
  const bool a_kfast = (p.sAk == 1) || (p.sAi != 1);
  const bool b_jfast = (p.sBj == 1) || (p.sBk != 1);
  T ra[TM], rb[TN];
  T acc[TM][TN];
#pragma unroll
  for (int r = 0; r < TM; ++r)
#pragma unroll
    for (int c = 0; c < TN; ++c) acc[r][c] = (T)0;

  const long long nchunk = (p.K + GM_K - 1) / GM_K;
  const long long steps = p.Sf * nchunk;

  auto fetch = [&](long long step) {
    const long long s = step / nchunk, k0 = (step % nchunk) * GM_K;
    const T* Ab = reinterpret_cast<const T*>(p.A) + b * p.sAb + (f + s) * p.sAf;
    const T* Bb = reinterpret_cast<const T*>(p.B) + b * p.sBb + (f + s) * p.sBf;
#pragma unroll
    for (int q = 0; q < TM; ++q) {
      const int e = tid + q * 256;  // 0 .. 16 BM - 1
      const int ak = a_kfast ? (e & 15) : (e / BM), ai = a_kfast ? (e >> 4) : (e % BM);
      const long long gi = i0 + ai, gk = k0 + ak;
      ra[q] = (gi < p.M && gk < p.K) ? Ab[gi * p.sAi + gk * p.sAk] : (T)0;
    }
#pragma unroll
    for (int q = 0; q < TN; ++q) {
      const int e = tid + q * 256;
      const int bk = b_jfast ? (e / BN) : (e & 15), bj = b_jfast ? (e % BN) : (e >> 4);
      const long long gj = j0 + bj, gk = k0 + bk;
      rb[q] = (gj < p.N && gk < p.K) ? Bb[gk * p.sBk + gj * p.sBj] : (T)0;
    }
  };
  auto stage = [&]() {
#pragma unroll
    for (int q = 0; q < TM; ++q) {
      const int e = tid + q * 256;
      const int ak = a_kfast ? (e & 15) : (e / BM), ai = a_kfast ? (e >> 4) : (e % BM);
      As[ak][ai] = ra[q];
    }
#pragma unroll
    for (int q = 0; q < TN; ++q) {
      const int e = tid + q * 256;
      const int bk = b_jfast ? (e / BN) : (e & 15), bj = b_jfast ? (e % BN) : (e >> 4);
      Bs[bk][bj] = rb[q];
    }
  };

  if (steps > 0) fetch(0);
  for (long long step = 0; step < steps; ++step) {
    __syncthreads();  // the previous chunk has been consumed
    stage();
    __syncthreads();
    if (step + 1 < steps) fetch(step + 1);
#pragma unroll
    for (int kk = 0; kk < GM_K; ++kk) {
      T a[TM], bb[TN];
#pragma unroll
      for (int r = 0; r < TM; ++r) a[r] = As[kk][gm_off<TM>(ty, r)];
#pragma unroll
      for (int c = 0; c < TN; ++c) bb[c] = Bs[kk][gm_off<TN>(tx, c)];
#pragma unroll
      for (int r = 0; r < TM; ++r)
#pragma unroll
        for (int c = 0; c < TN; ++c) acc[r][c] += a[r] * bb[c];
    }
  }

  const T alpha = (T)p.alpha, beta = (T)p.beta;
  T* Cb = reinterpret_cast<T*>(p.C) + b * p.sCb + f * p.sCf;
  const T* C0b = p.C0 != nullptr ? reinterpret_cast<const T*>(p.C0) + b * p.sC0b : nullptr;
#pragma unroll
  for (int r = 0; r < TM; ++r) {
    const long long gi = i0 + gm_off<TM>(ty, r);
    if (gi >= p.M) continue;
#pragma unroll
    for (int c = 0; c < TN; ++c) {
      const long long gj = j0 + gm_off<TN>(tx, c);
      if (gj >= p.N) continue;
      T v = alpha * acc[r][c];
      if (C0b != nullptr) v += beta * C0b[gi * p.sC0i + gj];
      Cb[gi * p.sCi + gj] = v;
    }
  }
}

}  // namespace b2ot
