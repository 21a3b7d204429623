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
  // Element q of a thread is element q = 0 plus q times a fixed (k, row) step, so a thread keeps ONE pointer per operand
  // that advances by a chunk, a row / column validity mask, and checks k only in the last chunk of a feature (the
  // per-element index arithmetic and bounds checks of the first version were as many instructions as the FMAs).
  const bool a_kfast = (p.sAk == 1) || (p.sAi != 1);
  const bool b_jfast = (p.sBj == 1) || (p.sBk != 1);
  const int a_k0 = a_kfast ? (tid & 15) : (tid / BM), a_i0 = a_kfast ? (tid >> 4) : (tid % BM);
  const int a_dk = a_kfast ? 0 : 256 / BM, a_di = a_kfast ? 16 : 0;
  const int b_k0 = b_jfast ? (tid / BN) : (tid & 15), b_j0 = b_jfast ? (tid % BN) : (tid >> 4);
  const int b_dk = b_jfast ? 256 / BN : 0, b_dj = b_jfast ? 0 : 16;
  const long long a_qs = (long long)a_dk * p.sAk + (long long)a_di * p.sAi;
  const long long b_qs = (long long)b_dk * p.sBk + (long long)b_dj * p.sBj;
  unsigned a_ok = 0, b_ok = 0;
#pragma unroll
  for (int q = 0; q < TM; ++q) a_ok |= (i0 + a_i0 + q * a_di < p.M) ? (1u << q) : 0u;
#pragma unroll
  for (int q = 0; q < TN; ++q) b_ok |= (j0 + b_j0 + q * b_dj < p.N) ? (1u << q) : 0u;
  const T* a_base = reinterpret_cast<const T*>(p.A) + b * p.sAb + f * p.sAf + (i0 + a_i0) * p.sAi + (long long)a_k0 * p.sAk;
  const T* b_base = reinterpret_cast<const T*>(p.B) + b * p.sBb + f * p.sBf + (j0 + b_j0) * p.sBj + (long long)b_k0 * p.sBk;
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
    const T* ap = a_base + s * p.sAf + k0 * p.sAk;
    const T* bp = b_base + s * p.sBf + k0 * p.sBk;
    if (k0 + GM_K <= p.K) {
#pragma unroll
      for (int q = 0; q < TM; ++q) ra[q] = (a_ok >> q & 1u) ? ap[q * a_qs] : (T)0;
#pragma unroll
      for (int q = 0; q < TN; ++q) rb[q] = (b_ok >> q & 1u) ? bp[q * b_qs] : (T)0;
    } else {
#pragma unroll
      for (int q = 0; q < TM; ++q) ra[q] = ((a_ok >> q & 1u) && k0 + a_k0 + q * a_dk < p.K) ? ap[q * a_qs] : (T)0;
#pragma unroll
      for (int q = 0; q < TN; ++q) rb[q] = ((b_ok >> q & 1u) && k0 + b_k0 + q * b_dk < p.K) ? bp[q * b_qs] : (T)0;
    }
  };
  auto stage = [&]() {
#pragma unroll
    for (int q = 0; q < TM; ++q) As[a_k0 + q * a_dk][a_i0 + q * a_di] = ra[q];
#pragma unroll
    for (int q = 0; q < TN; ++q) Bs[b_k0 + q * b_dk][b_j0 + q * b_dj] = rb[q];
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
