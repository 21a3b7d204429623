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
// Sums are accumulated in double for both dtypes (short trip sums in T, see mh_col_partial_kernel) and rounded once.
// Up to 8 histograms per thread; more run as further chunks (blockIdx.z).
//
// The same step serves the two other solvers that keep several histograms on one K, with the decisions on the device
// (MhCtl::mode; the state lives in the 16 doubles of `stats`, every kernel of a step returns at once when the solve has
// ended, so the host enqueues steps in chunks and polls the status word like the dense solver's):
//   MH_KNOPP  ot.bregman.sinkhorn_knopp, 2-D b (ot/bregman/_sinkhorn.py:626-663): v-then-u, guard (K^T u == 0, nan / inf
//             in u or v -> previous iterate, stop :633-645), joint error || v * K^T u - b ||_F every check_every
//             iterations (:649-653) - read off the NEXT step's column pass (mh_col_finish_kernel), so an iteration stays
//             two passes over K; `err < stop_thr` stops at the checked iterate;
//   MH_UNBAL  ot.unbalanced.sinkhorn_knopp_unbalanced, 2-D b (ot/unbalanced/_sinkhorn.py:749-787): u-then-v with the
//             exponents fi_1, fi_2, same guard, error (|u-uprev|_inf / max(|u|_inf, |uprev|_inf, 1) + same for v) / 2
//             over ALL histograms every iteration, `err < stop_thr` stops at the new iterate.
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int MH_SPLIT = 32;   // row blocks of the column pass

constexpr int MH_STATS = 0, MH_KNOPP = 1, MH_UNBAL = 2;
// stats[]: 0 max|U| 1 max|V| 2 max|U-Up| 3 max|Up| 4 max|V-Vp| 5 max|Vp| 6 nan in U 7 nan in V
//          8 sum (Vp * K^T Up - B)^2 (MH_KNOPP) 9 K^T u == 0 somewhere 10 inf in U 11 inf in V
//          12 status (0 running, 1 converged, 2 numerical guard) 13 iteration at which the solve ended
//          14 number of recorded errors 15 last error
constexpr int MH_ST_STATUS = 12, MH_ST_ITER = 13, MH_ST_NERR = 14, MH_ST_LAST = 15;

struct MhCtl {
  int mode;         // MH_STATS: no decisions, no gating (the host decides: sinkhorn_stabilized)
  int iter;         // index of this step
  int check_every;  // MH_KNOPP
  double stop_thr;
  double fi_u, fi_v;  // exponents of the u and v updates (1 = none)
  double* stats;      // 16 doubles (may be null for MH_STATS: no statistics)
  double* err_hist;   // MH_KNOPP / MH_UNBAL
  int err_cap;
};

__device__ __forceinline__ bool mh_ended(const MhCtl& c) {
  return c.mode != MH_STATS && __ldcg(&c.stats[MH_ST_STATUS]) != 0.0;
}

// HC = histograms per thread, EXACTLY the histograms [h_base + HC blockIdx.z, + HC) - no "h < H" test in the loops: with
// such a guard around every load + FMA pair the compiler emitted load, load, FMA, branch sequences (one load in flight
// per thread: 1.8 TB/s at 16384^2).  H > 8 runs as chunks of 8 (blockIdx.z, each re-reading K) plus one launch for the
// H % 8 rest.  Rows are taken MhTrip<T>::R at a time: the loads are issued before the first dependent FMA, the products
// of a trip are summed in T and the trip sum is added to the double accumulator.
template <typename T> struct MhTrip { static constexpr int R = 8; };
template <> struct MhTrip<double> { static constexpr int R = 4; };

constexpr int MH_TR = 128;  // rows of U staged per tile of the column pass

template <typename T, int HC>
__global__ void __launch_bounds__(256) mh_col_partial_kernel(const T* __restrict__ K, long long n, long long m,
                                                             long long ld, int H, int h_base,
                                                             const T* __restrict__ U, double* __restrict__ partial,
                                                             const MhCtl ctl) {
  constexpr int R = MhTrip<T>::R;
  // the U values of this block's rows, MH_TR rows at a time: HC operand loads per matrix element straight from global
  // memory made the pass load-issue bound (H = 8: 2.7 ms per step at 16384^2 against 0.49 ms for H = 1)
  __shared__ __align__(16) T Us[MH_TR * HC];
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const bool live = j < m;
  const int h0 = h_base + blockIdx.z * HC;
  if (mh_ended(ctl)) return;
  double acc[HC];
#pragma unroll
  for (int c = 0; c < HC; ++c) acc[c] = 0.0;
  const long long S = gridDim.y;
  const long long nrows = (n - (long long)blockIdx.y + S - 1) / S;  // rows blockIdx.y + t S, t < nrows
  const T* Kc = K + j;
  for (long long t0 = 0; t0 < nrows; t0 += MH_TR) {
    const int tl = (int)(nrows - t0 < MH_TR ? nrows - t0 : MH_TR);
    __syncthreads();  // the previous tile has been consumed
    for (int q = threadIdx.x; q < tl * HC; q += 256) {
      const int tt = q / HC, c = q - tt * HC;
      Us[q] = U[((long long)blockIdx.y + (t0 + tt) * S) * H + h0 + c];
    }
    __syncthreads();
    if (!live) continue;
    int tt = 0;
    for (; tt + R <= tl; tt += R) {
      T k[R], t[HC];
#pragma unroll
      for (int r = 0; r < R; ++r) k[r] = Kc[((long long)blockIdx.y + (t0 + tt + r) * S) * ld];
#pragma unroll
      for (int c = 0; c < HC; ++c) t[c] = (T)0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
#pragma unroll
        for (int c = 0; c < HC; ++c) t[c] += k[r] * Us[(tt + r) * HC + c];
      }
#pragma unroll
      for (int c = 0; c < HC; ++c) acc[c] += (double)t[c];
    }
    for (; tt < tl; ++tt) {
      const double k = (double)Kc[((long long)blockIdx.y + (t0 + tt) * S) * ld];
#pragma unroll
      for (int c = 0; c < HC; ++c) acc[c] += k * (double)Us[tt * HC + c];
    }
  }
  if (!live) return;
#pragma unroll
  for (int c = 0; c < HC; ++c) partial[((long long)blockIdx.y * H + h0 + c) * m + j] = acc[c];
}

// grid = (ceil(m / 256), H).  Besides the v update: per-block sums of (Vp * c - B)^2 (the column-marginal violation of
// the iterate the step started from, MH_KNOPP's error) and "c == 0" flags into blk[2][gridDim.y * gridDim.x].
template <typename T>
__global__ void __launch_bounds__(256) mh_col_finish_kernel(const double* __restrict__ partial, long long m, int H,
                                                            int S, const T* __restrict__ B, const T* __restrict__ Vp,
                                                            T* __restrict__ V, double* __restrict__ blk,
                                                            const MhCtl ctl) {
  __shared__ double sred[2][8];
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const int h = blockIdx.y;
  if (mh_ended(ctl)) return;
  double e2 = 0.0, cz = 0.0;
  if (j < m) {
    double s = 0.0;
    for (int q = 0; q < S; ++q) s += partial[((long long)q * H + h) * m + j];
    const T c = (T)s;
    T vn = B[j * H + h] / c;
    if (ctl.fi_v != 1.0) vn = pow_t(vn, (T)ctl.fi_v);
    V[j * H + h] = vn;
    if (c == (T)0) cz = 1.0;
    if (Vp != nullptr) {
      const double e = (double)Vp[j * H + h] * s - (double)B[j * H + h];
      e2 = e * e;
    }
  }
  if (blk == nullptr) return;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    e2 += __shfl_down_sync(0xffffffffu, e2, o);
    cz = fmax(cz, __shfl_down_sync(0xffffffffu, cz, o));
  }
  if ((threadIdx.x & 31) == 0) {
    sred[0][threadIdx.x >> 5] = e2;
    sred[1][threadIdx.x >> 5] = cz;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a2 = 0.0, az = 0.0;
    for (int w = 0; w < 8; ++w) {
      a2 += sred[0][w];
      az = fmax(az, sred[1][w]);
    }
    const long long nb = (long long)gridDim.x * gridDim.y, b = (long long)blockIdx.y * gridDim.x + blockIdx.x;
    blk[b] = a2;
    blk[nb + b] = az;
  }
}

// grid = (ceil(n / MH_ROWS), 1, chunks), MH_ROWS warps per block, one row per warp (every block streams all of V from
// L2 once: 16 rows per block keep that at HC / 16 of the K traffic).  V is staged through shared memory in tiles of
// MhTile<T>::J columns, transposed to [histogram][column] (+1 padding): a lane's operand reads are then conflict-free
// instead of HC global loads per element at a stride of H elements between lanes.
constexpr int MH_ROWS = 16;
template <typename T> struct MhTile { static constexpr int J = 1024; };
template <> struct MhTile<double> { static constexpr int J = 512; };

template <typename T, int HC>
__global__ void __launch_bounds__(MH_ROWS * 32) mh_row_kernel(const T* __restrict__ K, long long n, long long m, long long ld,
                                                     int H, int h_base, const T* __restrict__ V,
                                                     const T* __restrict__ a, T* __restrict__ U, const MhCtl ctl) {
  constexpr int R = MhTrip<T>::R;
  constexpr int TJ = MhTile<T>::J;
  __shared__ T Vs[HC][TJ + 1];
  const int lane = threadIdx.x & 31;
  const long long i = (long long)blockIdx.x * MH_ROWS + (threadIdx.x >> 5);
  const bool live = i < n;
  const int h0 = h_base + blockIdx.z * HC;
  if (mh_ended(ctl)) return;
  double acc[HC];
#pragma unroll
  for (int c = 0; c < HC; ++c) acc[c] = 0.0;
  const T* row = K + (live ? i : 0) * ld;
  for (long long j0 = 0; j0 < m; j0 += TJ) {
    const int tl = (int)(m - j0 < TJ ? m - j0 : TJ);
    __syncthreads();  // the previous tile has been consumed
    for (int q = threadIdx.x; q < tl * HC; q += MH_ROWS * 32) {
      const int jj = q / HC, c = q - jj * HC;
      Vs[c][jj] = V[(j0 + jj) * H + h0 + c];
    }
    __syncthreads();
    if (!live) continue;
    int jj = lane;
    for (; jj + 32 * (R - 1) < tl; jj += 32 * R) {  // R loads in flight per lane, trip sums in T (see the column pass)
      T k[R], t[HC];
#pragma unroll
      for (int r = 0; r < R; ++r) k[r] = row[j0 + jj + 32 * r];
#pragma unroll
      for (int c = 0; c < HC; ++c) t[c] = (T)0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
#pragma unroll
        for (int c = 0; c < HC; ++c) t[c] += k[r] * Vs[c][jj + 32 * r];
      }
#pragma unroll
      for (int c = 0; c < HC; ++c) acc[c] += (double)t[c];
    }
    for (; jj < tl; jj += 32) {
      const double k = (double)row[j0 + jj];
#pragma unroll
      for (int c = 0; c < HC; ++c) acc[c] += k * (double)Vs[c][jj];
    }
  }
  if (!live) return;
#pragma unroll
  for (int c = 0; c < HC; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[c] += __shfl_down_sync(0xffffffffu, acc[c], o);
  }
  if (lane == 0) {
    const T ai = a[i];
#pragma unroll
    for (int c = 0; c < HC; ++c) {
      T un = ai / (T)acc[c];
      if (ctl.fi_u != 1.0) un = pow_t(un, (T)ctl.fi_u);
      U[i * H + h0 + c] = un;
    }
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

// The statistics of a step (see the stats[] table above; the maxima skip nan entries, the host puts the nan back where
// numpy's max would propagate it) and, for MH_KNOPP / MH_UNBAL, the decisions.  One block.
template <typename T>
__global__ void __launch_bounds__(1024) mh_stats_kernel(const T* __restrict__ U, const T* __restrict__ Up,
                                                        long long nu, const T* __restrict__ V,
                                                        const T* __restrict__ Vp, long long nv,
                                                        const double* __restrict__ blk, long long nblk,
                                                        const MhCtl ctl) {
  double* stats = ctl.stats;
  if (mh_ended(ctl)) return;
  double mu = 0.0, mdu = 0.0, mup = 0.0, mv = 0.0, mdv = 0.0, mvp = 0.0, nanu = 0.0, nanv = 0.0, infu = 0.0, infv = 0.0;
  for (long long q = threadIdx.x; q < nu; q += blockDim.x) {
    const double x = (double)U[q], p = (double)Up[q];
    if (x != x) nanu = 1.0;
    if (isinf(x)) infu = 1.0;
    mu = fmax(mu, fabs(x));
    mup = fmax(mup, fabs(p));
    mdu = fmax(mdu, fabs((double)(T)(U[q] - Up[q])));
  }
  for (long long q = threadIdx.x; q < nv; q += blockDim.x) {
    const double x = (double)V[q], p = (double)Vp[q];
    if (x != x) nanv = 1.0;
    if (isinf(x)) infv = 1.0;
    mv = fmax(mv, fabs(x));
    mvp = fmax(mvp, fabs(p));
    mdv = fmax(mdv, fabs((double)(T)(V[q] - Vp[q])));
  }
  // block partials of the column pass: fixed order (strided per thread, then the warp / block trees)
  double e2 = 0.0, cz = 0.0;
  if (blk != nullptr) {
    for (long long q = threadIdx.x; q < nblk; q += blockDim.x) {
      e2 += blk[q];
      cz = fmax(cz, blk[nblk + q]);
    }
  }
  // all twelve reductions together: warp trees, one barrier, then warp w folds quantity w
  __shared__ double sq[12][33];
  double val[12] = {mu, mv, mdu, mup, mdv, mvp, nanu, nanv, infu, infv, cz, e2};
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    double x = val[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double y = __shfl_down_sync(0xffffffffu, x, o);
      x = (k == 11) ? x + y : fmax(x, y);
    }
    if ((threadIdx.x & 31) == 0) sq[k][threadIdx.x >> 5] = x;
  }
  __syncthreads();
  const int wq = threadIdx.x >> 5, lq = threadIdx.x & 31;
  if (wq < 12) {
    double x = lq < (int)(blockDim.x >> 5) ? sq[wq][lq] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double y = __shfl_down_sync(0xffffffffu, x, o);
      x = (wq == 11) ? x + y : fmax(x, y);
    }
    if (lq == 0) sq[wq][32] = x;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  mu = sq[0][32];
  mv = sq[1][32];
  mdu = sq[2][32];
  mup = sq[3][32];
  mdv = sq[4][32];
  mvp = sq[5][32];
  nanu = sq[6][32];
  nanv = sq[7][32];
  infu = sq[8][32];
  infv = sq[9][32];
  cz = sq[10][32];
  e2 = sq[11][32];
  stats[0] = mu;
  stats[1] = mv;
  stats[2] = mdu;
  stats[3] = mup;
  stats[4] = mdv;
  stats[5] = mvp;
  stats[6] = nanu;
  stats[7] = nanv;
  stats[8] = e2;
  stats[9] = cz;
  stats[10] = infu;
  stats[11] = infv;
  if (ctl.mode == MH_STATS) return;
  const bool bad = cz != 0.0 || nanu != 0.0 || nanv != 0.0 || infu != 0.0 || infv != 0.0;
  auto record = [&](double err) {
    const int k = (int)stats[MH_ST_NERR];
    if (k < ctl.err_cap) ctl.err_hist[k] = err;
    stats[MH_ST_NERR] = (double)(k + 1);
    stats[MH_ST_LAST] = err;
  };
  if (ctl.mode == MH_KNOPP) {
    // the column pass of this step measured the iterate of step iter - 1 (:649-653)
    if (ctl.iter >= 1 && (ctl.iter - 1) % ctl.check_every == 0) {
      const double err = sqrt(e2);
      record(err);
      if (err < ctl.stop_thr) {
        stats[MH_ST_STATUS] = 1.0;
        stats[MH_ST_ITER] = (double)(ctl.iter - 1);
        return;
      }
    }
    if (bad) {  // :633-645: previous iterate, stop
      stats[MH_ST_STATUS] = 2.0;
      stats[MH_ST_ITER] = (double)ctl.iter;
    }
  } else {  // MH_UNBAL
    if (bad) {  // ot/unbalanced/_sinkhorn.py:760-772
      stats[MH_ST_STATUS] = 2.0;
      stats[MH_ST_ITER] = (double)ctl.iter;
      return;
    }
    const double err = 0.5 * (mdu / fmax(fmax(mu, mup), 1.0) + mdv / fmax(fmax(mv, mvp), 1.0));  // :774-781
    record(err);
    if (err < ctl.stop_thr) {
      stats[MH_ST_STATUS] = 1.0;
      stats[MH_ST_ITER] = (double)ctl.iter;
    }
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
