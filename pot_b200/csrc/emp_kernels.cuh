// Online-cost (empirical) Sinkhorn kernels, sm_100a: the cost C_ij = max(|x_i|^2 + |y_j|^2 - 2 x_i.y_j, 0)
// (ot/utils.py:292-299) is rebuilt tile by tile from the point clouds and consumed in registers; the n x m
// cost matrix is never materialised (reference: block loop of ot/bregman/_empirical.py:218-233, which
// rebuilds cost blocks on the CPU).
//
// This header holds the generic register-tiled path (any d, any alignment, fp32 and fp64): a 64 x 64 output
// tile per step, 4 x 4 micro-tile per thread, with three epilogues on the finished tile:
//   EPI_LSE  : online (max, sum-exp) of  pot_j - C_ij / reg  along the row -> partial log-sum-exp per row
//   EPI_DIST : write C_ij                         (ot.dist, sqeuclidean / euclidean)
//   EPI_PLAN : write exp(f_i + g_j - C_ij / reg)  (lazy plan block, _empirical.py:61-67)
// The fp32 tensor-core path (tcgen05, 3xTF32) lives in emp_tc_kernels.cuh and shares the finalize kernels.
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int EPI_LSE = 0;
constexpr int EPI_DIST = 1;
constexpr int EPI_PLAN = 2;
constexpr int EPI_LOSS = 3;  // sum_ij C_ij exp(f_i + g_j - C_ij / reg) -> one double per block (empirical_sinkhorn2, :466-483)

constexpr int EMP_BM = 64, EMP_BN = 64, EMP_BK = 16, EMP_NT = 256;

struct EmpTileParams {
  const void* X;     // nx x d rows (leading dimension ldx)
  const void* Y;     // ny x d rows (leading dimension ldy)
  const void* xx;    // |x_i|^2 (nx)
  const void* yy;    // |y_j|^2 (ny)
  const void* potx;  // EPI_PLAN: f_i (nx)
  const void* poty;  // EPI_LSE: potential added along the reduced axis (ny) ; EPI_PLAN: g_j
  void* out;         // EPI_DIST / EPI_PLAN: nx x ny (leading dimension ldo)
  void* pmax;        // EPI_LSE: [nsplit][nx]
  void* psum;        // EPI_LSE: [nsplit][nx]
  double* blk_out;   // EPI_LOSS: one partial per block
  long long nx, ny, ldx, ldy, ldo;
  int d;
  int nsplit;        // EPI_LSE / EPI_LOSS: the ny range is cut into nsplit chunks (blockIdx.y)
  int metric;        // 0 sqeuclidean, 1 euclidean
  double reg;
  const int* status; // optional sticky status word: no-op unless *status == expect_status
  int expect_status;
  const int* pot_sel;      // optional: poty is a ring, slot *pot_sel (stride pot_stride elements) is the live one
  long long pot_stride;
};

template <typename T, int EPI>
__global__ void __launch_bounds__(EMP_NT) emp_tile_kernel(const EmpTileParams p) {
  __shared__ T Xs[EMP_BK][EMP_BM + 1];
  __shared__ T Ys[EMP_BK][EMP_BN + 1];
  if (p.status != nullptr && *p.status != p.expect_status) return;

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long row0 = (long long)blockIdx.x * EMP_BM;
  const T* X = reinterpret_cast<const T*>(p.X);
  const T* Y = reinterpret_cast<const T*>(p.Y);
  const T* xxg = reinterpret_cast<const T*>(p.xx);
  const T* yyg = reinterpret_cast<const T*>(p.yy);
  const T reg = (T)p.reg;
  const T* potyg = reinterpret_cast<const T*>(p.poty);
  if (p.pot_sel != nullptr) potyg += (long long)(*p.pot_sel) * p.pot_stride;

  // column range of this block
  long long c_begin = 0, c_end = p.ny;
  if (EPI == EPI_LSE || EPI == EPI_LOSS) {
    const long long tiles = (p.ny + EMP_BN - 1) / EMP_BN;
    const long long per = (tiles + p.nsplit - 1) / p.nsplit;
    c_begin = (long long)blockIdx.y * per * EMP_BN;
    c_end = c_begin + per * EMP_BN;
    if (c_end > p.ny) c_end = p.ny;
  } else {
    c_begin = (long long)blockIdx.y * EMP_BN;
    c_end = c_begin + EMP_BN;
    if (c_end > p.ny) c_end = p.ny;
  }

  T xr[4], fr[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const long long i = row0 + ty * 4 + r;
    xr[r] = (i < p.nx) ? xxg[i] : (T)0;
    fr[r] = ((EPI == EPI_PLAN || EPI == EPI_LOSS) && i < p.nx) ? reinterpret_cast<const T*>(p.potx)[i] : (T)0;
  }
  T rmax[4], rsum[4];
  double loss = 0.0;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    rmax[r] = neg_inf<T>();
    rsum[r] = (T)0;
  }

  for (long long col0 = c_begin; col0 < c_end; col0 += EMP_BN) {
    T acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[r][c] = (T)0;

    for (int k0 = 0; k0 < p.d; k0 += EMP_BK) {
      __syncthreads();
#pragma unroll
      for (int q = 0; q < (EMP_BM * EMP_BK) / EMP_NT; ++q) {
        const int e = tid + q * EMP_NT;
        const int k = e % EMP_BK, i = e / EMP_BK;
        const long long gi = row0 + i, gj = col0 + i;
        Xs[k][i] = (gi < p.nx && k0 + k < p.d) ? X[gi * p.ldx + k0 + k] : (T)0;
        Ys[k][i] = (gj < p.ny && k0 + k < p.d) ? Y[gj * p.ldy + k0 + k] : (T)0;
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < EMP_BK; ++k) {
        T xv[4], yv[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) xv[r] = Xs[k][ty * 4 + r];
#pragma unroll
        for (int c = 0; c < 4; ++c) yv[c] = Ys[k][tx + 16 * c];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[r][c] = fma(xv[r], yv[c], acc[r][c]);
      }
    }

    // ---- epilogue on the finished 4 x 4 micro-tile ----
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const long long j = col0 + tx + 16 * c;
      const bool jok = j < p.ny;
      const T yj = jok ? yyg[j] : (T)0;
      T pj = (T)0;
      if (EPI != EPI_DIST && jok) pj = potyg[j];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        // reference order: C = -2 x.y ; C += |x|^2 ; C += |y|^2 ; max(C, 0)   (ot/utils.py:292-299)
        T cst = max_t(((T)(-2) * acc[r][c] + xr[r]) + yj, (T)0);
        if (p.metric == 1) cst = sqrt(cst);
        if (EPI == EPI_DIST) {
          acc[r][c] = cst;
        } else if (EPI == EPI_PLAN) {
          acc[r][c] = exp_t((fr[r] + pj) - cst / reg);
        } else if (EPI == EPI_LOSS) {
          if (jok && row0 + ty * 4 + r < p.nx) loss += (double)(cst * exp_t((fr[r] + pj) - cst / reg));
        } else {
          acc[r][c] = jok ? (pj - cst / reg) : neg_inf<T>();
        }
      }
    }
    if (EPI == EPI_LSE) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const T m4 = max_t(max_t(acc[r][0], acc[r][1]), max_t(acc[r][2], acc[r][3]));
        if (m4 == neg_inf<T>()) continue;
        if (m4 > rmax[r]) {
          rsum[r] *= exp_t(rmax[r] - m4);  // exp(-inf) = 0 on the first hit
          rmax[r] = m4;
        }
        const T mref = rmax[r];
        rsum[r] += (exp_t(acc[r][0] - mref) + exp_t(acc[r][1] - mref)) +
                   (exp_t(acc[r][2] - mref) + exp_t(acc[r][3] - mref));
      }
    } else if (EPI != EPI_LOSS) {
      T* out = reinterpret_cast<T*>(p.out);
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const long long i = row0 + ty * 4 + r;
        if (i >= p.nx) continue;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const long long j = col0 + tx + 16 * c;
          if (j < p.ny) out[i * p.ldo + j] = acc[r][c];
        }
      }
    }
  }

  if (EPI == EPI_LOSS) {
    __shared__ double lred[EMP_NT];
    lred[tid] = loss;
    __syncthreads();
    for (int o = EMP_NT / 2; o > 0; o >>= 1) {
      if (tid < o) lred[tid] += lred[tid + o];
      __syncthreads();
    }
    if (tid == 0) p.blk_out[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = lred[0];
  }
  if (EPI == EPI_LSE) {
    // combine the 16 threads (tx) that share a row: they are 16 consecutive lanes
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const T om = __shfl_xor_sync(0xffffffffu, rmax[r], o);
        const T os = __shfl_xor_sync(0xffffffffu, rsum[r], o);
        const T nm = max_t(rmax[r], om);
        T s = (T)0;
        if (rmax[r] != neg_inf<T>()) s += rsum[r] * exp_t(rmax[r] - nm);
        if (om != neg_inf<T>()) s += os * exp_t(om - nm);
        rmax[r] = nm;
        rsum[r] = s;
      }
    }
    if (tx == 0) {
      T* pm = reinterpret_cast<T*>(p.pmax) + (long long)blockIdx.y * p.nx;
      T* ps = reinterpret_cast<T*>(p.psum) + (long long)blockIdx.y * p.nx;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const long long i = row0 + ty * 4 + r;
        if (i < p.nx) {
          pm[i] = rmax[r];
          ps[i] = rsum[r];
        }
      }
    }
  }
}

__global__ void emp_sum_blocks_kernel(const double* blk, int nblk, double* out) {
  __shared__ double sred[256];
  double x = 0.0;
  for (int i = threadIdx.x; i < nblk; i += 256) x += blk[i];
  sred[threadIdx.x] = x;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) sred[threadIdx.x] += sred[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = sred[0];
}

// |x_i|^2, one warp per row (ot/utils.py:292-293: einsum("ij,ij->i", X, X))
template <typename T>
__global__ void __launch_bounds__(256) sqnorm_kernel(const T* __restrict__ X, long long n, int d, long long ld,
                                                     T* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  T s = (T)0;
  for (int k = lane; k < d; k += 32) {
    const T x = X[i * ld + k];
    s = fma(x, x, s);
  }
  s = warp_sum(s);
  if (lane == 0) out[i] = s;
}

// ------------------------------------------------------------------------------------------------
// Solver state + finalize kernels of the online-cost solve (f then g, ot/bregman/_empirical.py:218-256).
//
// Iteration `it`:  A  row log-sum-exp partials of (g_j - C_ij/reg)           -> emp_fin_f
//                  B  column log-sum-exp partials of (f_i - C_ij/reg)        -> emp_fin_g
// The reference's check after iteration it ((it + 1) % 10 == 0) is  err = sum_i |sum_j exp(f_i + g_j - C_ij/reg) - a_i|
// = sum_i |exp(f_i + rowLSE_i(g)) - a_i|, and rowLSE(g) is exactly what pass A of iteration it + 1 produces:
// the check costs nothing extra and is evaluated by emp_fin_f BEFORE it commits the next f (so that a
// converged solve returns the (f, g) of iteration it, like the reference's `break`).
// ------------------------------------------------------------------------------------------------
struct __align__(16) EmpState {
  int status;
  int iter;      // index of the iteration whose pass A runs next
  int niter;     // reference's dict_log["niter"]
  int sel_f;     // slot of the current f in the 2-ring
  int n_err;
  unsigned int ticket;
  unsigned int ticket_g;
  int pad;
  double last_err;
  double err_local;  // this rank's part of the current check (row-sharded runs all-reduce it)
};

struct EmpFinParams {
  const void* pmax;   // [nsplit][len]
  const void* psum;
  int nsplit;
  long long len;      // rows (fin_f) or columns (fin_g) handled here
  const void* logw;   // log a (local rows) or log b
  const void* w;      // a (local rows), fin_f only
  void* fring;        // 2 x n_local
  void* g;            // m
  void* cbuf;         // fin_g multi-rank staging: 3 x m
  EmpState* st;
  double* blk_err;
  double* err_hist;
  double* err_xchg;   // one double that the all-reduce hook sums across ranks (null: single rank)
  int max_iter;
  double stop_thr;
  int stage;
};

template <typename T>
__device__ __forceinline__ void lse_of_partials(const T* pm, const T* ps, int nsplit, long long len, long long i,
                                                T& M, T& S) {
  M = neg_inf<T>();
  for (int q = 0; q < nsplit; ++q) M = max_t(M, pm[(long long)q * len + i]);
  S = (T)0;
  for (int q = 0; q < nsplit; ++q) {
    const T mq = pm[(long long)q * len + i];
    if (mq != neg_inf<T>()) S += ps[(long long)q * len + i] * exp_t(mq - M);
  }
}

__device__ __forceinline__ void emp_decide(EmpState* st, double err, double* err_hist, int max_iter, double stop_thr) {
  const int it = st->iter;
  const bool check_due = it > 0 && (it % 10) == 0;
  if (check_due) {
    err_hist[st->n_err] = err;
    st->n_err += 1;
    st->last_err = err;
    if (err <= stop_thr) {  // reference: `if err <= stopThr: break` (_empirical.py:255)
      st->status = ST_CONVERGED;
      st->niter = it - 1;
      return;
    }
  }
  if (it >= max_iter) {
    st->status = ST_MAXITER;
    st->niter = max_iter - 1;
    return;
  }
  st->sel_f ^= 1;  // commit f_it
}

// f_new = log a - rowLSE(g); also this rank's part of the pending marginal check.
template <typename T>
__global__ void __launch_bounds__(256) emp_fin_f_kernel(const EmpFinParams p) {
  __shared__ double sred[256];
  __shared__ int s_last;
  EmpState* st = p.st;
  if (st->status != ST_RUNNING) return;
  const int tid = threadIdx.x;
  const int it = st->iter;
  const bool check_due = it > 0 && (it % 10) == 0;
  const int sel = st->sel_f;
  const T* fold = reinterpret_cast<const T*>(p.fring) + (long long)sel * p.len;
  T* fnew = reinterpret_cast<T*>(p.fring) + (long long)(sel ^ 1) * p.len;
  double e = 0.0;
  for (long long i = (long long)blockIdx.x * 256 + tid; i < p.len; i += (long long)gridDim.x * 256) {
    T M, S;
    lse_of_partials<T>(reinterpret_cast<const T*>(p.pmax), reinterpret_cast<const T*>(p.psum), p.nsplit, p.len, i, M, S);
    const T r = M + log_t(S);
    fnew[i] = reinterpret_cast<const T*>(p.logw)[i] - r;
    if (check_due) {
      const T m1 = exp_t(fold[i] + r);
      const T dlt = m1 - reinterpret_cast<const T*>(p.w)[i];
      e += (double)(dlt < (T)0 ? -dlt : dlt);
    }
  }
  sred[tid] = e;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) sred[tid] += sred[tid + o];
    __syncthreads();
  }
  if (tid == 0) {
    p.blk_err[blockIdx.x] = sred[0];
    __threadfence();
    const unsigned int t = atomicAdd(&st->ticket, 1u);
    s_last = (t == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double x = 0.0;
  for (unsigned int i = tid; i < gridDim.x; i += 256) x += __ldcg(&p.blk_err[i]);
  sred[tid] = x;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) sred[tid] += sred[tid + o];
    __syncthreads();
  }
  if (tid != 0) return;
  st->ticket = 0;
  if (p.err_xchg != nullptr) {
    *p.err_xchg = sred[0];  // summed across ranks by the hook, then emp_decide_kernel
  } else {
    emp_decide(st, sred[0], p.err_hist, p.max_iter, p.stop_thr);
  }
  __threadfence();
}

__global__ void emp_decide_kernel(EmpState* st, const double* err_xchg, double* err_hist, int max_iter,
                                  double stop_thr) {
  if (st->status != ST_RUNNING) return;
  emp_decide(st, *err_xchg, err_hist, max_iter, stop_thr);
}

// g = log b - colLSE(f).  stage 0: single rank, everything; stages 1-3: row-sharded (see emp.cu).
template <typename T>
__global__ void __launch_bounds__(256) emp_fin_g_kernel(const EmpFinParams p) {
  __shared__ int s_last;
  EmpState* st = p.st;
  if (st->status != ST_RUNNING) return;
  const int tid = threadIdx.x;
  T* g = reinterpret_cast<T*>(p.g);
  T* cb = reinterpret_cast<T*>(p.cbuf);
  const long long m = p.len;
  for (long long j = (long long)blockIdx.x * 256 + tid; j < m; j += (long long)gridDim.x * 256) {
    if (p.stage == 0 || p.stage == 1) {
      T M, S;
      lse_of_partials<T>(reinterpret_cast<const T*>(p.pmax), reinterpret_cast<const T*>(p.psum), p.nsplit, m, j, M, S);
      if (p.stage == 0) {
        g[j] = reinterpret_cast<const T*>(p.logw)[j] - (M + log_t(S));
      } else {
        cb[j] = M;          // local maximum
        cb[m + j] = M;      // all-reduced (max) in place
        cb[2 * m + j] = S;  // local sum
      }
    } else if (p.stage == 2) {
      const T Ml = cb[j], Mg = cb[m + j];
      cb[2 * m + j] = (Ml == neg_inf<T>()) ? (T)0 : cb[2 * m + j] * exp_t(Ml - Mg);  // all-reduced (sum) in place
    } else {
      g[j] = reinterpret_cast<const T*>(p.logw)[j] - (cb[m + j] + log_t(cb[2 * m + j]));
    }
  }
  if (p.stage == 1 || p.stage == 2) return;
  // the iteration is complete once every block has written its g
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const unsigned int t = atomicAdd(&st->ticket_g, 1u);
    s_last = (t == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_last || tid != 0) return;
  st->ticket_g = 0;
  const int it = st->iter + 1;
  st->iter = it;
  if (it >= p.max_iter && (p.max_iter % 10) != 0) {  // no check pending for the last iteration
    st->status = ST_MAXITER;
    st->niter = p.max_iter - 1;
  }
  __threadfence();
}

template <typename T>
__global__ void emp_export_kernel(const void* fring, const void* g, long long n, long long m, const EmpState* st,
                                  void* f_out, void* g_out) {
  const T* f = reinterpret_cast<const T*>(fring) + (long long)st->sel_f * n;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (f_out != nullptr && i < n) reinterpret_cast<T*>(f_out)[i] = f[i];
  if (g_out != nullptr && i < m) reinterpret_cast<T*>(g_out)[i] = reinterpret_cast<const T*>(g)[i];
}

}  // namespace b2ot
