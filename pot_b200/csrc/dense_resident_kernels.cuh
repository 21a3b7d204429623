// Resident solver for SMALL dense problems (scaling domain and log domain), sm_100a.
//
// BASELINE config 1 (ot.sinkhorn, 1000 x 1000 fp64, 1000 iterations) is 8 MB: the streaming solver runs it as two
// dependent launches per iteration (fused sweep + finalize, 12.1 us per iteration, of which the L2-resident data
// movement is ~1.6 us).  Here ONE cooperative launch runs a whole chunk of iterations with K held in the shared memory of
// the grid - twice: CTA g keeps its block of ROWS [g R, (g + 1) R) (all columns) and its block of COLUMNS
// [g Q, (g + 1) Q) (all rows, stored transposed).  An iteration is then two vector all-gathers and nothing else:
//
//   phase A : u_i = a_i / (K v)_i for the CTA's rows (one warp per row, fixed lane-strided order + shuffle tree);
//   gather  : every CTA reads the whole u;
//   phase B : c_j = (K^T u)_j for the CTA's columns (one warp per column), and on them exactly finalize_kernel's
//             column step: marginal error term (v_j c_j - b_j)^2, v_j = b_j / c_j;
//   gather  : whole v (and, on a check iteration, the G error partials, summed in a fixed order) - then every CTA
//             applies finalize_kernel's state machine to its private copy of the solver state: all CTAs decide from the
//             same words, CTA 0 writes the state back at the end.  The guard flags need no exchange: "u not finite"
//             and "K^T u = 0 or v not finite" are read off the gathered vectors by every CTA.
//
// The gathers carry their own synchronisation: a value travels as 8-byte words {32 data bits, epoch} (one word for fp32,
// two for fp64; an aligned 8-byte store is single-copy atomic), written once by the producer and polled by the
// consumers with relaxed gpu-scope loads until the epoch matches - no flags, no fences, no atomics, one L2 write and one L2
// read of latency per gather.  Measured history on config 1 (profiles/r02_resident_experiments.md): two counter
// barriers per iteration (atomicAdd + spin by 143 CTAs) 10.0 us per iteration; flag words with st.release /
// ld.acquire 17.4 us (every gpu-scope acquire or fence costs an L1 invalidate); this form 7.0 us (fp64) / 6.0 us
// (fp32), of which the two gathers are 1.9 us each - the price of one producer -> L2 -> consumer hand-off among
// 148 CTAs - and the two block-wide passes 1.3 us each; the size of the problem hardly matters (400 x 400: 6.6 us).
// Results are bit-reproducible and do not depend on how iterate() calls chunk the solve.  Same update order
// (v-then-u, ot/bregman/_sinkhorn.py:629-631), error ||v * K^T u - b||_2 every check_every iterations (:646-663) and
// guard with the revert to the previous (u, v) (:633-645) as the streaming solver, so init / poll / export of the
// handle are unchanged (the plain u / v rings are written alongside).
// Log domain (MODE_LOG, ot/bregman/_sinkhorn.py:880-906): the same two phases with exact log-sum-exps (block_lses) in place
// of the dot products: u = log a - LSE_j(Mr + v) on the rows, L_j = LSE_i(Mr + u) on the columns, error term
// (exp(L_j + v_j) - b_j)^2 = the column marginal of exp(Mr + u + v), next v = log b - L_j; no guard, like the reference.
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int RES_THREADS = 256;
constexpr int RES_MAXG = 256;     // CTAs at most (>= SM count)

#ifdef B2OT_DIAG
#define RES_TRACE(slot) do { if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && it < 64) p.trace[it * 8 + (slot)] = clock64(); } while (0)
#else
#define RES_TRACE(slot) do { } while (0)
#endif

struct ResParams {
  long long* trace;  // diagnosis build: clock64 stamps of CTA 0, [iteration][8]
  const void* K;
  long long ld, n, m, mp;
  const void* a;     // KNOPP: a, LOG: log a
  const void* b;
  const void* logb;  // LOG only
  void* uring;       // 2 x n
  void* vring;       // 3 x mp
  void* cbuf;        // mp
  void* xbuf;        // exchange words, `replicas` copies of: u (n values), v (m values), error partials (G doubles);
                     // zeroed by init
  int replicas;      // every producer writes all copies, consumer CTA g reads copy g % replicas
  SolveState* st;
  double* err_hist;
  int max_iter, check_every;
  double stop_thr;
  int n_iters;       // iterations of this launch
  int rows_per_cta;  // R
  int cols_per_cta;  // Q
};

template <typename T>
__device__ __forceinline__ T warp_sum_down(T x) {  // lane 0 holds the total; one fixed association order
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
  return x;
}

// ---- exchange words: {32 data bits, epoch}, one aligned 8-byte store each; fp64 values are two words side by side ----
__device__ __forceinline__ void xw_store(unsigned long long* p, unsigned int bits, unsigned int epoch) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(bits), "r"(epoch) : "memory");
}
template <typename T> struct XWords;
template <> struct XWords<float> {
  static constexpr int N = 1;
  static __device__ __forceinline__ void put(unsigned long long* base, long long i, float x, unsigned int epoch) {
    xw_store(base + i, __float_as_uint(x), epoch);
  }
  // one poll: true when the word carries `epoch`
  static __device__ __forceinline__ bool poll(const unsigned long long* base, long long i, unsigned int epoch, float& out) {
    unsigned int bits, e;
    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(bits), "=r"(e) : "l"(base + i) : "memory");
    out = __uint_as_float(bits);
    return e == epoch;
  }
};
template <> struct XWords<double> {
  static constexpr int N = 2;
  static __device__ __forceinline__ void put(unsigned long long* base, long long i, double x, unsigned int epoch) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    xw_store(base + 2 * i, (unsigned int)b, epoch);
    xw_store(base + 2 * i + 1, (unsigned int)(b >> 32), epoch);
  }
  // both halves in one 16-byte load (each half is its own atomic word: both epochs are checked)
  static __device__ __forceinline__ bool poll(const unsigned long long* base, long long i, unsigned int epoch, double& out) {
    unsigned int lo, e0, hi, e1;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(lo), "=r"(e0), "=r"(hi), "=r"(e1)
                 : "l"(base + 2 * i)
                 : "memory");
    out = __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
    return e0 == epoch && e1 == epoch;
  }
};
// all-gather, consumer side: `count` values (producer CTA q owns [q S, (q + 1) S)) into shared memory.  First one thread
// per producer waits for the LAST value of that producer's slice (one polled word per producer and consumer, with a
// short back-off: letting every thread poll its own values made 150 k polls per round hammer the same 128 L2 lines and
// the producers' stores queued behind them, 3.8 us per gather); then every thread reads its values, four in flight, and
// re-reads any whose epoch is not there yet (values of one slice come from different warps).  Returns whether this thread
// saw a non-finite value.
template <typename T>
__device__ __forceinline__ int res_gather(const unsigned long long* base, long long count, long long S, unsigned int G,
                                          unsigned int epoch, T* dst) {
  for (unsigned int q = threadIdx.x; q < G; q += RES_THREADS) {
    const long long lo = (long long)q * S;
    if (lo >= count) break;
    const long long last = (lo + S < count ? lo + S : count) - 1;
    T tmp;
    while (!XWords<T>::poll(base, last, epoch, tmp)) __nanosleep(20);
  }
  __syncthreads();
  int bad = 0;
  for (long long i0 = threadIdx.x; i0 < count; i0 += 4 * RES_THREADS) {
    T x0 = (T)0, x1 = (T)0, x2 = (T)0, x3 = (T)0;
    const long long i1 = i0 + RES_THREADS, i2 = i0 + 2 * RES_THREADS, i3 = i0 + 3 * RES_THREADS;
    bool ok;
    do {
      ok = XWords<T>::poll(base, i0, epoch, x0);
      if (i1 < count) ok &= XWords<T>::poll(base, i1, epoch, x1);
      if (i2 < count) ok &= XWords<T>::poll(base, i2, epoch, x2);
      if (i3 < count) ok &= XWords<T>::poll(base, i3, epoch, x3);
    } while (!ok);
    dst[i0] = x0;
    bad |= !finite_t(x0);
    if (i1 < count) { dst[i1] = x1; bad |= !finite_t(x1); }
    if (i2 < count) { dst[i2] = x2; bad |= !finite_t(x2); }
    if (i3 < count) { dst[i3] = x3; bad |= !finite_t(x3); }
  }
  return bad;
}
// `nrows` dot products mat[r][:] . vec (shared memory, row length `len`) by the whole block: a thread owns the columns
// j = tid, tid + 256, ... of up to eight rows at a time, shuffle tree per warp, the eight warp partials are added in a
// fixed order.  out[r] (shared memory) is valid after the call.  (One warp per row left seven of eight warps idle on
// a 1000-long serial chain: 2200 clk per pass at n = m = 1000.)
template <typename T>
__device__ __forceinline__ void block_dots(const T* mat, const T* vec, int nrows, long long len, T* red, T* out) {
  constexpr int RB = 8;
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  for (int i0 = 0; i0 < nrows; i0 += RB) {
    T acc[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = (T)0;
    const int nb = nrows - i0 < RB ? nrows - i0 : RB;
    if (nb == RB) {
      for (long long j = tid; j < len; j += RES_THREADS) {
        const T x = vec[j];
#pragma unroll
        for (int r = 0; r < RB; ++r) acc[r] += mat[(size_t)(i0 + r) * len + j] * x;
      }
    } else {
      for (long long j = tid; j < len; j += RES_THREADS) {
        const T x = vec[j];
#pragma unroll
        for (int r = 0; r < RB; ++r)
          if (r < nb) acc[r] += mat[(size_t)(i0 + r) * len + j] * x;
      }
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) acc[r] = warp_sum_down(acc[r]);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < RB; ++r) red[wrp * RB + r] = acc[r];
    }
    __syncthreads();
    if (tid < nb) {
      T s = (T)0;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += red[w * RB + tid];
      out[i0 + tid] = s;
    }
    __syncthreads();
  }
}
// `nrows` log-sum-exps  out[r] = log sum_j exp(mat[r][j] + vec[j])  by the whole block (log domain): exact row maxima
// first (no exponential), then sum exp(x - max) - one exponential per element, fixed summation order.  A row of -inf
// gives -inf.  rmx: 8 values of shared memory.
template <typename T>
__device__ __forceinline__ void block_lses(const T* mat, const T* vec, int nrows, long long len, T* red, T* rmx, T* out) {
  constexpr int RB = 8;
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  for (int i0 = 0; i0 < nrows; i0 += RB) {
    const int nb = nrows - i0 < RB ? nrows - i0 : RB;
    T mx[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) mx[r] = neg_inf<T>();
    for (long long j = tid; j < len; j += RES_THREADS) {
      const T x = vec[j];
#pragma unroll
      for (int r = 0; r < RB; ++r)
        if (r < nb) mx[r] = max_t(mx[r], mat[(size_t)(i0 + r) * len + j] + x);
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx[r] = max_t(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], o));
    }
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < RB; ++r) red[wrp * RB + r] = mx[r];
    }
    __syncthreads();
    if (tid < RB) {
      T M = red[tid];
#pragma unroll
      for (int w = 1; w < 8; ++w) M = max_t(M, red[w * RB + tid]);
      rmx[tid] = M;
    }
    __syncthreads();
    T ref[RB], sm[RB];
#pragma unroll
    for (int r = 0; r < RB; ++r) {
      const T M = rmx[r];
      ref[r] = (M == neg_inf<T>()) ? (T)0 : M;
      sm[r] = (T)0;
    }
    for (long long j = tid; j < len; j += RES_THREADS) {
      const T x = vec[j];
#pragma unroll
      for (int r = 0; r < RB; ++r)
        if (r < nb) sm[r] += exp_t(mat[(size_t)(i0 + r) * len + j] + x - ref[r]);
    }
#pragma unroll
    for (int r = 0; r < RB; ++r) sm[r] = warp_sum_down(sm[r]);
    if (lane == 0) {
#pragma unroll
      for (int r = 0; r < RB; ++r) red[wrp * RB + r] = sm[r];
    }
    __syncthreads();
    if (tid < nb) {
      T t = (T)0;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w * RB + tid];
      const T M = rmx[tid];
      out[i0 + tid] = ((M == neg_inf<T>()) ? (T)0 : M) + log_t(t);
    }
    __syncthreads();
  }
}
// words of ONE copy of the exchange area (multiple of 16: copies start on 128-byte lines)
__host__ __device__ inline size_t res_xcopy_words(long long n, long long m, int elem_words) {
  return (((size_t)(n + m) * elem_words + 2 + 2 * (size_t)RES_MAXG) + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t res_xbuf_bytes(long long n, long long m, int elem_words, int replicas) {
  return res_xcopy_words(n, m, elem_words) * (size_t)replicas * sizeof(unsigned long long);
}


template <typename T, int MODE>
__global__ void __launch_bounds__(RES_THREADS, 1) resident_solve_kernel(const ResParams p) {
  extern __shared__ __align__(16) unsigned char res_smem[];
  const int tid = threadIdx.x, lane = tid & 31, wrp = tid >> 5;
  const unsigned int G = gridDim.x;
  const int cta = blockIdx.x;
  const long long n = p.n, m = p.m;
  const int R = p.rows_per_cta, Q = p.cols_per_cta;
  const long long r0 = (long long)cta * R;
  const int nr = (int)((r0 + R <= n) ? R : (n > r0 ? n - r0 : 0));
  const long long j0 = (long long)cta * Q;
  const int nq = (int)((j0 + Q <= m) ? Q : (m > j0 ? m - j0 : 0));

  T* Ks = reinterpret_cast<T*>(res_smem);   // R x m   : rows of this CTA
  T* Kt = Ks + (size_t)R * m;               // Q x n   : columns of this CTA, transposed
  T* vs = Kt + (size_t)Q * n;               // m       : current v
  T* us = vs + m;                           // n       : u of this iteration
  T* vn = us + n;                           // m       : v being produced
  T* as = vn + m;                           // R       : a on the CTA's rows
  T* bs = as + R;                           // Q       : b on the CTA's columns
  T* red = bs + Q;                          // 64      : warp partials of block_dots / block_lses
  T* rmx = red + 64;                        // 8       : row maxima of block_lses
  T* dots = rmx + 8;                        // max(R, Q)
  T* lbs = dots + (R > Q ? R : Q);          // Q       : log b on the CTA's columns (log domain)
  __shared__ double sred[8];

  const T* Kg = reinterpret_cast<const T*>(p.K);
  const T* a = reinterpret_cast<const T*>(p.a);
  const T* b = reinterpret_cast<const T*>(p.b);
  T* uring = reinterpret_cast<T*>(p.uring);
  T* vring = reinterpret_cast<T*>(p.vring);
  T* cbuf = reinterpret_cast<T*>(p.cbuf);
  SolveState* st = p.st;
  // Exchange area.  All G CTAs read every value, and 148 readers of one 128-byte line serialise in its L2 slice (a
  // gather cost 3.5 us with a single copy): every producer writes `replicas` copies, CTA g reads copy g % replicas.
  const size_t xstride = res_xcopy_words(n, m, XWords<T>::N);
  const int nrep = p.replicas;
  unsigned long long* xall = reinterpret_cast<unsigned long long*>(p.xbuf);
  const size_t off_v = (size_t)n * XWords<T>::N;
  const size_t off_e = ((size_t)(n + m) * XWords<T>::N + 1) & ~(size_t)1;   // 2 words per CTA (a double), 16-byte aligned
  const unsigned long long* xu = xall + (size_t)(cta % nrep) * xstride;     // the copy this CTA reads
  const unsigned long long* xv = xu + off_v;
  const unsigned long long* xe = xu + off_e;

  // private copy of the solver state (identical in every CTA: same reads, same decisions)
  int status = st->status, iter = st->iter, niter = st->niter, sel_u = st->sel_u, sel_v = st->sel_v;
  int cur_flags = st->cur_flags, n_err = st->n_err;
  double last_err = st->last_err;
  unsigned int epoch = st->ticket;   // exchanges completed by earlier launches of this solve (0 after init)

  if (status == ST_RUNNING && p.n_iters > 0) {
    // K blocks of this CTA -> shared memory, once per launch; v of the first iteration
    for (int i = 0; i < nr; ++i)
      for (long long j = tid; j < m; j += RES_THREADS) Ks[(size_t)i * m + j] = Kg[(r0 + i) * p.ld + j];
    for (long long i = wrp; i < n; i += 8)
      for (int q = lane; q < nq; q += 32) Kt[(size_t)q * n + i] = Kg[i * p.ld + j0 + q];
    const T* v0 = vring + (long long)sel_v * p.mp;
    for (long long j = tid; j < m; j += RES_THREADS) vs[j] = __ldcg(&v0[j]);
    for (int i = tid; i < nr; i += RES_THREADS) as[i] = a[r0 + i];
    for (int q = tid; q < nq; q += RES_THREADS) {
      bs[q] = b[j0 + q];
      if (MODE == MODE_LOG) lbs[q] = reinterpret_cast<const T*>(p.logb)[j0 + q];
    }
  }
  __syncthreads();

  for (int it = 0; it < p.n_iters && status == ST_RUNNING; ++it) {
    // ---------------- phase A: u = a / (K v) on the CTA's rows; all-gather of u ----------------
    T* unew = uring + (long long)(sel_u ^ 1) * n;
    epoch += 1;
    RES_TRACE(0);
    if (MODE == MODE_LOG)
      block_lses(Ks, vs, nr, m, red, rmx, dots);   // u = log a - LSE_j(Mr + v)   (_sinkhorn.py:882)
    else
      block_dots(Ks, vs, nr, m, red, dots);
    for (int i = tid; i < nr; i += RES_THREADS) {
      const T u = (MODE == MODE_LOG) ? as[i] - dots[i] : as[i] / dots[i];
      for (int r = 0; r < nrep; ++r) XWords<T>::put(xall + (size_t)r * xstride, r0 + i, u, epoch);
      unew[r0 + i] = u;
    }
    RES_TRACE(1);
    int uf = __syncthreads_or(res_gather<T>(xu, n, R, G, epoch, us));
    if (MODE == MODE_LOG) uf = 0;  // -inf is a legitimate log-scaling; sinkhorn_log has no numerical guard
    RES_TRACE(2);

    // ---------------- phase B: c = K^T u on the CTA's columns, v = b / c, error term; all-gather of v ----------------
    T* vnext = vring + (long long)((sel_v + 1) % 3) * p.mp;
    const int ii = iter;
    const bool check = (ii % p.check_every) == 0;
    epoch += 1;
    if (MODE == MODE_LOG)
      block_lses(Kt, us, nq, n, red, rmx, dots);   // L_j = LSE_i(Mr + u)
    else
      block_dots(Kt, us, nq, n, red, dots);
    double e2 = 0.0;
    for (int q = tid; q < nq; q += RES_THREADS) {
      const long long j = j0 + q;
      const T bj = bs[q];
      T c, vnew;
      if (MODE == MODE_LOG) {
        c = exp_t(dots[q] + vs[j]);    // column marginal of exp(Mr + u + v) with the new u (:884-889)
        vnew = lbs[q] - dots[q];       // v of the next iteration = log b - LSE_i(Mr + u)   (:881)
        const T d = c - bj;
        e2 += (double)d * (double)d;
      } else {
        c = dots[q];
        const T d = vs[j] * c - bj;    // column marginal of diag(u) K diag(v) minus b
        e2 += (double)d * (double)d;
        vnew = bj / c;
      }
      for (int r = 0; r < nrep; ++r) XWords<T>::put(xall + (size_t)r * xstride + off_v, j, vnew, epoch);
      cbuf[j] = c;
      vnext[j] = vnew;
    }
    if (check) {  // the CTA's error partial: warp trees, then the eight warp sums in order
      e2 = warp_sum_down(e2);
      if (lane == 0) sred[wrp] = e2;
      __syncthreads();
      if (tid == 0) {
        double e = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) e += sred[w];
        for (int r = 0; r < nrep; ++r) XWords<double>::put(xall + (size_t)r * xstride + off_e, cta, e, epoch);
      }
    }
    RES_TRACE(3);
    // K^T u = 0 or v nan / inf (FLAG_CZERO / FLAG_VBAD of finalize_kernel): b / 0 is not finite either
    const int vbad = res_gather<T>(xv, m, Q, G, epoch, vn);
    RES_TRACE(4);
    double err2 = 0.0;
    if (check && wrp == 0) {  // fixed order: lanes take q = lane, lane + 32, ...; tree; every CTA gets the same bits
      double x = 0.0;
      for (unsigned int q = lane; q < G; q += 32) {
        double e;
        while (!XWords<double>::poll(xe, q, epoch, e)) {
        }
        x += e;
      }
      x = warp_sum_down(x);
      if (lane == 0) sred[0] = x;
    }
    const int nflags = (__syncthreads_or(vbad) && MODE == MODE_KNOPP) ? FLAG_VBAD : 0;
    if (check) err2 = sred[0];
    RES_TRACE(5);

    // ---------------- decision (redundant, identical in every CTA) ----------------
    if (MODE == MODE_KNOPP && (cur_flags != 0 || uf != 0)) {
      // reference guard (_sinkhorn.py:633-645): come back to the previous (u, v) and stop
      status = ST_NUMERR;
      niter = ii;
      sel_v = (sel_v + 2) % 3;
    } else {
      sel_u ^= 1;
      bool stop = false;
      if (check) {
        const double err = sqrt(err2);
        if (cta == 0 && tid == 0) p.err_hist[n_err] = err;
        n_err += 1;
        last_err = err;
        if (err < p.stop_thr) {
          status = ST_CONVERGED;
          niter = ii;
          stop = true;
        }
      }
      if (!stop) {
        if (ii + 1 >= p.max_iter) {
          status = ST_MAXITER;
          niter = ii;
        } else {
          sel_v = (sel_v + 1) % 3;
          iter = ii + 1;
          cur_flags = nflags;
          T* t = vs;  // the gathered v becomes the current one
          vs = vn;
          vn = t;
        }
      }
    }
    __syncthreads();
  }

  // ---------------- leave: CTA 0 publishes the state ----------------
  if (cta == 0 && tid == 0) {
    st->status = status;
    st->iter = iter;
    st->niter = niter;
    st->sel_u = sel_u;
    st->sel_v = sel_v;
    st->cur_flags = cur_flags;
    st->n_err = n_err;
    st->last_err = last_err;
    st->next_flags = 0;
    st->uflag = 0;
    st->ticket = epoch;
  }
}

}  // namespace b2ot
