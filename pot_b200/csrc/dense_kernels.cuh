// Device kernels of the dense Sinkhorn solver (scaling domain and log domain), sm_100a.
//
// Hot kernel: fused_sweep_kernel — ONE pass over K per Sinkhorn iteration.
//   * A thread-block cluster owns whole rows: CTA c of the cluster owns the column slice
//     [c*W, (c+1)*W).  Row-group tiles (R rows x W columns) stream into a multi-stage
//     shared-memory ring with 1-D TMA bulk copies (cp.async.bulk + mbarrier tx), issued by one
//     elected thread as soon as a stage has been drained.
//   * All warps pull a tile into registers (freeing the stage at once), form the partial
//     row sums (K v)_i of their slice; every warp pushes its partials into all CTAs of the
//     cluster with one-sided DSMEM puts (st.async + mbarrier complete_tx: no fences, no CTA
//     barrier), each warp then forms u_i = a_i / (K v)_i and immediately accumulates the tile's
//     contribution u_i K_ij into per-thread column accumulators that live in registers for the
//     whole sweep.  K is read from HBM exactly once per iteration (the reference makes two GEMV
//     passes, ot/bregman/_sinkhorn.py:629-631).
//   * Log domain: same structure; the row pass is a (max, sum-exp) reduction and the column
//     accumulators collect the row-frame plan entries e_ij * f_i, which equal
//     exp(Mr_ij + u_i + v_j): their column sums are the column marginal, so the next
//     v = v + log b - log(colsum) needs no second exp per element (ot/bregman/_sinkhorn.py:881-889).
//
// Reference iteration order is "v then u"; one fused sweep P(v) -> (u, K^T u) covers the u update
// of iteration ii and the K^T u of iteration ii+1, the finalize kernel derives v, the marginal
// error ||v * K^T u - b|| (:653-654) and the numerical guard (:633-645) from the same column sums.
#pragma once

#include "common.cuh"
#include "comm.cuh"

namespace b2ot {

// The fused sweep comes in three CTA shapes:
//   NW = 8,  OCC = 1: 255-register budget, 64 K values per thread per row group (scaling domain);
//   NW = 16, OCC = 1: 128 registers, half the per-thread work, 4 warps per SM sub-partition;
//   NW = 8,  OCC = 2: 128 registers, half-width column slices, TWO independent CTAs per SM so that
//                     one CTA's MUFU-heavy row pass overlaps the other's exchange/combine latency.
constexpr int SMAX = 8;           // max pipeline stages
constexpr int CSMAX = 16;         // max cluster size
constexpr int NPMAX = 160;        // rows of the partial-sum scratch (>= clusters in a grid)

// rows per group: the register tile holds R x VPT vectors = 16 (NW = 8) or 8 (NW = 16) vectors per thread
__host__ __device__ constexpr int rows_per_group(int vpt, int nw, int occ) {
  return ((nw == 8 && occ == 1) ? 16 : 8) / vpt > 8 ? 8 : ((nw == 8 && occ == 1) ? 16 : 8) / vpt;
}

struct SweepParams {
  const void* K;        // n x m, leading dimension ld (elements)
  long long ld, n, m, mp;
  const void* a;        // KNOPP: a ; LOG: log a        (n)
  void* uring;          // 2 x n
  const void* vring;    // 3 x mp
  void* partials;       // NPMAX x mp   (row q = column sums produced by cluster q)
  SolveState* st;
  int W;                // columns per CTA slice (multiple of the vector width)
  int stages;
  int use_hint;         // L2 evict-first on the K stream (K much larger than L2)
  double fi;            // MODE_UNBAL only: u = (a / Kv)^fi (ot/unbalanced/_sinkhorn.py:756)
};

template <typename T, int VPT, int MODE, int NCW, int OCC>
__global__ void __launch_bounds__(NCW * 32, OCC) fused_sweep_kernel(const SweepParams p) {
  using V = typename Vec<T>::type;
  constexpr int VEC = Vec<T>::N;
  constexpr int NCT = NCW * 32;
  constexpr int R = rows_per_group(VPT, NCW, OCC);
  constexpr int PAY = (MODE == MODE_LOG) ? 2 : 1;
  constexpr int NQ = R * PAY;          // payload values per warp per row group (always even)
  constexpr int NCH = NQ / 2;          // ... sent as NCH two-element puts
  constexpr int WCAP = NCT * VPT * VEC;
  constexpr int NSRCMAX = (R > 2) ? NCW : CSMAX * NCW;  // R > 2 only occurs with a single-CTA "cluster"

  __shared__ __align__(16) uint64_t full_bar[SMAX];
  __shared__ __align__(16) uint64_t exch_bar[2];
  __shared__ __align__(16) T exch[2][NSRCMAX][NQ];      // [parity][source warp of the cluster][payload]
  extern __shared__ __align__(128) unsigned char dyn_smem[];
  T* const stage_base = reinterpret_cast<T*>(dyn_smem);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const uint32_t crank = cluster_ctarank();
  const uint32_t csize = cluster_nctarank();
  const long long cid = cluster_id_x();
  const long long ncl = cluster_count_x();
  const long long col0 = (long long)crank * p.W;
  const long long wrem = p.m - col0;
  const int wc = wrem <= 0 ? 0 : (wrem > p.W ? p.W : (int)wrem);  // valid columns of my slice
  const long long NG = (p.n + R - 1) / R;
  const int S = p.stages;
  const T* const Kg = reinterpret_cast<const T*>(p.K);
  const uint32_t row_bytes = (uint32_t)wc * (uint32_t)sizeof(T);
  const uint64_t pol = p.use_hint ? l2_policy_evict_first() : 0ull;
  const int nsrc = (int)csize * NCW;                              // warps in the cluster
  const uint32_t exch_bytes = (uint32_t)nsrc * NQ * (uint32_t)sizeof(T);

  // Producer step (one elected thread): stream row group g into stage s with TMA bulk copies.
  auto issue_group = [&](long long g, int s) {
    const long long row0 = g * R;
    const int rv = (p.n - row0) < R ? (int)(p.n - row0) : R;
    if (row_bytes == 0) {
      mbar_arrive(&full_bar[s]);
      return;
    }
    mbar_arrive_expect_tx(&full_bar[s], row_bytes * (uint32_t)rv);
    for (int r = 0; r < rv; ++r) {
      T* dst = stage_base + ((size_t)s * R + r) * WCAP;
      const T* src = Kg + (row0 + r) * p.ld + col0;
      if (p.use_hint)
        bulk_g2s_hint(dst, src, row_bytes, &full_bar[s], pol);
      else
        bulk_g2s(dst, src, row_bytes, &full_bar[s]);
    }
  };

  if (tid == 0) {
    for (int s = 0; s < S; ++s) mbar_init(&full_bar[s], 1);
    mbar_init(&exch_bar[0], 1);
    mbar_init(&exch_bar[1], 1);
    fence_mbar_init();
    // prologue: fill the ring
    for (int s = 0; s < S; ++s) {
      const long long g = cid + (long long)s * ncl;
      if (g < NG) issue_group(g, s);
    }
  }
  __syncthreads();
  cluster_sync_all();  // every CTA's barriers exist before any remote put

  // Programmatic dependent launch: everything above (barrier setup, the first S row groups of K already in flight)
  // overlaps the tail of the previous kernel of the stream (the finalize of the previous iteration); K does not
  // depend on it.  The solver state and v do: wait for them here.
  griddep_launch_dependents();
  griddep_wait();
  if (p.st->status != ST_RUNNING) {  // sticky stop (uniform over the grid): drain the copies in flight and leave
    for (int s = 0; s < S; ++s) {
      const long long g = cid + (long long)s * ncl;
      if (g < NG) mbar_wait(&full_bar[s], 0);
    }
    return;
  }
  const int sel_u = p.st->sel_u;
  const int sel_v = p.st->sel_v;

  const T* vcur = reinterpret_cast<const T*>(p.vring) + (long long)sel_v * p.mp + col0;
  T* unew = reinterpret_cast<T*>(p.uring) + (long long)(sel_u ^ 1) * p.n;
  const T* ag = reinterpret_cast<const T*>(p.a);

  V vreg[VPT];
  V acc[VPT];
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int e = (k * NCT + tid) * VEC;
    if (e < wc)
      vreg[k] = *reinterpret_cast<const V*>(vcur + e);
    else
      fill(vreg[k], (T)0);
    fill(acc[k], (T)0);
  }

  const bool full_cols = ((VPT - 1) * NCT + tid) * VEC + VEC <= wc;  // all of this thread's vectors are in range

  int cs_log2 = 0;
  while ((1u << cs_log2) < csize) ++cs_log2;
  // this lane's put: payload chunk `my_ch` to peer CTA `my_peer`
  const int my_peer = lane & ((int)csize - 1);
  const int my_ch = lane >> cs_log2;
  const bool sender = my_ch < NCH;
  const uint32_t my_slot = crank * NCW + (uint32_t)warp;   // where this warp's payload lands in every CTA

  int s = 0;          // ring slot of the current group
  uint32_t ph = 0;    // its mbarrier phase parity
  long long it = 0;
  for (long long g = cid; g < NG; g += ncl, ++it) {
    const int par = (int)(it & 1);
    const uint32_t eph = (uint32_t)((it >> 1) & 1);
    const long long row0 = g * R;
    const int rv = (p.n - row0) < R ? (int)(p.n - row0) : R;

    if (tid == 0) mbar_arrive_expect_tx(&exch_bar[par], exch_bytes);  // arm this group's exchange

    T areg[R];
#pragma unroll
    for (int r = 0; r < R; ++r) areg[r] = (r < rv) ? ag[row0 + r] : (T)0;

    // ---- tile: shared -> registers ----
    mbar_wait(&full_bar[s], ph);
    V kreg[R][VPT];
    if (full_cols && rv == R) {  // interior tile: unmasked loads
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const T* srow = stage_base + ((size_t)s * R + r) * WCAP + tid * VEC;
#pragma unroll
        for (int k = 0; k < VPT; ++k) kreg[r][k] = *reinterpret_cast<const V*>(srow + k * NCT * VEC);
      }
    } else {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const T* srow = stage_base + ((size_t)s * R + r) * WCAP;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
          const int e = (k * NCT + tid) * VEC;
          if (r < rv && e < wc)
            kreg[r][k] = *reinterpret_cast<const V*>(srow + e);
          else
            fill(kreg[r][k], MODE == MODE_LOG ? neg_inf<T>() : (T)0);
        }
      }
    }

    // ---- phase A: this warp's partial row reductions over its columns ----
    T pay[NQ];
    T wref[R];  // LOG: this warp's exp reference per row
    T wcorr[R]; // LOG fp32: first-order correction of the rounded reference (1 elsewhere)
#pragma unroll
    for (int r = 0; r < R; ++r) wcorr[r] = (T)1;
    if (MODE != MODE_LOG) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        T s0 = (T)0, s1 = (T)0;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
#pragma unroll
          for (int i = 0; i < VEC; i += 2) {
            s0 = fma(elem(kreg[r][k], i), elem(vreg[k], i), s0);
            s1 = fma(elem(kreg[r][k], i + 1), elem(vreg[k], i + 1), s1);
          }
        }
        pay[r] = s0 + s1;
        wref[r] = (T)0;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) pay[r] += __shfl_xor_sync(0xffffffffu, pay[r], o);
      }
    } else if constexpr (sizeof(T) == 4) {
      // fp32 log-domain row pass on packed fp32x2 arithmetic: x = Mr + v (FADD2), row maximum on four
      // independent FMNMX3 chains, then e = 2^(x log2e - ref log2e) as one FFMA2 + MUFU.EX2 per element.
      T mx[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float m0 = neg_inf<float>(), m1 = m0, m2 = m0, m3 = m0;
#pragma unroll
        for (int k = 0; k < VPT; k += 2) {
          float4& q0 = reinterpret_cast<float4&>(kreg[r][k]);
          float4& q1 = reinterpret_cast<float4&>(kreg[r][k + (VPT > 1 ? 1 : 0)]);
          const float4& w0 = reinterpret_cast<const float4&>(vreg[k]);
          const float4& w1 = reinterpret_cast<const float4&>(vreg[k + (VPT > 1 ? 1 : 0)]);
          add2(q0.x, q0.y, w0.x, w0.y);
          add2(q0.z, q0.w, w0.z, w0.w);
          if (VPT > 1) {
            add2(q1.x, q1.y, w1.x, w1.y);
            add2(q1.z, q1.w, w1.z, w1.w);
            m0 = max3(m0, q0.x, q1.x);
            m1 = max3(m1, q0.y, q1.y);
            m2 = max3(m2, q0.z, q1.z);
            m3 = max3(m3, q0.w, q1.w);
          } else {
            m0 = fmaxf(m0, q0.x);
            m1 = fmaxf(m1, q0.y);
            m2 = fmaxf(m2, q0.z);
            m3 = fmaxf(m3, q0.w);
          }
        }
        mx[r] = fmaxf(max3(m0, m1, m2), m3);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) mx[r] = max_t(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], o));
      }
      constexpr float L2E = 1.4426950408889634f;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float ref = (mx[r] == neg_inf<float>()) ? 0.f : (float)mx[r];
        const float nref = -ref * L2E;
        // the rounding of ref * log2e is a common factor 2^lo of the whole row: divide it back out of the sum here
        // and out of the row-frame factor f_i below (wcorr)
        const float lo = fmaf(ref, L2E, nref);  // exact product minus its rounding
        wcorr[r] = (T)(1.f - lo * 0.6931471805599453f);
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
          float4& q = reinterpret_cast<float4&>(kreg[r][k]);
          float t0, t1, t2, t3;
          fma2(t0, t1, q.x, q.y, L2E, L2E, nref, nref);
          fma2(t2, t3, q.z, q.w, L2E, L2E, nref, nref);
          q.x = ex2_ftz(t0);
          q.y = ex2_ftz(t1);
          q.z = ex2_ftz(t2);
          q.w = ex2_ftz(t3);
          add2(s0, s1, q.x, q.y);
          add2(s2, s3, q.z, q.w);
        }
        wref[r] = (T)ref;
        pay[2 * r] = mx[r];
        pay[2 * r + 1] = (T)(((s0 + s1) + (s2 + s3)) * (float)wcorr[r]);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) pay[2 * r + 1] += __shfl_xor_sync(0xffffffffu, pay[2 * r + 1], o);
      }
    } else {
      T mx[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        T m0 = neg_inf<T>();
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
#pragma unroll
          for (int i = 0; i < VEC; ++i) {
            const T x = elem(kreg[r][k], i) + elem(vreg[k], i);
            elem(kreg[r][k], i) = x;
            m0 = max_t(m0, x);
          }
        }
        mx[r] = m0;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) mx[r] = max_t(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], o));
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const T ref = (mx[r] == neg_inf<T>()) ? (T)0 : mx[r];
        T sm0 = (T)0, sm1 = (T)0;
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
#pragma unroll
          for (int i = 0; i < VEC; i += 2) {
            const T e0 = exp_elem(elem(kreg[r][k], i) - ref);
            const T e1 = exp_elem(elem(kreg[r][k], i + 1) - ref);
            elem(kreg[r][k], i) = e0;
            elem(kreg[r][k], i + 1) = e1;
            sm0 += e0;
            sm1 += e1;
          }
        }
        wref[r] = ref;
        pay[2 * r] = mx[r];
        pay[2 * r + 1] = sm0 + sm1;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) pay[2 * r + 1] += __shfl_xor_sync(0xffffffffu, pay[2 * r + 1], o);
      }
    }

    // ---- push this warp's payload into every CTA of the cluster (st.async + complete_tx) ----
    if (sender) {
      T x0 = pay[0], x1 = pay[1];
#pragma unroll
      for (int c = 1; c < NCH; ++c) {
        if (my_ch == c) {
          x0 = pay[2 * c];
          x1 = pay[2 * c + 1];
        }
      }
      const uint32_t dst = mapa_u32(smem_u32(&exch[par][my_slot][0]) + (uint32_t)(2 * my_ch * sizeof(T)), (uint32_t)my_peer);
      const uint32_t bar = mapa_u32(smem_u32(&exch_bar[par]), (uint32_t)my_peer);
      st_async_v2(dst, x0, x1, bar);
    }

    // ---- wait for the payloads of all warps of the cluster ----
    // Completion also proves that every warp of this CTA has drained stage s: refill it.
    mbar_wait(&exch_bar[par], eph);
    if (tid == 0) {
      const long long gn = g + (long long)S * ncl;
      if (gn < NG) issue_group(gn, s);
    }

    T urow[R];  // KNOPP: u_i ; LOG: f_i (row-frame factor for this warp)
    if (MODE != MODE_LOG) {
      T tot[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        T x = (T)0;
        for (int src = lane; src < nsrc; src += 32) x += exch[par][src][r];
        tot[r] = x;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) tot[r] += __shfl_xor_sync(0xffffffffu, tot[r], o);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        T u = areg[r] / tot[r];
        if constexpr (MODE == MODE_UNBAL) {  // unbalanced: u = (a / Kv)^fi_1 (ot/unbalanced/_sinkhorn.py:756)
          if (p.fi != 1.0) u = pow_row(u, (T)p.fi);
        }
        urow[r] = (r < rv) ? u : (T)0;
        if (crank == 0 && tid == r && r < rv) {
          unew[row0 + r] = u;
          if (!finite_t(u)) atomicOr(&p.st->uflag, 1);
        }
      }
    } else {
      T M[R], Ssum[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        T m0 = neg_inf<T>();
        for (int src = lane; src < nsrc; src += 32) m0 = max_t(m0, exch[par][src][2 * r]);
        M[r] = m0;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) M[r] = max_t(M[r], __shfl_xor_sync(0xffffffffu, M[r], o));
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        T x = (T)0;
        for (int src = lane; src < nsrc; src += 32) {
          const T ms = exch[par][src][2 * r];
          if (ms != neg_inf<T>()) x += exch[par][src][2 * r + 1] * exp_t(ms - M[r]);
        }
        Ssum[r] = x;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) Ssum[r] += __shfl_xor_sync(0xffffffffu, Ssum[r], o);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const T u = areg[r] - (M[r] + log_t(Ssum[r]));
        urow[r] = (r < rv) ? exp_t(wref[r] + u) * wcorr[r] : (T)0;
        if (crank == 0 && tid == r && r < rv) unew[row0 + r] = u;
      }
    }

    // ---- phase B: the resident tile feeds the column accumulators ----
    if constexpr (sizeof(T) == 4 && MODE == MODE_LOG) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float ur = (float)urow[r];
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
          const float4& q = reinterpret_cast<const float4&>(kreg[r][k]);
          float4& ac = reinterpret_cast<float4&>(acc[k]);
          fma2(ac.x, ac.y, q.x, q.y, ur, ur, ac.x, ac.y);
          fma2(ac.z, ac.w, q.z, q.w, ur, ur, ac.z, ac.w);
        }
      }
    } else {
#pragma unroll
      for (int r = 0; r < R; ++r) {
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
#pragma unroll
          for (int i = 0; i < VEC; ++i) elem(acc[k], i) = fma(elem(kreg[r][k], i), urow[r], elem(acc[k], i));
        }
      }
    }

    if (++s == S) {
      s = 0;
      ph ^= 1u;
    }
  }

  // ---- one flush per sweep: this cluster's column sums ----
  T* pout = reinterpret_cast<T*>(p.partials) + cid * p.mp + col0;
#pragma unroll
  for (int k = 0; k < VPT; ++k) {
    const int e = (k * NCT + tid) * VEC;
    if (e < wc) *reinterpret_cast<V*>(pout + e) = acc[k];
  }
  cluster_sync_all();  // no CTA leaves while a peer could still address its shared memory
}

// ------------------------------------------------------------------------------------------------
// Split (two-pass) sweep: any shape, any alignment.  Also provides the column-only pass that seeds
// the first v update.
// ------------------------------------------------------------------------------------------------
struct SplitParams {
  const void* K;
  long long ld, n, m, mp;
  const void* a;
  void* uring;
  const void* vring;
  void* partials;
  SolveState* st;
  int u_slot_xor;   // cols kernel: 0 = current u, 1 = the u just produced
  int nsplit;       // row chunks of the column kernel
  int expect_status;
  double fi;        // KNOPP rows kernel: u = (a / Kv)^fi (unbalanced); 1 = balanced
};

template <typename T, int MODE>
__global__ void __launch_bounds__(256) rows_kernel(const SplitParams p) {
  if (p.st->status != ST_RUNNING) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * 8 + warp;
  if (row >= p.n) return;
  const T* kr = reinterpret_cast<const T*>(p.K) + row * p.ld;
  const T* v = reinterpret_cast<const T*>(p.vring) + (long long)p.st->sel_v * p.mp;
  T* unew = reinterpret_cast<T*>(p.uring) + (long long)(p.st->sel_u ^ 1) * p.n;
  const T ai = reinterpret_cast<const T*>(p.a)[row];
  if (MODE == MODE_KNOPP) {
    T s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    long long j = lane;
    for (; j + 96 < p.m; j += 128) {
      s0 = fma(kr[j], v[j], s0);
      s1 = fma(kr[j + 32], v[j + 32], s1);
      s2 = fma(kr[j + 64], v[j + 64], s2);
      s3 = fma(kr[j + 96], v[j + 96], s3);
    }
    for (; j < p.m; j += 32) s0 = fma(kr[j], v[j], s0);
    const T tot = warp_sum((s0 + s1) + (s2 + s3));
    if (lane == 0) {
      T u = ai / tot;
      if (p.fi != 1.0) u = pow_row(u, (T)p.fi);
      unew[row] = u;
      if (!finite_t(u)) atomicOr(&p.st->uflag, 1);
    }
  } else {
    T mx = neg_inf<T>();
    for (long long j = lane; j < p.m; j += 32) mx = max_t(mx, kr[j] + v[j]);
    mx = warp_max(mx);
    const T ref = (mx == neg_inf<T>()) ? (T)0 : mx;
    T s0 = 0, s1 = 0;
    long long j = lane;
    for (; j + 32 < p.m; j += 64) {
      s0 += exp_t((kr[j] + v[j]) - ref);
      s1 += exp_t((kr[j + 32] + v[j + 32]) - ref);
    }
    for (; j < p.m; j += 32) s0 += exp_t((kr[j] + v[j]) - ref);
    const T tot = warp_sum(s0 + s1);
    if (lane == 0) unew[row] = ai - (ref + log_t(tot));
  }
}

// KNOPP: partial_j = sum_i K_ij u_i ; LOG: partial_j = sum_i exp(K_ij + u_i + v_j)
template <typename T, int MODE>
__global__ void __launch_bounds__(256) cols_kernel(const SplitParams p) {
  if (p.st->status != p.expect_status) return;
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long rows_per = (p.n + p.nsplit - 1) / p.nsplit;
  const long long r0 = (long long)blockIdx.y * rows_per;
  const long long r1 = (r0 + rows_per < p.n) ? r0 + rows_per : p.n;
  if (j >= p.m) return;
  const T* Kc = reinterpret_cast<const T*>(p.K) + j;
  const T* u = reinterpret_cast<const T*>(p.uring) + (long long)(p.st->sel_u ^ p.u_slot_xor) * p.n;
  T vj = (T)0;
  if (MODE == MODE_LOG) vj = (reinterpret_cast<const T*>(p.vring) + (long long)p.st->sel_v * p.mp)[j];
  T a0 = 0, a1 = 0, a2 = 0, a3 = 0;
  long long i = r0;
  if (MODE == MODE_KNOPP) {
    for (; i + 3 < r1; i += 4) {
      a0 = fma(Kc[i * p.ld], u[i], a0);
      a1 = fma(Kc[(i + 1) * p.ld], u[i + 1], a1);
      a2 = fma(Kc[(i + 2) * p.ld], u[i + 2], a2);
      a3 = fma(Kc[(i + 3) * p.ld], u[i + 3], a3);
    }
    for (; i < r1; ++i) a0 = fma(Kc[i * p.ld], u[i], a0);
  } else {
    for (; i + 1 < r1; i += 2) {
      a0 += exp_t((Kc[i * p.ld] + u[i]) + vj);
      a1 += exp_t((Kc[(i + 1) * p.ld] + u[i + 1]) + vj);
    }
    for (; i < r1; ++i) a0 += exp_t((Kc[i * p.ld] + u[i]) + vj);
  }
  reinterpret_cast<T*>(p.partials)[(long long)blockIdx.y * p.mp + j] = (a0 + a1) + (a2 + a3);
}

// Exact column log-sum-exp partials (online max): rows [0,nsplit) of `partials` = max, rows
// [nsplit, 2 nsplit) = sum.  Used for the first v update of the log domain (u given) and as the
// fallback when a row-frame column sum underflows.
template <typename T>
__global__ void __launch_bounds__(256) cols_lse_kernel(const SplitParams p) {
  if (p.st->status != p.expect_status) return;
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long rows_per = (p.n + p.nsplit - 1) / p.nsplit;
  const long long r0 = (long long)blockIdx.y * rows_per;
  const long long r1 = (r0 + rows_per < p.n) ? r0 + rows_per : p.n;
  if (j >= p.m) return;
  const T* Kc = reinterpret_cast<const T*>(p.K) + j;
  const T* u = reinterpret_cast<const T*>(p.uring) + (long long)(p.st->sel_u ^ p.u_slot_xor) * p.n;
  T mx = neg_inf<T>(), sm = (T)0;
  for (long long i = r0; i < r1; ++i) {
    const T x = Kc[i * p.ld] + u[i];
    if (x == neg_inf<T>()) continue;
    if (x > mx) {
      sm = sm * exp_t(mx - x) + (T)1;  // exp(-inf) = 0 on the first hit
      mx = x;
    } else {
      sm += exp_t(x - mx);
    }
  }
  T* P = reinterpret_cast<T*>(p.partials);
  P[(long long)blockIdx.y * p.mp + j] = mx;
  P[(long long)(p.nsplit + blockIdx.y) * p.mp + j] = sm;
}

struct LseParams {
  void* partials;
  long long m, mp;
  int nsplit;
  const void* logb;
  void* vring;
  SolveState* st;
  int expect_status;
  int advance_iter;  // 0: seeding pass (iteration counter untouched), 1: fallback inside iteration
  int max_iter;
};
// rows 2ns: local max, 2ns+1: max to be all-reduced(max), 2ns+2: rescaled sum to be all-reduced(sum)
template <typename T>
__global__ void __launch_bounds__(256) lse_combine_kernel(const LseParams p) {
  if (p.st->status != p.expect_status) return;
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  if (j >= p.m) return;
  T* P = reinterpret_cast<T*>(p.partials);
  T M = neg_inf<T>();
  for (int y = 0; y < p.nsplit; ++y) M = max_t(M, P[(long long)y * p.mp + j]);
  T S = (T)0;
  for (int y = 0; y < p.nsplit; ++y) {
    const T my = P[(long long)y * p.mp + j];
    if (my != neg_inf<T>()) S += P[(long long)(p.nsplit + y) * p.mp + j] * exp_t(my - M);
  }
  P[(long long)(2 * p.nsplit) * p.mp + j] = M;
  P[(long long)(2 * p.nsplit + 1) * p.mp + j] = M;
  P[(long long)(2 * p.nsplit + 2) * p.mp + j] = S;
}
template <typename T>
__global__ void __launch_bounds__(256) lse_rescale_kernel(const LseParams p) {
  if (p.st->status != p.expect_status) return;
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  if (j >= p.m) return;
  T* P = reinterpret_cast<T*>(p.partials);
  const T Ml = P[(long long)(2 * p.nsplit) * p.mp + j];
  const T Mg = P[(long long)(2 * p.nsplit + 1) * p.mp + j];
  T& S = P[(long long)(2 * p.nsplit + 2) * p.mp + j];
  S = (Ml == neg_inf<T>()) ? (T)0 : S * exp_t(Ml - Mg);
}
template <typename T>
__global__ void __launch_bounds__(256) lse_finish_kernel(const LseParams p) {
  if (p.st->status != p.expect_status) return;
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  if (j >= p.m) return;
  const T* P = reinterpret_cast<const T*>(p.partials);
  const T Mg = P[(long long)(2 * p.nsplit + 1) * p.mp + j];
  const T S = P[(long long)(2 * p.nsplit + 2) * p.mp + j];
  const int nxt = (p.st->sel_v + 1) % 3;
  T* vnext = reinterpret_cast<T*>(p.vring) + (long long)nxt * p.mp;
  vnext[j] = reinterpret_cast<const T*>(p.logb)[j] - (Mg + log_t(S));
}
__global__ void lse_advance_kernel(SolveState* st, int expect_status, int advance_iter) {
  if (st->status != expect_status) return;
  st->sel_v = (st->sel_v + 1) % 3;
  if (advance_iter) st->iter += 1;
  st->status = ST_RUNNING;
}

// ------------------------------------------------------------------------------------------------
// Finalize: column sums -> v update, marginal error, guard flags, stop decision (all on device).
// ------------------------------------------------------------------------------------------------
struct FinParams {
  const void* src;      // nsrc rows of column sums (stride src_ld)
  long long src_ld;
  int nsrc;
  const void* extra_flag;  // optional: T value(s) != 0 mean "some rank produced a non-finite u"
  long long extra_flag_ld; // ... flag_n values, extra_flag_ld elements apart
  int extra_flag_n;
  int use_comm;            // 1: `src` are the peer-exchange slots of sequence `seq`: wait for all ranks first
                           // 2: tagged-word exchange INSIDE this kernel (comm.cuh): this rank's partial rows `lsrc` are
                           //    reduced and stored to every peer, then every column is taken from the R slots
  unsigned long long seq;
  CommView comm;
  const void* lsrc;        // use_comm == 2: local partial rows (lnsrc rows, stride lsrc_ld)
  long long lsrc_ld;
  int lnsrc;
  const int* uflag_src;    // use_comm == 2: "this rank produced a non-finite u" (travels as element m)
  unsigned long long ll_off;  // byte offset of the tagged-word region inside a slot
  const void* b;
  const void* logb;
  void* vring;
  void* cbuf;
  long long m, mp;
  SolveState* st;
  double* blk_err;
  double* err_hist;
  int max_iter, check_every;
  double stop_thr;
  int init;             // KNOPP seeding pass: only v0 = b / (K^T u0) and its guard flags
  int qsplit;           // 1: blocks of 32 columns x 8 warps over the partial rows (small m, many partial rows)
  // stabilized scaling iteration (ot/bregman/_sinkhorn.py:1105-1382), KNOPP only, tau > 0 turns it on: after the sweep
  // the pair (u, v) of this iteration is tested against tau (:1274); when it is exceeded st->absorb_now is raised and
  // the gated kernels that follow (absorb_vec_kernel, stab_rebuild_kernel, this kernel with init = 1) fold (u, v) into
  // the potentials, rewrite K and re-seed the column pass (:1275-1288).  The error of such an iteration is the
  // reference's: measured on Gamma AFTER the reset u = 1/n, v = 1/m (:1303), i.e. on the plan scaled by 1/(n m).
  // Stop test `err <= stop_thr` (:1316).
  double tau;
  const void* uring;    // 2 x n, read for max|u| when tau > 0
  long long n;
  int gate_rebuild;     // 1 (init pass after an absorption): run only when st->rebuild is raised, clear it at the end
};

// STAB: the stabilized variant is its own instantiation: the extra state costs the fp32 kernel 8 registers, and the
// balanced headline runs exactly at the register limit where TWO resident finalize blocks (32 registers) and the
// programmatically launched next sweep CTA (192) share an SM: 40 registers pushed the sweep's early start behind the
// finalize blocks' exit, +30 us per 2.7 ms iteration at 65536^2 (profiles/r01c_ab_finalize_stabilized.jsonl).
// LL: the tagged-word exchange (use_comm == 2) is its own instantiation for the same reason (79 registers with the
// polling loops in; capped at 48 so that a finalize block and the early-started sweep CTA still share an SM).
template <typename T, int MODE, bool STAB, bool LL>
__device__ __forceinline__ void finalize_body(const FinParams& p) {
  __shared__ double sred[256];
  __shared__ int sflags;
  __shared__ int s_last;
  SolveState* st = p.st;
  griddep_wait();               // column partials of the sweep (or the peer put) are complete and visible
  griddep_launch_dependents();  // let the next sweep start its prologue (barriers, first K tiles) under this kernel
  const int status_in = st->status;
  const int tid = threadIdx.x;
  // fused "take" of the NVLink exchange.  The wait is NOT gated on the solver status: every sequence number is
  // waited for (comm.cuh), also by the no-op iterations enqueued after the solve stopped -- except after a timeout.
  __shared__ int s_tmo;
  const unsigned int tag = (unsigned int)p.seq;
  const int par = (int)(p.seq & 1ull);
  if (LL && p.use_comm == 2) {
    // Tagged-word exchange, put side: reduce the local partial rows of every column (and the u flag as element m) and
    // store the words into this rank's slot on every peer.  NOT gated on the solver status: every sequence number is
    // sent and taken by every rank (comm.cuh), a stopped solver sends zeros.
    if (tid == 0) s_tmo = 0;
    const T* lsrc = reinterpret_cast<const T*>(p.lsrc);
    for (long long j = (long long)blockIdx.x * 256 + tid; j <= p.m; j += (long long)gridDim.x * 256) {
      T c = (T)0;
      if (status_in == ST_RUNNING) {
        if (j < p.m)
          for (int q = 0; q < p.lnsrc; ++q) c += __ldcg(&lsrc[(long long)q * p.lsrc_ld + j]);
        else
          c = (T)(atomicOr(const_cast<int*>(p.uflag_src), 0) ? 1 : 0);
      }
      for (int r = 0; r < p.comm.nranks; ++r)
        LLWords<T>::put(comm_ll_slot(p.comm, r, par, p.comm.rank, p.ll_off), j, c, tag);
    }
    if (status_in != ST_RUNNING) {  // consume the peers' words of this sequence number all the same
      if (status_in != ST_COMM_TIMEOUT)
        for (long long j = (long long)blockIdx.x * 256 + tid; j <= p.m; j += (long long)gridDim.x * 256) {
          T c;
          if (!comm_ll_take<T>(p.comm, par, p.ll_off, j, tag, c)) break;
        }
      return;
    }
    __syncthreads();
  } else if (p.use_comm && status_in != ST_COMM_TIMEOUT && !comm_wait(p.comm, p.seq)) {
    if (tid == 0) st->status = ST_COMM_TIMEOUT;
    return;
  }
  if (status_in != ST_RUNNING) return;
  if (STAB && p.gate_rebuild && st->rebuild == 0) return;
  const int sel_v = st->sel_v;
  const int nxt = (sel_v + 1) % 3;
  const T* vcur = reinterpret_cast<const T*>(p.vring) + (long long)sel_v * p.mp;
  T* vnext = reinterpret_cast<T*>(p.vring) + (long long)nxt * p.mp;
  const T* src = reinterpret_cast<const T*>(p.src);
  const T* b = reinterpret_cast<const T*>(p.b);
  const T* logb = reinterpret_cast<const T*>(p.logb);
  T* cbuf = reinterpret_cast<T*>(p.cbuf);
  if (tid == 0) sflags = 0;
  __syncthreads();

  double e2 = 0.0, e2s = 0.0;
  int flags = 0;
  const bool stab = STAB && MODE == MODE_KNOPP && !p.init;
  const T tau = (T)p.tau;
  const T reset_scale = ((T)1 / (T)p.n) * ((T)1 / (T)p.m);  // u = 1/n, v = 1/m after an absorption (:1283-1287)
  if (STAB && stab) {  // max|u| of the u this sweep produced
    const T* unew = reinterpret_cast<const T*>(p.uring) + (long long)(st->sel_u ^ 1) * p.n;
    for (long long i = (long long)blockIdx.x * 256 + tid; i < p.n; i += (long long)gridDim.x * 256)
      if (abs_t(__ldcg(&unew[i])) > tau) flags |= 8;
  }
  auto column = [&](long long j, T c) {
    const T bj = b[j];
    const T vj = vcur[j];
    T vn;
    if (MODE == MODE_KNOPP) {
      const T d = vj * c - bj;  // column marginal of diag(u) K diag(v) minus b
      e2 += (double)d * (double)d;
      vn = bj / c;
      if (c == (T)0) flags |= FLAG_CZERO;
      if (!finite_t(vn)) flags |= FLAG_VBAD;
      if constexpr (STAB) {
        if (stab) {
          if (abs_t(vj) > tau) flags |= 8;
          const T ds = (vj * c) * reset_scale - bj;
          e2s += (double)ds * (double)ds;
        }
      }
    } else {
      const T d = c - bj;
      e2 += (double)d * (double)d;
      vn = vj + (logb[j] - log_t(c));
      if (!(c >= tiny_guard<T>())) flags |= 4;
    }
    cbuf[j] = c;
    vnext[j] = vn;
  };
  if (LL && p.use_comm == 2) {
    for (long long j = (long long)blockIdx.x * 256 + tid; j < p.m; j += (long long)gridDim.x * 256) {
      T c;
      if (!comm_ll_take<T>(p.comm, par, p.ll_off, j, tag, c)) {
        s_tmo = 1;
        break;
      }
      column(j, c);
    }
    __syncthreads();
    if (s_tmo) {
      if (tid == 0) st->status = ST_COMM_TIMEOUT;
      return;
    }
  } else if (!p.qsplit) {
    for (long long j = (long long)blockIdx.x * 256 + tid; j < p.m; j += (long long)gridDim.x * 256) {
      T c = (T)0;
      for (int q = 0; q < p.nsrc; ++q) c += __ldcg(&src[(long long)q * p.src_ld + j]);
      column(j, c);
    }
  } else {
    // Small problems: few columns, many partial rows (one per sweep cluster, up to NPMAX).  A block takes 32 columns;
    // warp w sums the partial rows q = w, w + 8, ... (coalesced 32-column loads, 8 chains of L2 latency instead of one
    // thread walking all rows: that serial walk made this kernel ~10 us at n = m = 1000), warp 0 combines the eight
    // partial sums in a fixed order and finishes the columns.
    __shared__ T sq[8][33];
    const int lane = tid & 31, qw = tid >> 5;
    for (long long j0 = (long long)blockIdx.x * 32; j0 < p.m; j0 += (long long)gridDim.x * 32) {
      const long long j = j0 + lane;
      T part = (T)0;
      if (j < p.m)
        for (int q = qw; q < p.nsrc; q += 8) part += __ldcg(&src[(long long)q * p.src_ld + j]);
      sq[qw][lane] = part;
      __syncthreads();
      if (qw == 0 && j < p.m) {
        const T c = ((sq[0][lane] + sq[1][lane]) + (sq[2][lane] + sq[3][lane])) +
                    ((sq[4][lane] + sq[5][lane]) + (sq[6][lane] + sq[7][lane]));
        column(j, c);
      }
      __syncthreads();
    }
  }
  sred[tid] = e2;
  if (flags) atomicOr(&sflags, flags);
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) sred[tid] += sred[tid + o];
    __syncthreads();
  }
  if (STAB && stab) {  // second total: the error as the reference reports it when this iteration absorbs
    const double first = sred[0];
    __syncthreads();
    sred[tid] = e2s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (tid < o) sred[tid] += sred[tid + o];
      __syncthreads();
    }
    if (tid == 0) {
      p.blk_err[512 + blockIdx.x] = sred[0];  // grid <= 512 blocks when tau > 0 (launch_finalize)
      sred[0] = first;
    }
    __syncthreads();
  }
  if (tid == 0) {
    p.blk_err[blockIdx.x] = sred[0];
    if (sflags & 3) atomicOr(&st->next_flags, sflags & 3);
    if (sflags & 4) atomicOr(&st->need_exact, 1);
    if (sflags & 8) atomicOr(&st->absorb, 1);
    __threadfence();
    const unsigned int t = atomicAdd(&st->ticket, 1u);
    s_last = (t == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;

  // ---- last block: deterministic error total + stop decision ----
  __threadfence();
  double x = 0.0;
  for (unsigned int i = tid; i < gridDim.x; i += 256) x += __ldcg(&p.blk_err[i]);
  sred[tid] = x;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (tid < o) sred[tid] += sred[tid + o];
    __syncthreads();
  }
  const int absorb = (STAB && stab) ? atomicOr(&st->absorb, 0) : 0;
  if (STAB && absorb) {  // block-uniform: total of the scaled error instead
    __syncthreads();
    double y = 0.0;
    for (unsigned int i = tid; i < gridDim.x; i += 256) y += __ldcg(&p.blk_err[512 + i]);
    sred[tid] = y;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (tid < o) sred[tid] += sred[tid + o];
      __syncthreads();
    }
  }
  if (tid != 0) return;
  const int nflags = atomicOr(&st->next_flags, 0);
  const int nexact = atomicOr(&st->need_exact, 0);
  int uflag = atomicOr(&st->uflag, 0);
  if (LL && p.use_comm == 2) {  // element m of the exchange: some rank produced a non-finite u
    T fsum;
    if (!comm_ll_take<T>(p.comm, par, p.ll_off, p.m, tag, fsum)) {
      st->status = ST_COMM_TIMEOUT;
      return;
    }
    if (fsum != (T)0) uflag = 1;
  } else if (p.extra_flag != nullptr) {
    for (int q = 0; q < p.extra_flag_n; ++q)
      if (__ldcg(reinterpret_cast<const T*>(p.extra_flag) + (long long)q * p.extra_flag_ld) != (T)0) uflag = 1;
  }
  const int ii = st->iter;
  if (p.init) {
    st->sel_v = nxt;
    st->cur_flags = nflags;
    if (STAB && p.gate_rebuild) st->rebuild = 0;
  } else if (MODE == MODE_KNOPP && (st->cur_flags != 0 || uflag != 0)) {
    // reference guard (_sinkhorn.py:633-645): come back to the previous (u, v) and stop
    st->status = ST_NUMERR;
    st->niter = ii;
    st->sel_v = (sel_v + 2) % 3;
  } else {
    st->sel_u ^= 1;  // the u of this sweep becomes current
    bool stop = false;
    if (ii % p.check_every == 0) {
      const double err = sqrt(sred[0]);
      p.err_hist[st->n_err] = err;
      st->n_err += 1;
      st->last_err = err;
      if (STAB ? (err <= p.stop_thr) : (err < p.stop_thr)) {
        st->status = ST_CONVERGED;
        st->niter = ii;
        stop = true;
      }
    }
    if (!stop) {
      if (ii + 1 >= p.max_iter) {
        st->status = ST_MAXITER;
        st->niter = ii;
      } else if (MODE == MODE_LOG && nexact) {
        st->status = ST_PAUSE_EXACT;  // v_next unusable for some columns: exact column pass needed
      } else if (STAB && MODE == MODE_KNOPP && absorb) {
        // stabilized: the absorption chain resets every slot of both rings and re-seeds v; only the index moves on
        st->iter = ii + 1;
        st->cur_flags = 0;
      } else {
        st->sel_v = nxt;
        st->iter = ii + 1;
        st->cur_flags = nflags;
      }
    }
  }
  if (STAB && absorb && st->status != ST_NUMERR) {  // also when this iteration is the last one (:1275 runs before the tests)
    st->absorb_now = 1;
    st->n_absorb += 1;
  }
  st->next_flags = 0;
  st->uflag = 0;
  st->need_exact = 0;
  st->absorb = 0;
  st->ticket = 0;
  __threadfence();
}

template <typename T, int MODE, bool STAB = false>
__global__ void __launch_bounds__(256) finalize_kernel(const FinParams p) {
  finalize_body<T, MODE, STAB, false>(p);
}
template <typename T, int MODE>
__global__ void __launch_bounds__(256, 5) finalize_ll_kernel(const FinParams p) {
  finalize_body<T, MODE, false, true>(p);
}

// ------------------------------------------------------------------------------------------------
// Stabilized scaling iteration, absorption chain (ot/bregman/_sinkhorn.py:1274-1288), all gated on device flags so
// that the chain can sit in the enqueued iteration stream (and in its CUDA graph) and costs three empty launches when
// nothing is to absorb.
//   absorb_vec_kernel   : alpha/reg += log u, beta/reg += log v (:1281; double accumulators for both dtypes), u = 1/n,
//                         v = 1/m in every ring slot (:1283-1287)
//   stab_rebuild_kernel : K = exp(-(M - alpha - beta)/reg) (:1288) from Mr = -M/reg, written over the handle's K, and in
//                         the same pass the column sums K^T u (u = 1/n) that seed the next v update (:1270)
//   finalize_kernel with init = 1, gate_rebuild = 1: v = b / K^T u
// An absorbing iteration costs 3 passes over an n x m matrix (sweep reads K; rebuild reads Mr, writes K) instead of the
// reference's 5 (two GEMVs, get_K reads M and writes K, get_Gamma for the error on check iterations on top).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) absorb_vec_kernel(SolveState* st, void* uring_, void* vring_, long long n,
                                                         long long m, long long mp, double* __restrict__ alpha_r,
                                                         double* __restrict__ beta_r) {
  if (st->absorb_now == 0) return;
  T* uring = reinterpret_cast<T*>(uring_);
  T* vring = reinterpret_cast<T*>(vring_);
  const int sel_u = st->sel_u, sel_v = st->sel_v;
  const T u0 = (T)1 / (T)n, v0 = (T)1 / (T)m;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    alpha_r[i] += log((double)uring[(long long)sel_u * n + i]);
    uring[i] = u0;
    uring[n + i] = u0;
  }
  for (long long j = (long long)blockIdx.x * 256 + threadIdx.x; j < m; j += (long long)gridDim.x * 256) {
    beta_r[j] += log((double)vring[(long long)sel_v * mp + j]);
    vring[j] = v0;
    vring[mp + j] = v0;
    vring[2 * mp + j] = v0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&st->ticket, 1u) == gridDim.x - 1) {  // every block has passed the gate: safe to drop it
      st->ticket = 0;
      st->absorb_now = 0;
      if (st->status == ST_RUNNING) st->rebuild = 1;
      __threadfence();
    }
  }
}

// grid (ceil(m / (256 VEC)), gy): a thread owns VEC columns and walks the rows y, y + gy, ... two at a time
template <typename T, bool VECT>
__global__ void __launch_bounds__(256) stab_rebuild_kernel(const T* __restrict__ Mr, long long ldM, T* __restrict__ K,
                                                           long long ldK, const double* __restrict__ alpha_r,
                                                           const double* __restrict__ beta_r, long long n, long long m,
                                                           long long mp, T* __restrict__ partials,
                                                           const SolveState* st) {
  if (st->rebuild == 0) return;
  constexpr int VEC = VECT ? Vec<T>::N : 1;
  const long long j = ((long long)blockIdx.x * 256 + threadIdx.x) * VEC;
  if (j >= m) return;
  const T u0 = (T)1 / (T)n;
  double bj[VEC];  // the potentials are accumulated and added in double for both dtypes: with T = float the exponent
  T acc[VEC];      // Mr + alpha/reg + beta/reg cancels terms of magnitude 1/reg, and alpha/reg grows by one log per absorption
#pragma unroll
  for (int e = 0; e < VEC; ++e) {
    bj[e] = beta_r[j + e];
    acc[e] = (T)0;
  }
  auto load = [&](long long i, T (&x)[VEC]) {
    if constexpr (VECT) {
      const typename Vec<T>::type q = __ldcs(reinterpret_cast<const typename Vec<T>::type*>(Mr + i * ldM + j));
#pragma unroll
      for (int e = 0; e < VEC; ++e) x[e] = elem(q, e);
    } else {
      x[0] = __ldcs(Mr + i * ldM + j);
    }
  };
  auto store = [&](long long i, const T (&y)[VEC]) {
    if constexpr (VECT) {
      typename Vec<T>::type q;
#pragma unroll
      for (int e = 0; e < VEC; ++e) elem(q, e) = y[e];
      *reinterpret_cast<typename Vec<T>::type*>(K + i * ldK + j) = q;  // plain store: the sweep reads K next (L2)
    } else {
      K[i * ldK + j] = y[0];
    }
  };
  for (long long i = blockIdx.y; i < n; i += 2LL * gridDim.y) {
    const long long i2 = i + gridDim.y;
    const bool two = i2 < n;
    T x0[VEC], x1[VEC], y0[VEC], y1[VEC];
    load(i, x0);
    if (two) load(i2, x1);
    const double a0 = alpha_r[i], a1 = alpha_r[two ? i2 : i];
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      y0[e] = exp_t((T)(((double)x0[e] + a0) + bj[e]));
      acc[e] = fma(y0[e], u0, acc[e]);
    }
    store(i, y0);
    if (two) {
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        y1[e] = exp_t((T)(((double)x1[e] + a1) + bj[e]));
        acc[e] = fma(y1[e], u0, acc[e]);
      }
      store(i2, y1);
    }
  }
#pragma unroll
  for (int e = 0; e < VEC; ++e) partials[(long long)blockIdx.y * mp + j + e] = acc[e];
}

// ------------------------------------------------------------------------------------------------
// Finalize of the UNBALANCED scaling iteration (ot/unbalanced/_sinkhorn.py:749-787).  The fused sweep has
// produced u = (a / Kv)^fi_1 and the column sums c = K^T u; here v = (b / c)^fi_2, the guard (c == 0,
// nan / inf in u or v -> keep the previous (u, v) and stop, :760-772), the error
// err = (|u - uprev|_inf / max(|u|_inf, |uprev|_inf, 1) + |v - vprev|_inf / max(|v|_inf, |vprev|_inf, 1)) / 2
// every iteration and the `<` stop test (:774-787).  No iteration skew: sweep i + this kernel = iteration i.
// ------------------------------------------------------------------------------------------------
struct FinUnbalParams {
  const void* src;      // nsrc rows of column sums (stride src_ld)
  long long src_ld;
  int nsrc;
  const void* b;
  void* vring;          // 3 x mp
  const void* uring;    // 2 x n
  void* cbuf;
  long long n, m, mp;
  SolveState* st;
  double* blk;          // [6][gridDim.x] block maxima: du, mu, mup, dv, mv, mvp
  double* err_hist;
  int max_iter;
  double stop_thr;
  double fi;            // fi_2
  int qsplit;           // 1: blocks of 32 columns x 8 warps over the partial rows (small m, many partial rows)
  // row-sharded runs: `src` is the all-reduced column-sum vector; the u statistics (|u - uprev|_inf, |u|_inf,
  // |uprev|_inf over ALL ranks' rows: 3 values of T, max-all-reduced, written by unbal_ustats_kernel) and the summed
  // "u not finite" flags of the ranks come from memory instead of this rank's rows / state word
  const void* ustats;
  const void* extra_flag;
};

// Row-sharded unbalanced runs: this rank's part of the u statistics of the error, 3 values of T for a MAX all-reduce.
// One block (n_local is a fraction of the rows; the kernel sits between two collectives anyway).
template <typename T>
__global__ void __launch_bounds__(1024) unbal_ustats_kernel(const void* uring, long long n, const SolveState* st,
                                                            void* out) {
  __shared__ double sred[3][32];
  if (st->status != ST_RUNNING) return;
  const int sel_u = st->sel_u;
  const T* ucur = reinterpret_cast<const T*>(uring) + (long long)sel_u * n;
  const T* unew = reinterpret_cast<const T*>(uring) + (long long)(sel_u ^ 1) * n;
  double mx[3] = {0.0, 0.0, 0.0};
  for (long long i = threadIdx.x; i < n; i += blockDim.x) {
    const T un = unew[i], uo = ucur[i];
    mx[0] = fmax(mx[0], (double)abs_t(un - uo));
    mx[1] = fmax(mx[1], (double)abs_t(un));
    mx[2] = fmax(mx[2], (double)abs_t(uo));
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double x = mx[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_xor_sync(0xffffffffu, x, o));
    if ((threadIdx.x & 31) == 0) sred[k][threadIdx.x >> 5] = x;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double x = 0.0;
    for (int w = 0; w < 32; ++w) x = fmax(x, sred[threadIdx.x][w]);
    reinterpret_cast<T*>(out)[threadIdx.x] = (T)x;  // maxima of |T| values: exact in T
  }
}

template <typename T>
__global__ void __launch_bounds__(256) finalize_unbal_kernel(const FinUnbalParams p) {
  __shared__ double sred[6][8];
  __shared__ int sflags;
  __shared__ int s_last;
  SolveState* st = p.st;
  griddep_wait();
  griddep_launch_dependents();
  if (st->status != ST_RUNNING) return;
  const int tid = threadIdx.x;
  const int sel_v = st->sel_v, sel_u = st->sel_u;
  const int nxt = (sel_v + 1) % 3;
  const T* vcur = reinterpret_cast<const T*>(p.vring) + (long long)sel_v * p.mp;
  T* vnext = reinterpret_cast<T*>(p.vring) + (long long)nxt * p.mp;
  const T* ucur = reinterpret_cast<const T*>(p.uring) + (long long)sel_u * p.n;
  const T* unew = reinterpret_cast<const T*>(p.uring) + (long long)(sel_u ^ 1) * p.n;
  const T* src = reinterpret_cast<const T*>(p.src);
  const T* b = reinterpret_cast<const T*>(p.b);
  T* cbuf = reinterpret_cast<T*>(p.cbuf);
  if (tid == 0) sflags = 0;
  __syncthreads();

  double mx[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  int flags = 0;
  auto column = [&](long long j, T c) {
    T vn = b[j] / c;
    if (p.fi != 1.0) vn = pow_t(vn, (T)p.fi);
    const T vj = vcur[j];
    if (c == (T)0) flags |= FLAG_CZERO;
    if (!finite_t(vn)) flags |= FLAG_VBAD;
    mx[3] = fmax(mx[3], (double)abs_t(vn - vj));
    mx[4] = fmax(mx[4], (double)abs_t(vn));
    mx[5] = fmax(mx[5], (double)abs_t(vj));
    cbuf[j] = c;
    vnext[j] = vn;
  };
  if (!p.qsplit) {
    for (long long j = (long long)blockIdx.x * 256 + tid; j < p.m; j += (long long)gridDim.x * 256) {
      T c = (T)0;
      for (int q = 0; q < p.nsrc; ++q) c += __ldcg(&src[(long long)q * p.src_ld + j]);
      column(j, c);
    }
  } else {  // small problems: 32 columns per block, 8 warps over the partial rows (see finalize_kernel)
    __shared__ T sq[8][33];
    const int lane = tid & 31, qw = tid >> 5;
    for (long long j0 = (long long)blockIdx.x * 32; j0 < p.m; j0 += (long long)gridDim.x * 32) {
      const long long j = j0 + lane;
      T part = (T)0;
      if (j < p.m)
        for (int q = qw; q < p.nsrc; q += 8) part += __ldcg(&src[(long long)q * p.src_ld + j]);
      sq[qw][lane] = part;
      __syncthreads();
      if (qw == 0 && j < p.m) {
        const T c = ((sq[0][lane] + sq[1][lane]) + (sq[2][lane] + sq[3][lane])) +
                    ((sq[4][lane] + sq[5][lane]) + (sq[6][lane] + sq[7][lane]));
        column(j, c);
      }
      __syncthreads();
    }
  }
  if (p.ustats == nullptr) {
    for (long long i = (long long)blockIdx.x * 256 + tid; i < p.n; i += (long long)gridDim.x * 256) {
      const T un = unew[i], uo = ucur[i];
      mx[0] = fmax(mx[0], (double)abs_t(un - uo));
      mx[1] = fmax(mx[1], (double)abs_t(un));
      mx[2] = fmax(mx[2], (double)abs_t(uo));
    }
  }
  if (flags) atomicOr(&sflags, flags);
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    double x = mx[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_xor_sync(0xffffffffu, x, o));
    if ((tid & 31) == 0) sred[k][tid >> 5] = x;
  }
  __syncthreads();
  if (tid < 6) {
    double x = 0.0;
    for (int w = 0; w < 8; ++w) x = fmax(x, sred[tid][w]);
    p.blk[(long long)tid * gridDim.x + blockIdx.x] = x;
  }
  __syncthreads();
  if (tid == 0) {
    if (sflags) atomicOr(&st->next_flags, sflags);
    __threadfence();
    const unsigned int t = atomicAdd(&st->ticket, 1u);
    s_last = (t == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;

  // ---- last block: maxima over the blocks, guard, error, stop decision ----
  __threadfence();
  if (tid < 6 * 32) {
    const int k = tid >> 5;
    double x = 0.0;
    for (unsigned int i = tid & 31; i < gridDim.x; i += 32) x = fmax(x, __ldcg(&p.blk[(long long)k * gridDim.x + i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_xor_sync(0xffffffffu, x, o));
    if ((tid & 31) == 0) sred[k][0] = x;
  }
  __syncthreads();
  if (tid != 0) return;
  const int nflags = atomicOr(&st->next_flags, 0);
  int uflag = atomicOr(&st->uflag, 0);
  if (p.extra_flag != nullptr && __ldcg(reinterpret_cast<const T*>(p.extra_flag)) > (T)0) uflag = 1;  // any rank
  if (p.ustats != nullptr) {
    const T* us = reinterpret_cast<const T*>(p.ustats);
    sred[0][0] = (double)__ldcg(&us[0]);
    sred[1][0] = (double)__ldcg(&us[1]);
    sred[2][0] = (double)__ldcg(&us[2]);
  }
  const int ii = st->iter;
  if (nflags != 0 || uflag != 0) {
    st->status = ST_NUMERR;  // previous (u, v) stay current
    st->niter = ii;
  } else {
    st->sel_u = sel_u ^ 1;
    st->sel_v = nxt;
    const double err_u = sred[0][0] / fmax(fmax(sred[1][0], sred[2][0]), 1.0);
    const double err_v = sred[3][0] / fmax(fmax(sred[4][0], sred[5][0]), 1.0);
    const double err = 0.5 * (err_u + err_v);
    p.err_hist[st->n_err] = err;
    st->n_err += 1;
    st->last_err = err;
    if (err < p.stop_thr) {
      st->status = ST_CONVERGED;
      st->niter = ii;
    } else if (ii + 1 >= p.max_iter) {
      st->status = ST_MAXITER;
      st->niter = ii;
    } else {
      st->iter = ii + 1;
    }
  }
  st->next_flags = 0;
  st->uflag = 0;
  st->ticket = 0;
  __threadfence();
}

// Row-sharded runs: collapse this rank's partial rows into the all-reduce buffer (m values + flag).
template <typename T>
__global__ void __launch_bounds__(256) reduce_partials_kernel(const void* partials, long long mp, int nsrc,
                                                              void* cbuf, long long m, SolveState* st,
                                                              int expect_status) {
  if (st->status != expect_status) return;
  const long long j = (long long)blockIdx.x * 256 + threadIdx.x;
  const T* P = reinterpret_cast<const T*>(partials);
  T* c = reinterpret_cast<T*>(cbuf);
  if (j < m) {
    T s = (T)0;
    for (int q = 0; q < nsrc; ++q) s += P[(long long)q * mp + j];
    c[j] = s;
  } else if (j == m) {
    c[m] = (T)(atomicOr(&st->uflag, 0) ? 1 : 0);
  }
}

template <typename T>
__global__ void export_kernel(const void* uring, const void* vring, long long n, long long m, long long mp,
                              const SolveState* st, void* u_out, void* v_out) {
  const T* u = reinterpret_cast<const T*>(uring) + (long long)st->sel_u * n;
  const T* v = reinterpret_cast<const T*>(vring) + (long long)st->sel_v * mp;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (u_out != nullptr && i < n) reinterpret_cast<T*>(u_out)[i] = u[i];
  if (v_out != nullptr && i < m) reinterpret_cast<T*>(v_out)[i] = v[i];
}

// ------------------------------------------------------------------------------------------------
// Elementwise: Gibbs kernel / scaled cost, plan, loss.
// ------------------------------------------------------------------------------------------------
template <typename T, int OP>  // OP 0: exp(M / (-reg)) ; 1: -M / reg
__global__ void __launch_bounds__(256) cost_transform_kernel(const T* __restrict__ M, T* __restrict__ out, long long n,
                                                             long long m, long long ldM, long long ldO, T reg) {
  const long long j0 = ((long long)blockIdx.y * 256 + threadIdx.x) * 4;
  for (long long i = blockIdx.x; i < n; i += gridDim.x) {
    const T* mr = M + i * ldM;
    T* orow = out + i * ldO;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long j = j0 + k;
      if (j < m) orow[j] = (OP == 0) ? exp_t(mr[j] / (-reg)) : (-mr[j]) / reg;
    }
  }
}

// 16-byte vector variants (rows and leading dimensions multiples of the vector width, 16-byte aligned bases): the scalar
// kernels above reach 55 % (transform) / 78 % (plan) of the HBM rate at 65536^2; two rows per trip keep >= 2 vector loads
// in flight per thread.  In-place transforms stay safe: every element is read and written by the same thread.
template <typename T, int OP>
__global__ void __launch_bounds__(256) cost_transform_vec_kernel(const T* M, T* out, long long n, long long m,
                                                                 long long ldM, long long ldO, T reg) {
  using V = typename Vec<T>::type;
  constexpr int VEC = Vec<T>::N;
  const long long nv = m / VEC;
  auto f = [&](const V& x) {
    V y;
#pragma unroll
    for (int e = 0; e < VEC; ++e) elem(y, e) = (OP == 0) ? exp_t(elem(x, e) / (-reg)) : (-elem(x, e)) / reg;
    return y;
  };
  for (long long i = blockIdx.y; i < n; i += 2LL * gridDim.y) {
    const long long i2 = i + gridDim.y;
    const V* r0 = reinterpret_cast<const V*>(M + i * ldM);
    V* o0 = reinterpret_cast<V*>(out + i * ldO);
    const V* r1 = reinterpret_cast<const V*>(M + (i2 < n ? i2 : i) * ldM);
    V* o1 = reinterpret_cast<V*>(out + (i2 < n ? i2 : i) * ldO);
    for (long long j = (long long)blockIdx.x * 256 + threadIdx.x; j < nv; j += (long long)gridDim.x * 256) {
      const V x0 = __ldcs(r0 + j);
      if (i2 < n) {
        const V x1 = __ldcs(r1 + j);
        __stcs(o0 + j, f(x0));
        __stcs(o1 + j, f(x1));
      } else {
        __stcs(o0 + j, f(x0));
      }
    }
  }
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) plan_vec_kernel(const T* __restrict__ K, const T* __restrict__ u,
                                                       const T* __restrict__ v, T* __restrict__ out, long long n,
                                                       long long m, long long ldK, long long ldO) {
  using V = typename Vec<T>::type;
  constexpr int VEC = Vec<T>::N;
  const long long nv = m / VEC;
  for (long long i = blockIdx.y; i < n; i += 2LL * gridDim.y) {
    const long long i2 = i + gridDim.y;
    const bool two = i2 < n;
    const T u0 = u[i], u1 = u[two ? i2 : i];
    const V* r0 = reinterpret_cast<const V*>(K + i * ldK);
    const V* r1 = reinterpret_cast<const V*>(K + (two ? i2 : i) * ldK);
    V* o0 = reinterpret_cast<V*>(out + i * ldO);
    V* o1 = reinterpret_cast<V*>(out + (two ? i2 : i) * ldO);
    for (long long j = (long long)blockIdx.x * 256 + threadIdx.x; j < nv; j += (long long)gridDim.x * 256) {
      const V vv = *reinterpret_cast<const V*>(v + j * VEC);
      const V x0 = __ldcs(r0 + j);
      V x1 = x0;
      if (two) x1 = __ldcs(r1 + j);
      V y0, y1;
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        elem(y0, e) = (MODE == MODE_KNOPP) ? (u0 * elem(x0, e)) * elem(vv, e) : exp_t((elem(x0, e) + u0) + elem(vv, e));
        elem(y1, e) = (MODE == MODE_KNOPP) ? (u1 * elem(x1, e)) * elem(vv, e) : exp_t((elem(x1, e) + u1) + elem(vv, e));
      }
      __stcs(o0 + j, y0);
      if (two) __stcs(o1 + j, y1);
    }
  }
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) plan_kernel(const T* __restrict__ K, const T* __restrict__ u,
                                                   const T* __restrict__ v, T* __restrict__ out, long long n,
                                                   long long m, long long ldK, long long ldO) {
  const long long j0 = ((long long)blockIdx.y * 256 + threadIdx.x) * 4;
  for (long long i = blockIdx.x; i < n; i += gridDim.x) {
    const T ui = u[i];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long j = j0 + k;
      if (j < m) {
        const T kij = K[i * ldK + j];
        out[i * ldO + j] = (MODE == MODE_KNOPP) ? (ui * kij) * v[j] : exp_t((kij + ui) + v[j]);
      }
    }
  }
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) loss_kernel(const T* __restrict__ K, const T* __restrict__ M,
                                                   const T* __restrict__ u, const T* __restrict__ v, long long n,
                                                   long long m, long long ldK, long long ldM, double* blk_out) {
  __shared__ double sred[256];
  double acc = 0.0;
  for (long long i = blockIdx.x; i < n; i += gridDim.x) {
    const T ui = u[i];
    for (long long j = threadIdx.x; j < m; j += 256) {
      const T kij = K[i * ldK + j];
      const T g = (MODE == MODE_KNOPP) ? (ui * kij) * v[j] : exp_t((kij + ui) + v[j]);
      acc += (double)(M[i * ldM + j] * g);
    }
  }
  sred[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) sred[threadIdx.x] += sred[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) blk_out[blockIdx.x] = sred[0];
}
__global__ void sum_blocks_kernel(const double* blk, int nblk, double* out) {
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

}  // namespace b2ot
