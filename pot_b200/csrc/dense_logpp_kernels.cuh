// Log-domain fused sweep, "ping-pong" variant (fp32, sm_100a).
//
// Same contract as fused_sweep_kernel<float, ., MODE_LOG, ...> (one pass over Mr = -M/reg per iteration, row pass =
// (max, sum-exp), row-frame exponentials re-used for the column sums; ot/bregman/_sinkhorn.py:881-889), different
// schedule.  The per-row-group work of the log domain is a chain of phases that use different units
// (FMA/ALU: x = Mr + v and the row maximum; shuffles; MUFU: 2^x; DSMEM: cluster exchange; FMA: column update) and
// in the one-CTA-in-lockstep kernel every warp of the SM is in the same phase at the same time (ncu: issue slots
// 38 % busy, MUFU 24 %, DRAM 61 %).  Here the CTA is split into TWO independent halves of 4 warps ("virtual CTAs"):
// half h owns the rows g = 2 cluster_id + h (mod 2 n_clusters) over the full column slice of the CTA, with its own
// TMA stage ring, its own exchange barriers and its own column accumulators.  Each SM sub-partition hosts one warp
// of each half, so one half's MUFU phase or exchange wait overlaps the other half's FMA phases.
//   * one row per group and half (64 K values per thread at a slice of 8192 columns), v kept in shared memory
//     (read-only, shared by both halves) to make room for 64 accumulator registers per thread;
//   * the halves' column accumulators are added through shared memory once per sweep, so the per-cluster partial
//     layout seen by finalize_kernel is unchanged.
#pragma once

#include "common.cuh"
#include "dense_kernels.cuh"

namespace b2ot {

constexpr int PP_NH = 2;      // halves per CTA
constexpr int PP_NWH = 4;     // warps per half
constexpr int PP_NTH = 128;   // threads per half
constexpr int PP_SMAX = 4;    // max stages per half

__host__ __device__ constexpr int pp_wcap(int vpt) { return PP_NTH * vpt * 4; }
inline size_t pp_smem_bytes(int vpt, int stages) {
  return (size_t)(PP_NH * stages + 1) * pp_wcap(vpt) * sizeof(float);  // stage rings + the v slice
}

// OCC = 2: two CTAs per SM (narrow slices, VPT <= 8): four independent halves per SM instead of two
template <int VPT, int OCC = 1>
__global__ void __launch_bounds__(PP_NH* PP_NTH, OCC) log_pp_sweep_kernel(const SweepParams p) {
  constexpr int WCAP = PP_NTH * VPT * 4;
  constexpr float L2E = 1.4426950408889634f;

  __shared__ __align__(16) uint64_t full_bar[PP_NH][PP_SMAX];
  __shared__ __align__(16) uint64_t exch_bar[PP_NH][2];
  __shared__ __align__(16) float exch[PP_NH][2][CSMAX * PP_NWH][2];  // [half][parity][source warp][max, sum]
  extern __shared__ __align__(128) unsigned char dyn_smem[];
  float* const stage_base = reinterpret_cast<float*>(dyn_smem);
  const int S = p.stages;
  float* const vs = stage_base + (size_t)PP_NH * S * WCAP;  // v slice of this CTA

  const int tid = threadIdx.x;
  const int hf = tid >> 7;          // half
  const int th = tid & (PP_NTH - 1);
  const int wh = th >> 5;           // warp within the half
  const int lane = tid & 31;
  const uint32_t crank = cluster_ctarank();
  const uint32_t csize = cluster_nctarank();
  const long long cid = cluster_id_x();
  const long long ncl = cluster_count_x();
  const long long col0 = (long long)crank * p.W;
  const long long wrem = p.m - col0;
  const int wc = wrem <= 0 ? 0 : (wrem > p.W ? p.W : (int)wrem);
  const long long NG = p.n;                         // one row per group
  const long long g0 = cid * PP_NH + hf;            // first row of this half
  const long long gstride = ncl * PP_NH;
  const float* const Kg = reinterpret_cast<const float*>(p.K);
  const uint32_t row_bytes = (uint32_t)wc * 4u;
  const uint64_t pol = p.use_hint ? l2_policy_evict_first() : 0ull;
  const int nsrc = (int)csize * PP_NWH;             // warps of this half across the cluster (<= 64)
  const uint32_t exch_bytes = (uint32_t)nsrc * 2u * 4u;
  float* const my_stage = stage_base + (size_t)hf * S * WCAP;

  auto issue_row = [&](long long g, int s) {
    if (row_bytes == 0) {
      mbar_arrive(&full_bar[hf][s]);
      return;
    }
    mbar_arrive_expect_tx(&full_bar[hf][s], row_bytes);
    float* dst = my_stage + (size_t)s * WCAP;
    const float* src = Kg + g * p.ld + col0;
    if (p.use_hint)
      bulk_g2s_hint(dst, src, row_bytes, &full_bar[hf][s], pol);
    else
      bulk_g2s(dst, src, row_bytes, &full_bar[hf][s]);
  };

  if (tid == 0) {
    for (int h = 0; h < PP_NH; ++h) {
      for (int s = 0; s < S; ++s) mbar_init(&full_bar[h][s], 1);
      mbar_init(&exch_bar[h][0], 1);
      mbar_init(&exch_bar[h][1], 1);
    }
    fence_mbar_init();
  }
  __syncthreads();
  if (th == 0) {  // each half's producer fills its ring (K does not depend on the previous kernel)
    for (int s = 0; s < S; ++s) {
      const long long g = g0 + (long long)s * gstride;
      if (g < NG) issue_row(g, s);
    }
  }
  cluster_sync_all();  // every CTA's barriers exist before any remote put

  griddep_launch_dependents();
  griddep_wait();
  if (p.st->status != ST_RUNNING) {  // sticky stop: drain the copies in flight and leave
    for (int s = 0; s < S; ++s) {
      const long long g = g0 + (long long)s * gstride;
      if (g < NG) mbar_wait(&full_bar[hf][s], 0);
    }
    return;
  }
  const int sel_u = p.st->sel_u;
  const int sel_v = p.st->sel_v;
  const float* vcur = reinterpret_cast<const float*>(p.vring) + (long long)sel_v * p.mp + col0;
  float* unew = reinterpret_cast<float*>(p.uring) + (long long)(sel_u ^ 1) * p.n;
  const float* ag = reinterpret_cast<const float*>(p.a);

  for (int e = tid * 4; e < WCAP; e += PP_NH * PP_NTH * 4) {
    float4 v4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e < wc) v4 = *reinterpret_cast<const float4*>(vcur + e);
    *reinterpret_cast<float4*>(vs + e) = v4;
  }
  __syncthreads();

  float4 acc[VPT];
#pragma unroll
  for (int k = 0; k < VPT; ++k) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool full_cols = ((VPT - 1) * PP_NTH + th) * 4 + 4 <= wc;
  const uint32_t my_slot = crank * PP_NWH + (uint32_t)wh;
  int s = 0;
  uint32_t ph = 0;
  long long it = 0;
  for (long long g = g0; g < NG; g += gstride, ++it) {
    const int par = (int)(it & 1);
    const uint32_t eph = (uint32_t)((it >> 1) & 1);
    if (th == 0) mbar_arrive_expect_tx(&exch_bar[hf][par], exch_bytes);  // arm this row's exchange
    const float a_r = ag[g];

    // ---- row slice: shared -> registers (frees the stage as soon as every warp of the half has it) ----
    mbar_wait(&full_bar[hf][s], ph);
    float4 kreg[VPT];
    const float* srow = my_stage + (size_t)s * WCAP;
    if (full_cols) {
#pragma unroll
      for (int k = 0; k < VPT; ++k) kreg[k] = *reinterpret_cast<const float4*>(srow + (k * PP_NTH + th) * 4);
    } else {
#pragma unroll
      for (int k = 0; k < VPT; ++k) {
        const int e = (k * PP_NTH + th) * 4;
        if (e < wc)
          kreg[k] = *reinterpret_cast<const float4*>(srow + e);
        else
          kreg[k] = make_float4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
      }
    }

    // ---- x = Mr + v, row maximum (4 chains), e = 2^((x - max) log2e), row sum ----
    float m0 = -CUDART_INF_F, m1 = m0, m2 = m0, m3 = m0;
#pragma unroll
    for (int k = 0; k < VPT; k += 2) {
      const float4 w0 = *reinterpret_cast<const float4*>(vs + (k * PP_NTH + th) * 4);
      const float4 w1 = *reinterpret_cast<const float4*>(vs + ((k + 1) * PP_NTH + th) * 4);
      add2(kreg[k].x, kreg[k].y, w0.x, w0.y);
      add2(kreg[k].z, kreg[k].w, w0.z, w0.w);
      add2(kreg[k + 1].x, kreg[k + 1].y, w1.x, w1.y);
      add2(kreg[k + 1].z, kreg[k + 1].w, w1.z, w1.w);
      m0 = max3(m0, kreg[k].x, kreg[k + 1].x);
      m1 = max3(m1, kreg[k].y, kreg[k + 1].y);
      m2 = max3(m2, kreg[k].z, kreg[k + 1].z);
      m3 = max3(m3, kreg[k].w, kreg[k + 1].w);
    }
    const float mx = warp_max_redux(fmaxf(max3(m0, m1, m2), m3));
    const float ref = (mx == -CUDART_INF_F) ? 0.f : mx;
    const float nref = -ref * L2E;
    // the rounding of ref * log2e is a common factor 2^lo of the whole row: divided back out of the sum and of f_i
    const float corr = 1.f - fmaf(ref, L2E, nref) * 0.6931471805599453f;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      float t0, t1, t2, t3;
      fma2(t0, t1, kreg[k].x, kreg[k].y, L2E, L2E, nref, nref);
      fma2(t2, t3, kreg[k].z, kreg[k].w, L2E, L2E, nref, nref);
      kreg[k].x = ex2_ftz(t0);
      kreg[k].y = ex2_ftz(t1);
      kreg[k].z = ex2_ftz(t2);
      kreg[k].w = ex2_ftz(t3);
      add2(s0, s1, kreg[k].x, kreg[k].y);
      add2(s2, s3, kreg[k].z, kreg[k].w);
    }
    const float wsum = warp_sum((s0 + s1) + (s2 + s3)) * corr;

    // ---- one-sided put of (max, sum) into every CTA of the cluster, then wait for all warps of this half ----
    if (lane < (int)csize) {
      const uint32_t dst = mapa_u32(smem_u32(&exch[hf][par][my_slot][0]), (uint32_t)lane);
      const uint32_t bar = mapa_u32(smem_u32(&exch_bar[hf][par]), (uint32_t)lane);
      st_async_v2(dst, mx, wsum, bar);
    }
    mbar_wait(&exch_bar[hf][par], eph);
    if (th == 0) {  // completion also proves that the 4 warps of this half have drained stage s: refill it
      const long long gn = g + (long long)S * gstride;
      if (gn < NG) issue_row(gn, s);
    }
    // every lane reads source warp (lane mod nsrc): all groups of nsrc lanes hold the same data, so log2(nsrc)
    // butterfly rounds leave the total in every lane
    // (clusters of 16 have 64 source warps: every lane then folds sources lane and lane + 32 first)
    const int nl = nsrc < 32 ? nsrc : 32;
    const int srcw = lane & (nl - 1);
    float ms = exch[hf][par][srcw][0];
    float sm = exch[hf][par][srcw][1];
    if (nsrc > 32) {
      const float ms2 = exch[hf][par][srcw + 32][0];
      const float sm2 = exch[hf][par][srcw + 32][1];
      const float mm = fmaxf(ms, ms2);
      const float t1 = (ms != -CUDART_INF_F) ? sm * ex2_ftz((ms - mm) * L2E) : 0.f;
      const float t2 = (ms2 != -CUDART_INF_F) ? sm2 * ex2_ftz((ms2 - mm) * L2E) : 0.f;
      ms = mm;
      sm = t1 + t2;
    }
    const float M = warp_max_redux(ms);
    float Ssum = (ms != -CUDART_INF_F) ? sm * ex2_ftz((ms - M) * L2E) : 0.f;
    for (int o = nl >> 1; o > 0; o >>= 1) Ssum += __shfl_xor_sync(0xffffffffu, Ssum, o);
    const float u = a_r - fmaf(lg2_ftz(Ssum), 0.6931471805599453f, M);
    const float ur = ex2_ftz((ref + u) * L2E) * corr;  // row-frame factor f_i for this warp's reference
    if (crank == 0 && th == 0) unew[g] = u;

    // ---- column accumulators: sum_i exp(Mr_ij + u_i + v_j) ----
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      fma2(acc[k].x, acc[k].y, kreg[k].x, kreg[k].y, ur, ur, acc[k].x, acc[k].y);
      fma2(acc[k].z, acc[k].w, kreg[k].z, kreg[k].w, ur, ur, acc[k].z, acc[k].w);
    }
    if (++s == S) {
      s = 0;
      ph ^= 1u;
    }
  }

  // ---- add the halves' column sums through shared memory and flush this cluster's partial row ----
  __syncthreads();  // both rings idle
  float* xbuf = stage_base;
  if (hf == 1) {
#pragma unroll
    for (int k = 0; k < VPT; ++k) *reinterpret_cast<float4*>(xbuf + (k * PP_NTH + th) * 4) = acc[k];
  }
  __syncthreads();
  if (hf == 0) {
    float* pout = reinterpret_cast<float*>(p.partials) + cid * p.mp + col0;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
      const int e = (k * PP_NTH + th) * 4;
      if (e < wc) {
        const float4 o = *reinterpret_cast<const float4*>(xbuf + e);
        float4 r = acc[k];
        r.x += o.x;
        r.y += o.y;
        r.z += o.z;
        r.w += o.w;
        *reinterpret_cast<float4*>(pout + e) = r;
      }
    }
  }
  cluster_sync_all();  // no CTA leaves while a peer could still address its shared memory
}

}  // namespace b2ot
