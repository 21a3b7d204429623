// Batched Sinkhorn, one WHOLE 256-column fp32 problem per CTA: half of the tile in registers, half in shared memory.
//
// batch_solve_kernel (batch_kernels.cuh) needs a cluster of 2 CTAs for a 256 x 256 fp32 problem (256 KiB > the 227 KiB
// of shared memory), and every iteration then pays a DSMEM exchange of the column totals between the two CTAs with
// nothing else resident on the SM to hide it (1 CTA per SM: the register file is full), plus 40 butterfly shuffles per
// group of four rows.  BASELINE config 5 ran at 3.2x its MUFU floor (log domain) and at 18 % of the FMA peak (scaling
// domain) that way (profiles/r01c_ncu_batch_256_*_regtile.txt).  This kernel keeps the problem on ONE SM:
//
//   * 512 threads; rows 0..127 live in registers for the whole solve (64 values per thread, as before), rows
//     128..n-1 (n = 192 or 256) stay in shared memory (<= 128 KiB) and are re-read once per iteration with 16-byte
//     loads (128 B/clk: 1024 clk per iteration, under the MUFU time of 4096 clk) -- no cluster, no DSMEM, two
//     __syncthreads per iteration;
//   * a lane owns columns {4 lane .. 4 lane + 3} and {128 + 4 lane .. + 3} (16-byte shared-memory accesses), a warp
//     takes four rows at a time; the four row sums are reduced with a TRANSPOSING butterfly (2 + 1 + 3 shuffles
//     instead of 4 x 5: after the first two steps every lane carries one row), the row factor comes back with one
//     shuffle per row;
//   * log domain: everything in log2 units, x2 = K log2e + v log2e (one packed FMA per two elements), and the
//     exponent reference of a row is the log-sum-exp the PREVIOUS iteration found for it (kept in shared memory), so
//     the hot loop has no max pass: e = 2^(x2 - lse_prev), s = sum e ~ 1.  A row group whose s leaves [2^-64, 2^64]
//     (first iteration of a launch, or a violent change of v) is redone with the exact row maximum (warp-uniform
//     branch), so the result is the same online log-sum-exp as before; 3 issue slots + 1 MUFU per element;
//   * the column update, the underflow fallback (exact two-pass column log-sum-exp), the batch-wide stop test and the
//     epilogue (plan, log-plan, <M, T>, entropy / KL term) follow batch_solve_kernel (reference: ot/batch/_utils.py:
//     46-290, ot/batch/_linear.py:317-345).
#pragma once

#include "batch_kernels.cuh"

namespace b2ot {

constexpr int HYB_M = 256;       // columns (exactly)
constexpr int HYB_NW = 16;       // warps
constexpr int HYB_NT = 512;      // threads
#ifndef HYB_RG_LOG
#define HYB_RG_LOG 4
#endif
constexpr int HYB_RREG = 128;    // rows held in registers: 2 trips x 16 warps x 4 rows
constexpr float HYB_LOG2E = 1.4426950408889634f;
constexpr float HYB_LN2 = 0.6931471805599453f;

__host__ __device__ inline size_t hyb_smem_bytes(int n) {
  // tile rows in shared memory (128 rows: the register half passes through it on the way in) + cross-warp column
  // partials + 5 column vectors (g, b, log b, candidate v, g log2e) + 4 row vectors (f, a, log a, lse2)
  return (size_t)128 * HYB_M * 4 + (size_t)HYB_NW * HYB_M * 4 + 5 * (size_t)HYB_M * 4 + 4 * (size_t)n * 4;
}
inline bool hyb_eligible(int dtype_is_f32, long long n, long long m) {
  return dtype_is_f32 && m == HYB_M && (n == 192 || n == 256);
}

// Transposing butterfly over RG rows (RG = 2 or 4): s[g] = this lane's partial of row g  ->  total of row
// hyb_gsel<RG>(lane), in every lane of that row's group of 32 / RG lanes.  RG = 4: 2 + 1 + 3 shuffles, RG = 2: 1 + 4.
template <int RG> __device__ __forceinline__ int hyb_gsel(int lane) { return RG == 4 ? (lane >> 3) & 3 : (lane >> 4) & 1; }
template <int RG, bool MAX>
__device__ __forceinline__ float hyb_reduce_rows(const float (&s)[RG], int lane) {
  auto op = [](float a, float b) { return MAX ? fmaxf(a, b) : a + b; };
  const bool b4 = (lane & 16) != 0;
  float v;
  if constexpr (RG == 4) {
    const bool b3 = (lane & 8) != 0;
    float t0 = b4 ? s[2] : s[0], t1 = b4 ? s[3] : s[1];
    const float o0 = b4 ? s[0] : s[2], o1 = b4 ? s[1] : s[3];
    t0 = op(t0, __shfl_xor_sync(0xffffffffu, o0, 16));
    t1 = op(t1, __shfl_xor_sync(0xffffffffu, o1, 16));
    v = b3 ? t1 : t0;
    const float o = b3 ? t0 : t1;
    v = op(v, __shfl_xor_sync(0xffffffffu, o, 8));
  } else {
    v = b4 ? s[1] : s[0];
    const float o = b4 ? s[0] : s[1];
    v = op(v, __shfl_xor_sync(0xffffffffu, o, 16));
    v = op(v, __shfl_xor_sync(0xffffffffu, v, 8));
  }
  v = op(v, __shfl_xor_sync(0xffffffffu, v, 4));
  v = op(v, __shfl_xor_sync(0xffffffffu, v, 2));
  v = op(v, __shfl_xor_sync(0xffffffffu, v, 1));
  return v;
}

// NS = trips (of 64 rows) that live in shared memory: n = 128 + 64 NS
template <int MODE, int NS>
__global__ void __launch_bounds__(HYB_NT, 1) batch_hyb_kernel(const BatchParams p) {
  // rows a warp takes together: 4 in the scaling domain; 2 in the log domain, where the exponentials of a group stay in
  // registers across the row reduction (4 rows x 8 columns on top of the 64 resident values spilled ~100 registers)
  constexpr int RG = (MODE == MODE_LOG) ? HYB_RG_LOG : 4;
  constexpr int TR = HYB_NW * RG;                 // rows per trip
  constexpr int M = HYB_M, NW = HYB_NW, NR = HYB_RREG / TR, N = 128 + 64 * NS, NTRIPS = N / TR;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t load_bar;
  __shared__ double red[HYB_NW][2];
  __shared__ int s_flag;

  if (p.expect_running && p.st->status != ST_RUNNING) return;  // sticky stop, uniform over the grid

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long prob = blockIdx.x;
  const int gsel = hyb_gsel<RG>(lane);           // row of a group this lane finishes
  const bool writer = (lane & (32 / RG - 1)) == 0;  // one lane per row writes the row's results
  constexpr int GL = 32 / RG;                    // lanes per row group: row g's results live in lanes [g GL, (g + 1) GL)

  float* Ks = reinterpret_cast<float*>(smem_raw);  // 128 x 256 (rows 128.. of the problem after the load phase)
  float* part = Ks + 128 * M;                      // [warp][256]
  float* gs = part + NW * M;                       // g / v
  float* bs = gs + M;
  float* lbs = bs + M;
  float* vtmp = lbs + M;
  float* gl2s = vtmp + M;                          // v log2e (log domain)
  float* fs = gl2s + M;                            // f / u                      [N]
  float* as = fs + N;
  float* las = as + N;
  float* lse2s = las + N;                          // log2-sum-exp of the previous iteration, per row

  const float* Xg = reinterpret_cast<const float*>(p.X) + (size_t)prob * N * M;
  const bool bulk_ok = (reinterpret_cast<uintptr_t>(Xg) & 15) == 0;
  const float reg = (float)p.reg;
  auto transform = [&](float x) -> float {
    return p.transform == 0 ? x : (p.transform == 1 ? expf((-x) / reg) : (-x) / reg);
  };

  // ---- load phase A: rows 0..127 through shared memory into registers ----
  if (bulk_ok) {
    if (tid == 0) {
      mbar_init(&load_bar, 1);
      fence_mbar_init();
      mbar_arrive_expect_tx(&load_bar, 128u * M * 4u);
      for (int o = 0; o < 4; ++o)
        bulk_g2s(reinterpret_cast<unsigned char*>(Ks) + o * 32768, reinterpret_cast<const unsigned char*>(Xg) + o * 32768,
                 32768u, &load_bar);
    }
  } else {
    for (int i = tid; i < 128 * M; i += HYB_NT) Ks[i] = Xg[i];
  }
  {
    const float* ag = reinterpret_cast<const float*>(p.a) + (size_t)prob * N;
    const float* bg = reinterpret_cast<const float*>(p.b) + (size_t)prob * M;
    const float* pbg = reinterpret_cast<const float*>(p.pb) + (size_t)prob * M;
    const float* pag = reinterpret_cast<const float*>(p.pa) + (size_t)prob * N;
    for (int r = tid; r < N; r += HYB_NT) {
      const float av = ag[r];
      as[r] = av;
      las[r] = logf(av);
      fs[r] = p.load_state ? pag[r] : ((MODE == MODE_LOG) ? 0.f : 1.f);
    }
    for (int j = tid; j < M; j += HYB_NT) {
      const float bv = bg[j];
      bs[j] = bv;
      lbs[j] = logf(bv);
      const float g0 = p.load_state ? pbg[j] : ((MODE == MODE_LOG) ? 0.f : 1.f);
      gs[j] = g0;
      gl2s[j] = g0 * HYB_LOG2E;
    }
  }
  __syncthreads();
  if (bulk_ok) mbar_wait(&load_bar, 0);
  float kr[NR][RG][8];  // this thread's share of rows 0..127, resident for the whole launch
#pragma unroll
  for (int t = 0; t < NR; ++t) {
#pragma unroll
    for (int g = 0; g < RG; ++g) {
      const int r = t * TR + warp * RG + g;
      const float4 lo = *reinterpret_cast<const float4*>(Ks + r * M + 4 * lane);
      const float4 hi = *reinterpret_cast<const float4*>(Ks + r * M + 128 + 4 * lane);
      kr[t][g][0] = transform(lo.x);  kr[t][g][1] = transform(lo.y);
      kr[t][g][2] = transform(lo.z);  kr[t][g][3] = transform(lo.w);
      kr[t][g][4] = transform(hi.x);  kr[t][g][5] = transform(hi.y);
      kr[t][g][6] = transform(hi.z);  kr[t][g][7] = transform(hi.w);
    }
  }
  __syncthreads();  // every thread has its registers: the staging rows may be overwritten
  // ---- load phase B: rows 128..N-1 stay in shared memory ----
  {
    const float* Xb = Xg + 128 * M;
    constexpr int NB = 64 * NS * M;
    if (bulk_ok) {
      if (tid == 0) {
        mbar_arrive_expect_tx(&load_bar, (uint32_t)NB * 4u);
        for (int o = 0; o < NB * 4; o += 32768)
          bulk_g2s(reinterpret_cast<unsigned char*>(Ks) + o, reinterpret_cast<const unsigned char*>(Xb) + o, 32768u,
                   &load_bar);
      }
      mbar_wait(&load_bar, 1);
    } else {
      for (int i = tid; i < NB; i += HYB_NT) Ks[i] = Xb[i];
      __syncthreads();
    }
    if (p.transform != 0) {
      for (int i = tid; i < NB; i += HYB_NT) Ks[i] = transform(Ks[i]);
    }
    __syncthreads();
  }

  // k[g][c] of trip `trip`: registers (trip < 2) or shared memory
  auto load_trip = [&](int trip, float (&k)[RG][8]) {
#pragma unroll
    for (int g = 0; g < RG; ++g) {
      if (trip < NR) {
#pragma unroll
        for (int c = 0; c < 8; ++c) k[g][c] = kr[trip < NR ? trip : 0][g][c];
      } else {
        const int r = (trip - NR) * TR + warp * RG + g;
        const float4 lo = *reinterpret_cast<const float4*>(Ks + r * M + 4 * lane);
        const float4 hi = *reinterpret_cast<const float4*>(Ks + r * M + 128 + 4 * lane);
        k[g][0] = lo.x;  k[g][1] = lo.y;  k[g][2] = lo.z;  k[g][3] = lo.w;
        k[g][4] = hi.x;  k[g][5] = hi.y;  k[g][6] = hi.z;  k[g][7] = hi.w;
      }
    }
  };
  auto load_cols = [&](const float* v, float (&o)[8]) {
    const float4 lo = *reinterpret_cast<const float4*>(v + 4 * lane);
    const float4 hi = *reinterpret_cast<const float4*>(v + 128 + 4 * lane);
    o[0] = lo.x;  o[1] = lo.y;  o[2] = lo.z;  o[3] = lo.w;
    o[4] = hi.x;  o[5] = hi.y;  o[6] = hi.z;  o[7] = hi.w;
  };
  auto store_part = [&](const float (&acc)[8]) {
    *reinterpret_cast<float4*>(part + warp * M + 4 * lane) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(part + warp * M + 128 + 4 * lane) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  };
  auto selg = [&](const float (&v)[RG]) -> float {
    float r = v[0];
#pragma unroll
    for (int g = 1; g < RG; ++g) r = (gsel == g) ? v[g] : r;
    return r;
  };

  // CTA-wide reduction of two doubles (every thread gets the totals)
  auto reduce2 = [&](double x0, double x1, double& o0, double& o1) {
    x0 = warp_sum_d(x0);
    x1 = warp_sum_d(x1);
    __syncthreads();
    if (lane == 0) {
      red[warp][0] = x0;
      red[warp][1] = x1;
    }
    __syncthreads();
    double t0 = 0.0, t1 = 0.0;
    for (int w = 0; w < NW; ++w) {
      t0 += red[w][0];
      t1 += red[w][1];
    }
    o0 = t0;
    o1 = t1;
  };

  // Exact column log-sum-exp (two passes over the tile): v_j = log b_j - LSE_i(K_ij + u_i).  Rare (underflow of a
  // row-frame column total, or the test hook).
  auto exact_columns = [&]() {
    float cm[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) cm[c] = -CUDART_INF_F;
#pragma unroll
    for (int trip = 0; trip < NTRIPS; ++trip) {
      float k[RG][8];
      load_trip(trip, k);
#pragma unroll
      for (int g = 0; g < RG; ++g) {
        const float u = fs[trip * TR + warp * RG + g];
#pragma unroll
        for (int c = 0; c < 8; ++c) cm[c] = fmaxf(cm[c], k[g][c] + u);
      }
    }
    store_part(cm);
    __syncthreads();
    if (tid < M) {
      float mx = -CUDART_INF_F;
      for (int w = 0; w < NW; ++w) mx = fmaxf(mx, part[w * M + tid]);
      vtmp[tid] = mx;
    }
    __syncthreads();
    float cmx[8], acc[8];
    load_cols(vtmp, cmx);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      acc[c] = 0.f;
      cmx[c] = (cmx[c] == -CUDART_INF_F) ? 0.f : cmx[c];
    }
#pragma unroll
    for (int trip = 0; trip < NTRIPS; ++trip) {
      float k[RG][8];
      load_trip(trip, k);
#pragma unroll
      for (int g = 0; g < RG; ++g) {
        const float u = fs[trip * TR + warp * RG + g];
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] += expf((k[g][c] + u) - cmx[c]);
      }
    }
    __syncthreads();  // vtmp / part of the first pass consumed
    store_part(acc);
    __syncthreads();
    if (tid < M) {
      float S = 0.f;
      for (int w = 0; w < NW; ++w) S += part[w * M + tid];
      const float mx = vtmp[tid];
      const float v = lbs[tid] - (mx + logf(S));
      gs[tid] = v;
      gl2s[tid] = v * HYB_LOG2E;
    }
    __syncthreads();
  };

  // ------------------------------------------------------------------------------------------
  // iterations (f then g, ot/batch/_utils.py:141-143, :268-270)
  //
  // The register trips are unrolled (static register indices), the shared-memory trips run in a rolled loop: the hot
  // loop stays a few KB of code (the first version unrolled everything, 8 trip bodies with the exact-maximum redo
  // inlined in each: 25 k static instructions and 14-20 % of the stall samples on instruction fetch, ncu
  // profiles/r02g_ncu_batch_hyb_*.txt).  Per-row scalar work (1 / s, log2 s) uses the MUFU approximations: it is
  // executed by all 32 lanes for RG rows, so the exact division / logf (10 instructions each) weighed as much as the
  // RG x 8 elements of useful work they follow.
  // ------------------------------------------------------------------------------------------
  auto rcp_fast = [](float x) -> float {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  };
  for (int it = p.it_begin; it < p.it_end; ++it) {
    float acc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[c] = 0.f;
    if (MODE == MODE_KNOPP) {
      float greg[8];
      load_cols(gs, greg);
      auto knopp_trip = [&](int rb, float (&k)[RG][8]) {
        float s[RG];
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int c = 0; c < 8; c += 2) fma2(s0, s1, k[g][c], k[g][c + 1], greg[c], greg[c + 1], s0, s1);
          s[g] = s0 + s1;
        }
        const float srow = hyb_reduce_rows<RG, false>(s, lane);
        float f = as[rb + gsel] * rcp_fast(srow);
        f = fmaf(fmaf(-f, srow, as[rb + gsel]), rcp_fast(srow), f);  // one Newton step: a / s to ~1 ulp
        if (writer) fs[rb + gsel] = f;
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          const float fg = __shfl_sync(0xffffffffu, f, g * GL);
#pragma unroll
          for (int c = 0; c < 8; c += 2) fma2(acc[c], acc[c + 1], k[g][c], k[g][c + 1], fg, fg, acc[c], acc[c + 1]);
        }
      };
#pragma unroll
      for (int trip = 0; trip < NR; ++trip) {
        float k[RG][8];
        load_trip(trip, k);
        knopp_trip(trip * TR + warp * RG, k);
      }
#pragma unroll 1
      for (int trip = NR; trip < NTRIPS; ++trip) {
        float k[RG][8];
        load_trip(trip, k);
        knopp_trip(trip * TR + warp * RG, k);
      }
      store_part(acc);
      __syncthreads();
      if (tid < M) {
        float tot = 0.f;
#pragma unroll 4
        for (int w = 0; w < NW; ++w) tot += part[w * M + tid];
        gs[tid] = bs[tid] / tot;
      }
      __syncthreads();
    } else {
      // One row pass.  EXACT = false: exponent reference = last iteration's row log-sum-exp (no max pass); a row whose
      // sum leaves [2^-64, 2^64] votes for a redo of the WHOLE pass with the exact row maxima (rare: first iteration
      // of a launch takes the exact pass directly, later ones only after a violent change of v).
      int redo = 0;
      auto log_trip = [&](const bool EXACT, int rb, float (&e)[RG][8]) {  // EXACT is a literal at every (inlined) call
        float gl2[8], ref2[RG], s[RG];
        load_cols(gl2s, gl2);
#pragma unroll
        for (int g = 0; g < RG; ++g)
#pragma unroll
          for (int c = 0; c < 8; c += 2)
            fma2(e[g][c], e[g][c + 1], e[g][c], e[g][c + 1], HYB_LOG2E, HYB_LOG2E, gl2[c], gl2[c + 1]);  // x2 = K log2e + v log2e
        if (EXACT) {
          float mx[RG];
#pragma unroll
          for (int g = 0; g < RG; ++g) {
            float m0 = fmaxf(e[g][0], e[g][1]);
#pragma unroll
            for (int c = 2; c < 8; c += 2) m0 = max3(m0, e[g][c], e[g][c + 1]);
            mx[g] = m0;
          }
          float mrow = hyb_reduce_rows<RG, true>(mx, lane);
          mrow = (mrow == -CUDART_INF_F) ? 0.f : mrow;
#pragma unroll
          for (int g = 0; g < RG; ++g) ref2[g] = __shfl_sync(0xffffffffu, mrow, g * GL);
        } else {
#pragma unroll
          for (int g = 0; g < RG; ++g) ref2[g] = lse2s[rb + g];
        }
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          const float nr = -ref2[g];
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int c = 0; c < 8; c += 2) {
            add2(e[g][c], e[g][c + 1], nr, nr);
            e[g][c] = ex2_ftz(e[g][c]);
            e[g][c + 1] = ex2_ftz(e[g][c + 1]);
            add2(s0, s1, e[g][c], e[g][c + 1]);
          }
          s[g] = s0 + s1;
        }
        const float srow = hyb_reduce_rows<RG, false>(s, lane);
        if (!EXACT) redo |= !(srow > 5.5e-20f && srow < 1.8e19f);
        const int r = rb + gsel;
        const float lse2 = selg(ref2) + lg2_ftz(srow);
        const float w = as[r] * rcp_fast(srow);          // a_i / s_i in this row's exponent frame
        if (writer) {
          fs[r] = fmaf(-lse2, HYB_LN2, las[r]);          // u_i = log a_i - LSE_j(K_ij + v_j)
          lse2s[r] = lse2;
        }
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          const float wg = __shfl_sync(0xffffffffu, w, g * GL);
#pragma unroll
          for (int c = 0; c < 8; c += 2) fma2(acc[c], acc[c + 1], e[g][c], e[g][c + 1], wg, wg, acc[c], acc[c + 1]);
        }
      };
      auto row_pass = [&](const bool exact_tag) {
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[c] = 0.f;
#pragma unroll
        for (int trip = 0; trip < NR; ++trip) {
          float e[RG][8];
          load_trip(trip, e);
          log_trip(exact_tag, trip * TR + warp * RG, e);
        }
#pragma unroll 1
        for (int trip = NR; trip < NTRIPS; ++trip) {
          float e[RG][8];
          load_trip(trip, e);
          log_trip(exact_tag, trip * TR + warp * RG, e);
        }
      };
      if (it > p.it_begin) {
        row_pass(false);
        if (__syncthreads_or(redo)) row_pass(true);  // (all warps have left the stale pass: fs / lse2s are rewritten)
      } else {
        row_pass(true);
      }
      store_part(acc);
      __syncthreads();
      int bad = p.force_exact;
      if (tid < M) {
        float tot = 0.f;
#pragma unroll 4
        for (int w = 0; w < NW; ++w) tot += part[w * M + tid];
        // tot = exp(v_j) * sum_i exp(K_ij + u_i): the column marginal of the current plan
        if (!(tot >= tiny_guard<float>())) bad = 1;
        vtmp[tid] = gs[tid] + (lbs[tid] - logf(tot));
      }
      bad = __syncthreads_or(bad);
      if (!bad) {
        if (tid < M) {
          const float v = vtmp[tid];
          gs[tid] = v;
          gl2s[tid] = v * HYB_LOG2E;
        }
        __syncthreads();
      } else {
        exact_columns();
      }
    }
  }

  // row sums of the current plan for the rows of one trip: T_ij = (K f) g  /  exp((K + u) + v)
  auto plan_trip = [&](int trip, const float (&greg)[8], float (&t)[RG][8]) {
    load_trip(trip, t);
#pragma unroll
    for (int g = 0; g < RG; ++g) {
      const float fi = fs[trip * TR + warp * RG + g];
#pragma unroll
      for (int c = 0; c < 8; ++c)
        t[g][c] = (MODE == MODE_KNOPP) ? (t[g][c] * fi) * greg[c] : expf((t[g][c] + fi) + greg[c]);
    }
  };

  // ------------------------------------------------------------------------------------------
  // stop test: L2 violation of the row marginal (with the g / v just produced), max over the batch
  // ------------------------------------------------------------------------------------------
  if (p.check) {
    double e2 = 0.0;
    float greg[8];
    load_cols(gs, greg);
#pragma unroll  // (static trip index: kr must stay in registers)
    for (int trip = 0; trip < NTRIPS; ++trip) {
      float t[RG][8], s[RG];
      plan_trip(trip, greg, t);
#pragma unroll
      for (int g = 0; g < RG; ++g) {
        float s0 = 0.f;
#pragma unroll
        for (int c = 0; c < 8; ++c) s0 += t[g][c];
        s[g] = s0;
      }
      const float srow = hyb_reduce_rows<RG, false>(s, lane);
      const float d = srow - as[trip * TR + warp * RG + gsel];
      if (writer) e2 += (double)(d * d);
    }
    double t0, t1;
    reduce2(e2, 0.0, t0, t1);
    if (tid == 0) {
      p.err_b[prob] = (double)(float)sqrt((double)(float)t0);
      __threadfence();
      const unsigned int tk = atomicAdd(&p.st->ticket, 1u);
      s_flag = (tk == (unsigned int)(p.B - 1)) ? 1 : 0;
    }
    __syncthreads();
    if (s_flag) {  // last problem of the batch: the reference's err = max over problems
      __threadfence();
      double mxe = 0.0;
      bool nan = false;
      for (long long i = tid; i < p.B; i += HYB_NT) {
        const double e = __ldcg(&p.err_b[i]);
        if (e != e) nan = true;
        mxe = fmax(mxe, e);
      }
      if (nan) mxe = CUDART_NAN;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double y = __shfl_xor_sync(0xffffffffu, mxe, o);
        mxe = (mxe != mxe || y != y) ? CUDART_NAN : fmax(mxe, y);
      }
      __syncthreads();
      if (lane == 0) red[warp][0] = mxe;
      __syncthreads();
      if (tid == 0) {
        double e = 0.0;
        for (int w = 0; w < NW; ++w) {
          const double y = red[w][0];
          e = (e != e || y != y) ? CUDART_NAN : fmax(e, y);
        }
        p.st->max_err = e;
        if (e < p.tol) {
          p.st->status = ST_CONVERGED;
          p.st->n_iters = p.it_end - 1;
        }
        p.st->ticket = 0;
        __threadfence();
      }
    }
  }

  // ---- potentials out (state for the next launch) ----
  {
    float* pag = reinterpret_cast<float*>(p.pa) + (size_t)prob * N;
    for (int r = tid; r < N; r += HYB_NT) pag[r] = fs[r];
    float* pbg = reinterpret_cast<float*>(p.pb) + (size_t)prob * M;
    for (int j = tid; j < M; j += HYB_NT) pbg[j] = gs[j];
  }

  // ------------------------------------------------------------------------------------------
  // epilogue: plan, <M, T>, regularisation term  (ot/batch/_linear.py:338-345, _utils.py:4-22)
  // ------------------------------------------------------------------------------------------
  if (p.epilogue) {
    double vl = 0.0, vr = 0.0;
    float* Tg = p.T_out ? reinterpret_cast<float*>(p.T_out) + (size_t)prob * N * M : nullptr;
    float* Lg = p.logT_out ? reinterpret_cast<float*>(p.logT_out) + (size_t)prob * N * M : nullptr;
    const float* Mg = p.Mval ? reinterpret_cast<const float*>(p.Mval) + (size_t)prob * N * M : nullptr;
    float greg[8], breg[8];
    load_cols(gs, greg);
    load_cols(bs, breg);
    const bool vec_ok = ((Tg == nullptr || (reinterpret_cast<uintptr_t>(Tg) & 15) == 0) &&
                         (Lg == nullptr || (reinterpret_cast<uintptr_t>(Lg) & 15) == 0) &&
                         (Mg == nullptr || (reinterpret_cast<uintptr_t>(Mg) & 15) == 0));
#pragma unroll  // (static trip index: kr must stay in registers)
    for (int trip = 0; trip < NTRIPS; ++trip) {
      float k[RG][8];
      load_trip(trip, k);
#pragma unroll
      for (int g = 0; g < RG; ++g) {
        const int r = trip * TR + warp * RG + g;
        const float fi = fs[r], ai = as[r];
        float t[8], lt[8], mv[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          if (MODE == MODE_KNOPP) {
            t[c] = (k[g][c] * fi) * greg[c];
            lt[c] = 0.f;
          } else {
            lt[c] = (k[g][c] + fi) + greg[c];
            t[c] = expf(lt[c]);
          }
          mv[c] = 0.f;
        }
        const size_t base = (size_t)r * M + 4 * lane;
        if (vec_ok) {
          if (Tg) {
            *reinterpret_cast<float4*>(Tg + base) = make_float4(t[0], t[1], t[2], t[3]);
            *reinterpret_cast<float4*>(Tg + base + 128) = make_float4(t[4], t[5], t[6], t[7]);
          }
          if (Lg) {
            *reinterpret_cast<float4*>(Lg + base) = make_float4(lt[0], lt[1], lt[2], lt[3]);
            *reinterpret_cast<float4*>(Lg + base + 128) = make_float4(lt[4], lt[5], lt[6], lt[7]);
          }
          if (Mg) {
            const float4 m0 = __ldcs(reinterpret_cast<const float4*>(Mg + base));
            const float4 m1 = __ldcs(reinterpret_cast<const float4*>(Mg + base + 128));
            mv[0] = m0.x;  mv[1] = m0.y;  mv[2] = m0.z;  mv[3] = m0.w;
            mv[4] = m1.x;  mv[5] = m1.y;  mv[6] = m1.z;  mv[7] = m1.w;
          }
        } else {
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const size_t idx = base + (c & 3) + (c >> 2) * 128;
            if (Tg) Tg[idx] = t[c];
            if (Lg) Lg[idx] = lt[c];
            if (Mg) mv[c] = Mg[idx];
          }
        }
        // per-row partials in fp32 (8 terms), folded into the double accumulators once per row
        if (Mg) {
          float sl = 0.f;
#pragma unroll
          for (int c = 0; c < 8; ++c) sl = fmaf(mv[c], t[c], sl);
          vl += (double)sl;
        }
        if (p.reg_type != 0) {
          float sr = 0.f;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const float arg = (p.reg_type == 1) ? t[c] + 1e-16f : t[c] / (ai * breg[c]) + 1e-16f;
            sr = fmaf(t[c], logf(arg), sr);
          }
          vr += (double)sr;
        }
      }
    }
    double t0, t1;
    reduce2(vl, vr, t0, t1);
    if (tid == 0) {
      if (p.value_linear) p.value_linear[prob] = t0;
      if (p.value_reg) p.value_reg[prob] = t1;
    }
  }
}

}  // namespace b2ot
