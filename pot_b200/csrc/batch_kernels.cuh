// Batched Sinkhorn (ot.batch): every problem of the batch is solved ON CHIP, sm_100a.
//
// One thread-block cluster of CS CTAs owns one problem of the batch for a whole launch: CTA c keeps
// rows [c*RPC, (c+1)*RPC) of the n x m matrix in shared memory (loaded once with a TMA bulk copy, the
// Gibbs transform exp(-M/reg) or -M/reg applied in place), so HBM sees the cost once on the way in and the
// plan once on the way out while the iterations run from shared memory and registers.
//
// Iteration (reference order "f then g", ot/batch/_utils.py:141-143 and :268-270), ONE pass over the
// resident tile: a warp pulls a group of rows into registers, forms the row sums (K g)_i with a butterfly
// shuffle, f_i = a_i / (K g)_i, and immediately accumulates f_i K_ij into per-lane column accumulators;
// the column sums are combined across warps through shared memory and across the CTAs of the cluster
// through DSMEM, giving g_j = b_j / (K^T f)_j.  Log domain: the row pass is a (max, sum-exp) reduction;
// the row-frame exponentials e_ij = exp(K_ij + v_j - max_i) are re-used for the column update
// (sum_i e_ij a_i / s_i = exp(v_j) sum_i exp(K_ij + u_i)), with an exact online-max column pass as the
// fallback when such a sum underflows.
//
// The stop test of the reference couples the whole batch (max over problems of the L2 row-marginal
// violation, every 10th iteration, :144-149 / :273-278).  It is evaluated on the device: each cluster
// writes its problem's violation, the last one to finish takes the maximum and flips a sticky status word.
#pragma once

#include "common.cuh"

namespace b2ot {

// threads per CTA: 16 warps for narrow problems, 8 warps when a lane owns >= 16 columns (halves the
// cross-warp column scratch, which scales with warps x padded columns)
__host__ __device__ constexpr int batch_threads(int cpl) { return cpl >= 16 ? 256 : 512; }
constexpr int BATCH_NWMAX = 16;
constexpr int BATCH_CSMAX = 8;

struct __align__(16) BatchState {
  int status;        // ST_RUNNING / ST_CONVERGED
  int n_iters;       // iteration index at which the stop test fired
  unsigned int ticket;
  int pad;
  double max_err;
};

struct BatchParams {
  const void* X;         // (B, n, m): cost M (transform 1, 2) or the kernel / log-kernel itself (transform 0)
  const void* Mval;      // (B, n, m) cost for value_linear, or null
  const void* a;         // (B, n)
  const void* b;         // (B, m)
  void* pa;              // (B, n) potentials f / u : output (and state between launches)
  void* pb;              // (B, m) potentials g / v : state in (load_state) and out
  void* T_out;           // (B, n, m) or null
  void* logT_out;        // (B, n, m) or null (log domain)
  double* value_linear;  // (B) or null
  double* value_reg;     // (B) or null : sum T log(T + 1e-16) (entropy) or sum T log(T / (a b) + 1e-16) (kl)
  double* err_b;         // (B) scratch for the stop test
  BatchState* st;
  long long B;
  int n, m, rpc;         // rpc = rows per CTA
  int transform;         // 0 none, 1 exp(-x/reg), 2 -x/reg
  double reg;
  int it_begin, it_end;  // iterations [it_begin, it_end) run in this launch
  int load_state;        // 1: g / v read from pb ; 0: initial value (1 / 0)
  int check;             // 1: stop test after the last iteration of this launch
  int epilogue;          // 1: write plan / values
  int reg_type;          // 0 none, 1 entropy, 2 kl
  int expect_running;    // 1: no-op unless status == ST_RUNNING
  int force_exact;       // test hook: always take the exact column pass (log domain)
  double tol;
};

__device__ __forceinline__ double warp_sum_d(double x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}

// rows a warp handles together (independent shuffle chains hide the butterfly latency)
__host__ __device__ constexpr int batch_rows_per_group(int cpl) { return cpl <= 8 ? 4 : (cpl == 16 ? 2 : 1); }

struct BatchSmem {
  size_t off_part, off_vtmp, off_vec, off_xch, total;
};
__host__ __device__ inline BatchSmem batch_smem_layout(int rpc, int m, int mp, int cs, int esz, int mode) {
  const int nw = batch_threads(mp / 32) / 32;
  BatchSmem L;
  size_t o = ((size_t)rpc * m * esz + 15) / 16 * 16;
  L.off_part = o;
  o += (size_t)nw * mp * esz;
  L.off_vtmp = o;
  o += (mode == 1) ? (size_t)mp * esz : 0;
  L.off_vec = o;
  o += (size_t)(3 * mp + 3 * rpc) * esz;
  o = (o + 15) / 16 * 16;
  L.off_xch = o;
  o += (cs > 1) ? (size_t)2 * cs * mp * esz : 0;
  L.total = o;
  return L;
}

// NTRIP > 0: register-resident fast path of the iteration loop.  Requires m == 32 CPL and exactly
// NTRIP * warps * RG rows in every CTA (e.g. 256 x 256 fp32 on a cluster of 2: 128 rows = 2 trips of 16 warps x 4
// rows): every thread keeps its NTRIP x RG x CPL (<= 64) tile values in registers for the whole solve, so an
// iteration re-reads nothing from shared memory but the vectors, and the loop carries no bounds checks
// (the generic loop spends 12 thread-instructions per element, ncu profiles/r01c_ncu_batch_256_v1.txt).
template <typename T, int MODE, int CPL, int NTRIP = 0>
__global__ void __launch_bounds__(batch_threads(CPL), 1) batch_solve_kernel(const BatchParams p) {
  constexpr int MP = CPL * 32;  // padded column count
  constexpr bool FAST = NTRIP > 0;
  constexpr int BATCH_NT = batch_threads(CPL);
  constexpr int BATCH_NW = BATCH_NT / 32;
  constexpr int RG = batch_rows_per_group(CPL * (int)(sizeof(T) / 4));
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t load_bar;
  __shared__ double red[BATCH_NWMAX][2];
  __shared__ __align__(16) double cl_red[BATCH_CSMAX][2];
  __shared__ int s_flag;
  __shared__ __align__(8) uint64_t xbar[2];  // per-parity mbarriers of the per-iteration column exchange

  if (p.expect_running && p.st->status != ST_RUNNING) return;  // sticky stop, uniform over the grid

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t crank = cluster_ctarank();
  const uint32_t csize = cluster_nctarank();
  const long long prob = cluster_id_x();
  const int n = p.n, m = p.m;
  const int row0 = (int)crank * p.rpc;
  const int rows = (row0 >= n) ? 0 : ((n - row0) < p.rpc ? (n - row0) : p.rpc);

  // ---- shared-memory carve-up (mirrored by batch_smem_bytes on the host) ----
  const BatchSmem L = batch_smem_layout(p.rpc, m, MP, (int)csize, (int)sizeof(T), MODE);
  T* Ks = reinterpret_cast<T*>(smem_raw);                 // rpc x m
  T* part = reinterpret_cast<T*>(smem_raw + L.off_part);  // [warp][MP] column partials
  T* vtmp = reinterpret_cast<T*>(smem_raw + L.off_vtmp);  // [MP] candidate v (log domain)
  T* gs = reinterpret_cast<T*>(smem_raw + L.off_vec);     // g / v
  T* bs = gs + MP;                                        // b
  T* lbs = bs + MP;                                       // log b
  T* fs = lbs + MP;                                       // f / u
  T* as = fs + p.rpc;                                     // a
  T* las = as + p.rpc;                                    // log a
  T* xch = reinterpret_cast<T*>(smem_raw + L.off_xch);    // [2 parities][CS][MP] cluster exchange
  const int CSX = (int)csize;

  const T* Xg = reinterpret_cast<const T*>(p.X) + ((size_t)prob * n + row0) * m;
  const size_t tile_elems = (size_t)rows * m;
  const size_t tile_bytes = tile_elems * sizeof(T);

  // ---- load this CTA's rows: one TMA bulk stream when alignment allows, else cooperative loads ----
  const bool bulk_ok = (tile_bytes % 16 == 0) && ((reinterpret_cast<uintptr_t>(Xg) & 15) == 0) && tile_bytes > 0;
  if (bulk_ok) {
    if (tid == 0) {
      mbar_init(&load_bar, 1);
      fence_mbar_init();
      mbar_arrive_expect_tx(&load_bar, (uint32_t)tile_bytes);
      const size_t CH = 32768;
      for (size_t o = 0; o < tile_bytes; o += CH) {
        const uint32_t nb = (uint32_t)((tile_bytes - o) < CH ? (tile_bytes - o) : CH);
        bulk_g2s(reinterpret_cast<unsigned char*>(Ks) + o, reinterpret_cast<const unsigned char*>(Xg) + o, nb, &load_bar);
      }
    }
  } else {
    for (size_t i = tid; i < tile_elems; i += BATCH_NT) Ks[i] = Xg[i];
  }
  // vectors (overlap with the bulk copy)
  {
    const T* ag = reinterpret_cast<const T*>(p.a) + (size_t)prob * n + row0;
    const T* bg = reinterpret_cast<const T*>(p.b) + (size_t)prob * m;
    const T* pbg = reinterpret_cast<const T*>(p.pb) + (size_t)prob * m;
    for (int r = tid; r < rows; r += BATCH_NT) {
      const T av = ag[r];
      as[r] = av;
      las[r] = log_t(av);
      fs[r] = (MODE == MODE_LOG) ? (T)0 : (T)1;
    }
    for (int j = tid; j < MP; j += BATCH_NT) {
      const T bv = (j < m) ? bg[j] : (T)1;
      bs[j] = bv;
      lbs[j] = log_t(bv);
      T g0 = (MODE == MODE_LOG) ? (T)0 : (T)1;
      if (p.load_state && j < m) g0 = pbg[j];
      gs[j] = g0;
    }
    if (p.load_state) {
      const T* pag = reinterpret_cast<const T*>(p.pa) + (size_t)prob * n + row0;
      for (int r = tid; r < rows; r += BATCH_NT) fs[r] = pag[r];
    }
  }
  __syncthreads();
  if (bulk_ok) mbar_wait(&load_bar, 0);
  if (p.transform != 0) {
    const T reg = (T)p.reg;
    for (size_t i = tid; i < tile_elems; i += BATCH_NT) {
      const T x = Ks[i];
      Ks[i] = (p.transform == 1) ? exp_t((-x) / reg) : (-x) / reg;
    }
  }
  __syncthreads();
  if (csize > 1) {
    if (tid == 0) {
      mbar_init(&xbar[0], 1);
      mbar_init(&xbar[1], 1);
      fence_mbar_init();
    }
    __syncthreads();
    cluster_sync_all();  // peers' shared memory and barriers are live before the first DSMEM store
  }
  uint32_t xph[2] = {0u, 0u};  // phase parity of xbar[0], xbar[1]

  int par = 0;

  // Cluster-wide (or CTA-wide) reduction of two doubles to every thread of rank 0 / of the CTA.
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
    for (int w = 0; w < BATCH_NW; ++w) {
      t0 += red[w][0];
      t1 += red[w][1];
    }
    if (csize > 1) {
      if (tid == 0) {
        const uint32_t d0 = mapa_u32(smem_u32(&cl_red[crank][0]), 0);
        st_cluster(d0, t0);
        st_cluster(d0 + 8, t1);
      }
      cluster_sync_all();
      t0 = t1 = 0.0;
      if (crank == 0) {
        for (uint32_t q = 0; q < csize; ++q) {
          t0 += cl_red[q][0];
          t1 += cl_red[q][1];
        }
      }
      cluster_sync_all();  // rank 0 has consumed cl_red before anyone may overwrite it
    }
    o0 = t0;
    o1 = t1;
  };

  // Column totals of `part` (and optionally `part2`) across warps and cluster ranks, fixed summation order.
  // Returns through `fn(j, total)` for the columns this thread owns.
  auto column_totals = [&](auto&& fn) {
    if (csize == 1) {
      for (int j = tid; j < m; j += BATCH_NT) {
        T tot = (T)0;
#pragma unroll 4
        for (int w = 0; w < BATCH_NW; ++w) tot += part[w * MP + j];
        fn(j, tot);
      }
    } else {
      // One-sided exchange: every CTA puts its column totals into slot `crank` of every CTA's buffer with st.async,
      // the byte count completes the receiver's mbarrier -- no cluster barrier, no membar per iteration.  Buffer `par`
      // is rewritten two exchanges later; a peer can only send that far ahead after it has received this CTA's next
      // exchange, which every thread here issues after the __syncthreads that follows the reads below.
      const uint32_t mine_addr = smem_u32(xch + (size_t)(par * CSX + crank) * MP);
      const uint32_t bar_addr = smem_u32(&xbar[par]);
      if (tid == 0) mbar_arrive_expect_tx(&xbar[par], (uint32_t)(csize * (uint32_t)m * sizeof(T)));
      for (int j = tid; j < m; j += BATCH_NT) {
        T tot = (T)0;
#pragma unroll 4
        for (int w = 0; w < BATCH_NW; ++w) tot += part[w * MP + j];
        for (uint32_t q = 0; q < csize; ++q)
          st_async_1(mapa_u32(mine_addr + (uint32_t)(j * sizeof(T)), q), tot, mapa_u32(bar_addr, q));
      }
      mbar_wait(&xbar[par], xph[par]);
      xph[par] ^= 1u;
      for (int j = tid; j < m; j += BATCH_NT) {
        T tot = (T)0;
        for (uint32_t q = 0; q < csize; ++q) tot += xch[(size_t)(par * CSX + q) * MP + j];
        fn(j, tot);
      }
      par ^= 1;
    }
  };

  // Exact column log-sum-exp (online max) of K_ij + u_i: v_j = log b_j - LSE_i(K_ij + u_i).
  // Cluster: two exchanges (the column maxima, then the sums rescaled to the common maximum).
  auto exact_columns = [&]() {
    constexpr int NJ = (MP + BATCH_NT - 1) / BATCH_NT;
    T mxl[NJ], sml[NJ];
#pragma unroll
    for (int q = 0; q < NJ; ++q) {
      const int j = tid + q * BATCH_NT;
      T mx = neg_inf<T>(), sm = (T)0;
      if (j < m) {
        for (int r = 0; r < rows; ++r) {
          const T x = Ks[(size_t)r * m + j] + fs[r];
          if (x == neg_inf<T>()) continue;
          if (x > mx) {
            sm = sm * exp_t(mx - x) + (T)1;
            mx = x;
          } else {
            sm += exp_t(x - mx);
          }
        }
        if (csize == 1) gs[j] = lbs[j] - (mx + log_t(sm));
      }
      mxl[q] = mx;
      sml[q] = sm;
    }
    if (csize > 1) {
      T Mg[NJ];
      {
        const uint32_t mine_addr = smem_u32(xch + (size_t)(par * CSX + crank) * MP);
#pragma unroll
        for (int q = 0; q < NJ; ++q) {
          const int j = tid + q * BATCH_NT;
          if (j < m)
            for (uint32_t c = 0; c < csize; ++c) st_cluster(mapa_u32(mine_addr + (uint32_t)(j * sizeof(T)), c), mxl[q]);
        }
        cluster_sync_all();
#pragma unroll
        for (int q = 0; q < NJ; ++q) {
          const int j = tid + q * BATCH_NT;
          T M = neg_inf<T>();
          if (j < m)
            for (uint32_t c = 0; c < csize; ++c) M = max_t(M, xch[(size_t)(par * CSX + c) * MP + j]);
          Mg[q] = M;
        }
        par ^= 1;
      }
      {
        const uint32_t mine_addr = smem_u32(xch + (size_t)(par * CSX + crank) * MP);
#pragma unroll
        for (int q = 0; q < NJ; ++q) {
          const int j = tid + q * BATCH_NT;
          if (j < m) {
            const T sc = (mxl[q] == neg_inf<T>()) ? (T)0 : sml[q] * exp_t(mxl[q] - Mg[q]);
            for (uint32_t c = 0; c < csize; ++c) st_cluster(mapa_u32(mine_addr + (uint32_t)(j * sizeof(T)), c), sc);
          }
        }
        cluster_sync_all();
#pragma unroll
        for (int q = 0; q < NJ; ++q) {
          const int j = tid + q * BATCH_NT;
          if (j < m) {
            T S = (T)0;
            for (uint32_t c = 0; c < csize; ++c) S += xch[(size_t)(par * CSX + c) * MP + j];
            gs[j] = lbs[j] - (Mg[q] + log_t(S));
          }
        }
        par ^= 1;
      }
    }
    __syncthreads();
  };

  // ------------------------------------------------------------------------------------------
  // iterations
  // ------------------------------------------------------------------------------------------
  constexpr int NTR = FAST ? NTRIP : 1;
  T kr[NTR][RG][CPL];  // FAST: this thread's share of the tile, resident for the whole solve
  if constexpr (FAST) {
#pragma unroll
    for (int t = 0; t < NTRIP; ++t) {
#pragma unroll
      for (int g = 0; g < RG; ++g) {
        const int r = warp * RG + t * BATCH_NW * RG + g;
#pragma unroll
        for (int c = 0; c < CPL; ++c) kr[t][g][c] = Ks[(size_t)r * MP + lane + 32 * c];
      }
    }
  }
  for (int it = p.it_begin; it < p.it_end; ++it) {
    T greg[CPL], acc[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) {
      greg[c] = gs[lane + 32 * c];
      acc[c] = (T)0;
    }
    const int n_trips = FAST ? NTRIP : (rows - warp * RG + BATCH_NW * RG - 1) / (BATCH_NW * RG);
#pragma unroll(FAST ? NTRIP : 1)
    for (int trip = 0; trip < n_trips; ++trip) {
      const int rb = warp * RG + trip * BATCH_NW * RG;
      T k[RG][CPL];
#pragma unroll
      for (int g = 0; g < RG; ++g) {
        const int r = rb + g;
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
          const int col = lane + 32 * c;
          if constexpr (FAST)
            k[g][c] = kr[trip < NTR ? trip : 0][g][c];
          else
            k[g][c] = (r < rows && col < m) ? Ks[(size_t)r * m + col] : (MODE == MODE_LOG ? neg_inf<T>() : (T)0);
        }
      }
      if (MODE == MODE_KNOPP) {
        T s[RG];
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          T s0 = (T)0;
#pragma unroll
          for (int c = 0; c < CPL; ++c) s0 = fma(k[g][c], greg[c], s0);
          s[g] = s0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int g = 0; g < RG; ++g) s[g] += __shfl_xor_sync(0xffffffffu, s[g], o);
        }
        if constexpr (FAST) {
          // lane l finishes row (l mod RG) only (one division per lane instead of RG), then the RG scalings are
          // broadcast back with one shuffle each
          const int gsel = lane & (RG - 1);
          T ssel = s[0];
#pragma unroll
          for (int g = 1; g < RG; ++g) ssel = (gsel == g) ? s[g] : ssel;
          const T fsel = as[rb + gsel] / ssel;
          if (lane < RG) fs[rb + lane] = fsel;
#pragma unroll
          for (int g = 0; g < RG; ++g) {
            const T f = __shfl_sync(0xffffffffu, fsel, g);
#pragma unroll
            for (int c = 0; c < CPL; ++c) acc[c] = fma(k[g][c], f, acc[c]);
          }
        } else {
#pragma unroll
          for (int g = 0; g < RG; ++g) {
            const int r = rb + g;
            if (r < rows) {
              const T f = as[r] / s[g];
              if (lane == 0) fs[r] = f;
#pragma unroll
              for (int c = 0; c < CPL; ++c) acc[c] = fma(k[g][c], f, acc[c]);
            }
          }
        }
      } else {
        T mx[RG], s[RG];
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          T m0 = neg_inf<T>();
#pragma unroll
          for (int c = 0; c < CPL; ++c) {
            k[g][c] += greg[c];
            m0 = max_t(m0, k[g][c]);
          }
          mx[g] = m0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int g = 0; g < RG; ++g) mx[g] = max_t(mx[g], __shfl_xor_sync(0xffffffffu, mx[g], o));
        }
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          const T ref = (mx[g] == neg_inf<T>()) ? (T)0 : mx[g];
          T s0 = (T)0;
          if constexpr (FAST && sizeof(T) == 4) {
            // e = 2^(x log2e - rnd(ref log2e)): one FFMA + MUFU.EX2 per element; the rounding of ref log2e is a
            // common factor of the row, divided back out of s and of the row factor below (corr)
            const float nref = -(float)ref * 1.4426950408889634f;
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
              k[g][c] = ex2_ftz(fmaf((float)k[g][c], 1.4426950408889634f, nref));
              s0 += k[g][c];
            }
          } else {
#pragma unroll
            for (int c = 0; c < CPL; ++c) {
              k[g][c] = exp_elem(k[g][c] - ref);
              s0 += k[g][c];
            }
          }
          s[g] = s0;
          mx[g] = ref;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int g = 0; g < RG; ++g) s[g] += __shfl_xor_sync(0xffffffffu, s[g], o);
        }
        if constexpr (FAST) {
          const int gsel = lane & (RG - 1);
          T ssel = s[0], msel = mx[0];
#pragma unroll
          for (int g = 1; g < RG; ++g) {
            ssel = (gsel == g) ? s[g] : ssel;
            msel = (gsel == g) ? mx[g] : msel;
          }
          T corr = (T)1;
          if constexpr (sizeof(T) == 4) {
            const float nref = -(float)msel * 1.4426950408889634f;
            corr = (T)(1.f - fmaf((float)msel, 1.4426950408889634f, nref) * 0.6931471805599453f);
          }
          const T la = las[rb + gsel];
          const T ls = log_t(ssel * corr);
          if (lane < RG) fs[rb + lane] = la - (msel + ls);    // u_i = log a_i - LSE_j(K_ij + v_j)
          const T wsel = exp_t(la - ls) * corr;               // a_i / s_i in this row's exponent frame
#pragma unroll
          for (int g = 0; g < RG; ++g) {
            const T w = __shfl_sync(0xffffffffu, wsel, g);
#pragma unroll
            for (int c = 0; c < CPL; ++c) acc[c] = fma(k[g][c], w, acc[c]);
          }
        } else {
#pragma unroll
          for (int g = 0; g < RG; ++g) {
            const int r = rb + g;
            if (r < rows) {
              const T ls = log_t(s[g]);
              if (lane == 0) fs[r] = las[r] - (mx[g] + ls);   // u_i = log a_i - LSE_j(K_ij + v_j)
              const T w = exp_t(las[r] - ls);                 // a_i / s_i
#pragma unroll
              for (int c = 0; c < CPL; ++c) acc[c] = fma(k[g][c], w, acc[c]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < CPL; ++c) part[warp * MP + lane + 32 * c] = acc[c];
    __syncthreads();

    if (MODE == MODE_KNOPP) {
      column_totals([&](int j, T tot) { gs[j] = bs[j] / tot; });
      __syncthreads();
    } else {
      int bad = p.force_exact;
      column_totals([&](int j, T tot) {
        // tot = exp(v_j) * sum_i exp(K_ij + u_i): the column marginal of the current plan
        if (!(tot >= tiny_guard<T>())) bad = 1;
        vtmp[j] = gs[j] + (lbs[j] - log_t(tot));
      });
      bad = __syncthreads_or(bad);
      if (!bad) {
        for (int j = tid; j < m; j += BATCH_NT) gs[j] = vtmp[j];
        __syncthreads();
      } else {
        exact_columns();  // every CTA of the cluster saw the same totals, so the branch is cluster-uniform
      }
    }
  }

  // ------------------------------------------------------------------------------------------
  // stop test: L2 violation of the row marginal (with the g / v just produced), max over the batch
  // ------------------------------------------------------------------------------------------
  if (p.check) {
    double e2 = 0.0;
    T greg[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) greg[c] = gs[lane + 32 * c];
    for (int r = warp; r < rows; r += BATCH_NW) {
      T s0 = (T)0;
      const T fi = fs[r];
#pragma unroll
      for (int c = 0; c < CPL; ++c) {
        const int col = lane + 32 * c;
        if (col < m) {
          const T kij = Ks[(size_t)r * m + col];
          if (MODE == MODE_KNOPP)
            s0 += (kij * fi) * greg[c];
          else
            s0 += exp_t((kij + fi) + greg[c]);
        }
      }
      s0 = warp_sum(s0);
      const T d = s0 - as[r];
      if (lane == 0) e2 += (double)(d * d);
    }
    double t0, t1;
    reduce2(e2, 0.0, t0, t1);
    if (crank == 0 && tid == 0) {
      p.err_b[prob] = (double)(T)sqrt((double)(T)t0);
      __threadfence();
      const unsigned int t = atomicAdd(&p.st->ticket, 1u);
      s_flag = (t == (unsigned int)(p.B - 1)) ? 1 : 0;
    }
    __syncthreads();
    if (crank == 0 && s_flag) {  // last problem of the batch: reference's err = max over problems
      __threadfence();
      double mxe = 0.0;
      bool nan = false;
      for (long long i = tid; i < p.B; i += BATCH_NT) {
        const double e = __ldcg(&p.err_b[i]);
        if (e != e) nan = true;
        mxe = fmax(mxe, e);
      }
      if (nan) mxe = CUDART_NAN;
      // block max (nan propagates like np.max)
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
        for (int w = 0; w < BATCH_NW; ++w) {
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
    T* pag = reinterpret_cast<T*>(p.pa) + (size_t)prob * n + row0;
    for (int r = tid; r < rows; r += BATCH_NT) pag[r] = fs[r];
    if (crank == 0) {
      T* pbg = reinterpret_cast<T*>(p.pb) + (size_t)prob * m;
      for (int j = tid; j < m; j += BATCH_NT) pbg[j] = gs[j];
    }
  }

  // ------------------------------------------------------------------------------------------
  // epilogue: plan, <M, T>, regularisation term  (ot/batch/_linear.py:338-345, _utils.py:4-22)
  // ------------------------------------------------------------------------------------------
  if (p.epilogue) {
    double vl = 0.0, vr = 0.0;
    T* Tg = p.T_out ? reinterpret_cast<T*>(p.T_out) + ((size_t)prob * n + row0) * m : nullptr;
    T* Lg = p.logT_out ? reinterpret_cast<T*>(p.logT_out) + ((size_t)prob * n + row0) * m : nullptr;
    const T* Mg = p.Mval ? reinterpret_cast<const T*>(p.Mval) + ((size_t)prob * n + row0) * m : nullptr;
    T greg[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) greg[c] = gs[lane + 32 * c];
    for (int r = warp; r < rows; r += BATCH_NW) {
      const T fi = fs[r];
      const T ai = as[r];
#pragma unroll
      for (int c = 0; c < CPL; ++c) {
        const int col = lane + 32 * c;
        if (col < m) {
          const size_t idx = (size_t)r * m + col;
          const T kij = Ks[idx];
          T t;
          if (MODE == MODE_KNOPP) {
            t = (kij * fi) * greg[c];
          } else {
            const T lt = (kij + fi) + greg[c];
            if (Lg) Lg[idx] = lt;
            t = exp_t(lt);
          }
          if (Tg) Tg[idx] = t;
          if (Mg) vl += (double)(Mg[idx] * t);
          if (p.reg_type == 1)
            vr += (double)(t * log_t(t + (T)1e-16));
          else if (p.reg_type == 2)
            vr += (double)(t * log_t(t / (ai * bs[col]) + (T)1e-16));
        }
      }
    }
    double t0, t1;
    reduce2(vl, vr, t0, t1);
    if (crank == 0 && tid == 0) {
      if (p.value_linear) p.value_linear[prob] = t0;
      if (p.value_reg) p.value_reg[prob] = t1;
    }
  }
  if (csize > 1) cluster_sync_all();  // nobody exits while a peer may still address its shared memory
}

}  // namespace b2ot
