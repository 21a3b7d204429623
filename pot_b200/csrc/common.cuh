// Common device helpers for the b2ot kernels (sm_100a only).
//
// PTX wrappers for the Blackwell/Hopper async machinery used by the sweep
// kernels: mbarrier (local + remote arrive), 1-D bulk async copies
// (cp.async.bulk -> SASS UBLKCP, the TMA engine's linear mode), thread-block
// cluster identity, DSMEM stores (mapa + st.shared::cluster) and warp reductions.
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace b2ot {

// ----------------------------------------------------------------------------------------------
// Solver state living in device memory: every kernel of a solve reads it first and turns into a
// no-op once `status` leaves ST_RUNNING (sticky), so the host never has to sync per iteration.
// ----------------------------------------------------------------------------------------------
enum : int {
  ST_RUNNING = 0,      // iterations are being applied
  ST_CONVERGED = 1,    // err < stopThr at iteration `niter`
  ST_NUMERR = 2,       // numerical guard fired at iteration `niter`; (u, v) reverted
  ST_MAXITER = 3,      // loop exhausted
  ST_PAUSE_EXACT = 4   // log domain: a column sum underflowed; host must run the exact column pass
};

struct __align__(16) SolveState {
  int status;
  int iter;        // index ii of the iteration the next sweep belongs to
  int niter;       // reference's log["niter"] (last loop index)
  int sel_u;       // slot of the current u in the 2-ring
  int sel_v;       // slot of the current v in the 3-ring
  int cur_flags;   // guard flags attached to the current (c, v): bit0 c==0, bit1 v nan/inf
  int next_flags;  // accumulated by finalize blocks for the v being produced
  int uflag;       // accumulated by the sweep: nan/inf in the u being produced
  int need_exact;  // accumulated by finalize (log domain): column sum below the underflow guard
  unsigned int ticket;
  int n_err;       // entries written to err_hist
  int absorb;      // accumulated by finalize blocks (stabilized): some |u_i| or |v_j| above tau
  double last_err;
  // stabilized scaling iteration (b2ot_dense_set_stabilized)
  int absorb_now;  // set by the finalize of an iteration whose max|u| / max|v| exceeded tau; cleared by absorb_vec_kernel
  int rebuild;     // set by absorb_vec_kernel when the solve goes on: K must be rebuilt and the column pass re-seeded
  int n_absorb;    // absorptions so far
  int pad1;
};

constexpr int MODE_KNOPP = 0;  // scaling domain
constexpr int MODE_LOG = 1;    // log domain
constexpr int MODE_UNBAL = 2;  // fused sweep only: scaling domain with the unbalanced row exponent u = (a / Kv)^fi

constexpr int FLAG_CZERO = 1;
constexpr int FLAG_VBAD = 2;

template <typename T> struct Vec;
template <> struct Vec<float> { using type = float4; static constexpr int N = 4; };
template <> struct Vec<double> { using type = double2; static constexpr int N = 2; };

template <typename T> __device__ __forceinline__ T neg_inf();
template <> __device__ __forceinline__ float neg_inf<float>() { return -CUDART_INF_F; }
template <> __device__ __forceinline__ double neg_inf<double>() { return -CUDART_INF; }

// Underflow guard for the row-frame column sums of the log-domain sweep.
template <typename T> __device__ __forceinline__ T tiny_guard();
template <> __device__ __forceinline__ float tiny_guard<float>() { return 1e-30f; }
template <> __device__ __forceinline__ double tiny_guard<double>() { return 1e-290; }

__device__ __forceinline__ float exp_t(float x) { return expf(x); }
__device__ __forceinline__ double exp_t(double x) { return exp(x); }
// Per-element exponential of the log-domain row pass, argument d = x - rowmax <= 0.  fp32: one FMUL
// + MUFU.EX2 (ex2.approx); its error grows with |d|, i.e. it is at the 1e-7 level for the entries
// that carry the sum and only reaches ~1e-5 relative on entries that are e^-100 of it.
__device__ __forceinline__ float exp_elem(float d) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(d * 1.4426950408889634f));
  return y;
}
__device__ __forceinline__ double exp_elem(double d) { return exp(d); }
__device__ __forceinline__ float log_t(float x) { return logf(x); }
__device__ __forceinline__ double log_t(double x) { return log(x); }
__device__ __forceinline__ float pow_t(float x, float y) { return powf(x, y); }
__device__ __forceinline__ double pow_t(double x, double y) { return pow(x, y); }
__device__ __forceinline__ float abs_t(float x) { return fabsf(x); }
__device__ __forceinline__ double abs_t(double x) { return fabs(x); }
__device__ __forceinline__ float max_t(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ double max_t(double a, double b) { return fmax(a, b); }
__device__ __forceinline__ bool finite_t(float x) { return isfinite(x); }
__device__ __forceinline__ bool finite_t(double x) { return isfinite(x); }

// ---- packed fp32x2 arithmetic (sm_100: FADD2 / FFMA2, one issue slot for two lanes of work) ----
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// (x0, x1) += (y0, y1)
__device__ __forceinline__ void add2(float& x0, float& x1, float y0, float y1) {
  unsigned long long r;
  asm("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pack2(x0, x1)), "l"(pack2(y0, y1)));
  unpack2(r, x0, x1);
}
// (d0, d1) = (a0, a1) * (b0, b1) + (c0, c1)
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pack2(a0, a1)), "l"(pack2(b0, b1)), "l"(pack2(c0, c1)));
  unpack2(r, d0, d1);
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
// warp-wide maximum in one instruction (sm_100a: CREDUX.MAX.F32); every lane gets the result
__device__ __forceinline__ float warp_max_redux(float x) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float warp_min_redux(float x) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float lg2_ftz(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Row exponent x^y of the unbalanced sweep (x = a / Kv > 0, 0 < y <= 1).  It sits on the serial path of every row
// group of the fused sweep, where the accurate powf (~60 dependent instructions) costs 20 % of the sweep.  fp32: split
// x = 2^e m, m in [1, 2): y log2 x = y e (two-float product, integer part peeled off exactly) + y lg2.approx(m), one
// ex2.approx on an argument of magnitude <= 1.5, exact 2^ip scaling.  Relative error ~2e-7 typical, <= ~5e-7
// (lg2.approx 2^-22 absolute on [1, 2), ex2.approx 2 ulp) -- the documented bound of powf itself is 4 ulp.
// Zero, denormal, inf, nan, negative inputs take powf so that the numerical guard sees the same special values.
__device__ __forceinline__ float pow_row(float x, float y) {
  if (!(x >= 1.17549435e-38f && x <= 3.4028234e38f)) return powf(x, y);
  const int xi = __float_as_int(x);
  const float ef = (float)(((xi >> 23) & 0xff) - 127);
  const float m = __int_as_float((xi & 0x007fffff) | 0x3f800000);
  const float lm = lg2_ftz(m);
  const float t_hi = y * ef;
  const float t_lo = fmaf(y, ef, -t_hi);
  const float ip = rintf(t_hi);
  const float arg = (t_hi - ip) + fmaf(y, lm, t_lo);
  const float r = ex2_ftz(arg);
  return r * __int_as_float(((int)ip + 127) << 23);
}
__device__ __forceinline__ double pow_row(double x, double y) { return pow(x, y); }

// ---- element access on the 16-byte vector types ----
__device__ __forceinline__ float& elem(float4& v, int i) { return (&v.x)[i]; }
__device__ __forceinline__ double& elem(double2& v, int i) { return (&v.x)[i]; }
__device__ __forceinline__ const float& elem(const float4& v, int i) { return (&v.x)[i]; }
__device__ __forceinline__ const double& elem(const double2& v, int i) { return (&v.x)[i]; }
__device__ __forceinline__ void fill(float4& v, float x) { v.x = v.y = v.z = v.w = x; }
__device__ __forceinline__ void fill(double2& v, double x) { v.x = v.y = x; }

// ---- warp reductions (butterfly: every lane ends with the result) ----
template <typename T> __device__ __forceinline__ T warp_sum(T x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
  return x;
}
template <typename T> __device__ __forceinline__ T warp_max(T x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x = max_t(x, __shfl_xor_sync(0xffffffffu, x, o));
  return x;
}

// ---- shared-memory addresses ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ----
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// acquire at cluster scope: pairs with the remote release-arrive of a peer CTA
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait_cluster(bar, parity)) {
  }
}
// arrive on the mbarrier of another CTA of the cluster (address from mapa)
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

// ---- cluster identity / barrier / DSMEM ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_id_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_count_x() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void st_cluster(uint32_t addr, double v) {
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}

// One-sided DSMEM put: stores two values into a peer CTA's shared memory and signals that CTA's
// mbarrier with the byte count (complete_tx) -- no fence, no separate arrive.
__device__ __forceinline__ void st_async_v2(uint32_t dst_cluster_addr, float x, float y, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(
                   dst_cluster_addr),
               "f"(x), "f"(y), "r"(bar_cluster_addr)
               : "memory");
}
__device__ __forceinline__ void st_async_v2(uint32_t dst_cluster_addr, double x, double y, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];" ::"r"(
                   dst_cluster_addr),
               "d"(x), "d"(y), "r"(bar_cluster_addr)
               : "memory");
}

__device__ __forceinline__ void st_async_1(uint32_t dst_cluster_addr, float x, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f32 [%0], %1, [%2];" ::"r"(dst_cluster_addr),
               "f"(x), "r"(bar_cluster_addr)
               : "memory");
}
__device__ __forceinline__ void st_async_1(uint32_t dst_cluster_addr, double x, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(dst_cluster_addr),
               "d"(x), "r"(bar_cluster_addr)
               : "memory");
}

// ---- bulk async copy global -> shared (TMA linear mode; completes on an mbarrier) ----
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void bulk_g2s_hint(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar,
                                              uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// ---- programmatic dependent launch (no-ops when the kernel was launched without the attribute) ----
// wait: the preceding kernel of the stream has completed and its memory is visible
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// the next kernel of the stream may start launching (it still blocks in its own griddep_wait)
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace b2ot
