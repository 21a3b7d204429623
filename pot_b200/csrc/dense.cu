// Host driver + C ABI of the dense Sinkhorn solver (see include/b2ot.h, dense_kernels.cuh).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <tuple>
#include <vector>

#include "../../include/b2ot.h"
#include "comm_host.h"
#include "dense_kernels.cuh"
#include "dense_resident_kernels.cuh"
#include "dense_logpp_kernels.cuh"
#include "host_util.h"

using namespace b2ot;

namespace b2ot {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
std::atomic<unsigned long long>& launch_counter() {
  static std::atomic<unsigned long long> c{0};
  return c;
}

static std::mutex g_pin_mu;
static std::vector<void*> g_pin_free;
void* pinned_acquire(size_t bytes) {
  if (bytes > kPinnedBlock) return nullptr;
  {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    if (!g_pin_free.empty()) {
      void* p = g_pin_free.back();
      g_pin_free.pop_back();
      return p;
    }
  }
  void* p = nullptr;
  if (cudaMallocHost(&p, kPinnedBlock) != cudaSuccess) return nullptr;
  return p;
}
void pinned_release(void* p) {
  if (p == nullptr) return;
  std::lock_guard<std::mutex> lk(g_pin_mu);
  if (g_pin_free.size() < 64)
    g_pin_free.push_back(p);
  else
    cudaFreeHost(p);
}

// cudaOccupancyMaxActiveClusters takes ~1 ms: cache its answer per (kernel, device, cluster size, shared memory)
static std::mutex g_occ_mu;
static std::map<std::tuple<const void*, int, int, size_t>, int> g_occ_cache;
template <typename Kern>
static int cached_max_clusters(Kern kern, const cudaLaunchConfig_t* cfg, int cs, int* out) {
  int dev = 0;
  cudaGetDevice(&dev);
  const auto key = std::make_tuple((const void*)kern, dev, cs, (size_t)cfg->dynamicSmemBytes);
  {
    std::lock_guard<std::mutex> lk(g_occ_mu);
    auto it = g_occ_cache.find(key);
    if (it != g_occ_cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  int maxcl = 0;
  B2OT_CUDA(cudaOccupancyMaxActiveClusters(&maxcl, kern, cfg));
  std::lock_guard<std::mutex> lk(g_occ_mu);
  g_occ_cache[key] = maxcl;
  *out = maxcl;
  return 0;
}

}  // namespace b2ot

struct b2ot_dense {
  int mode, dtype;
  long long n, m, ld, mp;
  const void *K, *a, *b, *logb;
  int max_iter, check_every;
  double stop_thr;
  int variant;  // resolved: FUSED or SPLIT
  cudaStream_t stream;
  // workspace carve-up
  SolveState* st;
  double* blk_err;
  double* err_hist;
  int err_cap;
  void *uring, *vring, *cbuf, *partials;
  // fused sweep configuration
  int cs, ncl, stages, vpt, nw, occ, W, use_hint;
  int pp_vpt;  // > 0: fp32 log-domain sweeps run the ping-pong kernel with this many vectors per thread
  int pp_occ;  // CTAs per SM of the ping-pong kernel (2: narrow slices, four independent halves per SM)
  size_t dyn_smem;
  // split sweep / column-only pass configuration
  int nsplit, nsplit_lse;
  int fin_blocks;
  b2ot_allreduce_fn ar;
  void* ar_ctx;
  b2ot_comm* comm;  // NVLink peer-memory exchange (takes precedence over the hook)
  int comm_ll;      // 1: the column sums travel as tagged words inside the finalize kernel (slots large enough)
  unsigned long long comm_ll_off;  // byte offset of the tagged-word region inside a slot
  SolveState* h_state;  // pinned host mirror for polls
  cudaGraphExec_t gexec;  // graph of kGraphIters iterations (single-device solves), built at the first iterate call
  int graph_state;        // 0 not built, 1 ready, -1 unavailable
  int graph_launches;     // kernels per replay
  int unbal;            // 1: unbalanced scaling iteration u = (a/Kv)^fi1, v = (b/K^T u)^fi2 (b2ot_dense_set_unbalanced)
  double fi1, fi2;
  double tau;           // > 0: stabilized scaling iteration (b2ot_dense_set_stabilized)
  const void* stab_mr;  // Mr = -M/reg (n x m, leading dimension stab_ld), source of the K rebuilds
  long long stab_ld;
  void *stab_alpha, *stab_beta;  // caller-owned alpha/reg (n), beta/reg (m) as DOUBLES, updated by the absorptions
  // resident solver of small problems (dense_resident_kernels.cuh): K lives in the grid's shared memory
  int res_ok;        // shape fits and the variant allows it
  int res_rows;      // rows per CTA
  int res_cols;      // columns per CTA
  int res_rep;       // copies of the exchange area
  int res_grid;      // CTAs (<= SM count: the launch is cooperative)
  size_t res_smem;
};

static void drop_graph(b2ot_dense* h);
static inline size_t esize(int dtype) { return dtype == B2OT_F64 ? 8 : 4; }
static inline long long padded_m(long long m) { return (m + 1 + 31) / 32 * 32; }  // room for the flag slot

extern "C" int b2ot_version(void) { return 100; }
extern "C" const char* b2ot_last_error(void) { return b2ot::g_err; }
extern "C" unsigned long long b2ot_launch_count(void) { return launch_counter().load(); }

extern "C" int b2ot_device_info(int device, int* sm_count, int* cc_major, int* cc_minor) {
  cudaDeviceProp prop;
  B2OT_CUDA(cudaGetDeviceProperties(&prop, device));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  return 0;
}

// ------------------------------------------------------------------------------------------------
// elementwise entry points
// ------------------------------------------------------------------------------------------------
template <typename T, int OP>
static int launch_cost_transform(const void* M, void* out, long long n, long long m, long long ldM, long long ldO,
                                 double reg, cudaStream_t s) {
  if (n <= 0 || m <= 0) return 0;
  constexpr int VEC = Vec<T>::N;
  if (m % VEC == 0 && ldM % VEC == 0 && ldO % VEC == 0 && ((uintptr_t)M % 16) == 0 && ((uintptr_t)out % 16) == 0) {
    const long long nv = m / VEC;
    const unsigned gx = (unsigned)std::min<long long>((nv + 255) / 256, 32);
    const unsigned gy = (unsigned)std::max<long long>(1, std::min<long long>((n + 1) / 2, (148 * 16) / gx));
    cost_transform_vec_kernel<T, OP><<<dim3(gx, gy), 256, 0, s>>>((const T*)M, (T*)out, n, m, ldM, ldO, (T)reg);
    B2OT_CHECK_LAUNCH();
    return 0;
  }
  dim3 grid((unsigned)std::min<long long>(n, 1 << 20), (unsigned)((m + 1023) / 1024));
  cost_transform_kernel<T, OP><<<grid, 256, 0, s>>>((const T*)M, (T*)out, n, m, ldM, ldO, (T)reg);
  B2OT_CHECK_LAUNCH();
  return 0;
}
extern "C" int b2ot_kexp(const void* M, void* K, int64_t n, int64_t m, int64_t ldM, int64_t ldK, double reg, int dtype,
                         void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_kexp: bad dtype %d", dtype);
  return dtype == B2OT_F32 ? launch_cost_transform<float, 0>(M, K, n, m, ldM, ldK, reg, (cudaStream_t)stream)
                           : launch_cost_transform<double, 0>(M, K, n, m, ldM, ldK, reg, (cudaStream_t)stream);
}
extern "C" int b2ot_negscale(const void* M, void* Mr, int64_t n, int64_t m, int64_t ldM, int64_t ldMr, double reg,
                             int dtype, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_negscale: bad dtype %d", dtype);
  return dtype == B2OT_F32 ? launch_cost_transform<float, 1>(M, Mr, n, m, ldM, ldMr, reg, (cudaStream_t)stream)
                           : launch_cost_transform<double, 1>(M, Mr, n, m, ldM, ldMr, reg, (cudaStream_t)stream);
}

template <typename T>
static int launch_plan(int mode, const void* K, const void* u, const void* v, void* out, long long n, long long m,
                       long long ldK, long long ldO, cudaStream_t s) {
  if (n <= 0 || m <= 0) return 0;
  constexpr int VEC = Vec<T>::N;
  if (m % VEC == 0 && ldK % VEC == 0 && ldO % VEC == 0 && ((uintptr_t)K % 16) == 0 && ((uintptr_t)out % 16) == 0 &&
      ((uintptr_t)v % 16) == 0) {
    const long long nv = m / VEC;
    const unsigned gx = (unsigned)std::min<long long>((nv + 255) / 256, 32);
    const unsigned gy = (unsigned)std::max<long long>(1, std::min<long long>((n + 1) / 2, (148 * 16) / gx));
    if (mode == B2OT_MODE_KNOPP)
      plan_vec_kernel<T, MODE_KNOPP><<<dim3(gx, gy), 256, 0, s>>>((const T*)K, (const T*)u, (const T*)v, (T*)out, n, m,
                                                                 ldK, ldO);
    else
      plan_vec_kernel<T, MODE_LOG><<<dim3(gx, gy), 256, 0, s>>>((const T*)K, (const T*)u, (const T*)v, (T*)out, n, m,
                                                               ldK, ldO);
    B2OT_CHECK_LAUNCH();
    return 0;
  }
  dim3 grid((unsigned)std::min<long long>(n, 1 << 20), (unsigned)((m + 1023) / 1024));
  if (mode == B2OT_MODE_KNOPP)
    plan_kernel<T, MODE_KNOPP><<<grid, 256, 0, s>>>((const T*)K, (const T*)u, (const T*)v, (T*)out, n, m, ldK, ldO);
  else
    plan_kernel<T, MODE_LOG><<<grid, 256, 0, s>>>((const T*)K, (const T*)u, (const T*)v, (T*)out, n, m, ldK, ldO);
  B2OT_CHECK_LAUNCH();
  return 0;
}
extern "C" int b2ot_dense_plan(int mode, int dtype, const void* K, const void* u, const void* v, void* out, int64_t n,
                               int64_t m, int64_t ldK, int64_t ldOut, void* stream) {
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || mode == B2OT_MODE_LOG, "b2ot_dense_plan: bad mode %d", mode);
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_dense_plan: bad dtype %d", dtype);
  return dtype == B2OT_F32 ? launch_plan<float>(mode, K, u, v, out, n, m, ldK, ldOut, (cudaStream_t)stream)
                           : launch_plan<double>(mode, K, u, v, out, n, m, ldK, ldOut, (cudaStream_t)stream);
}

static const int kLossBlocks = 1184;  // 148 SMs x 8
extern "C" size_t b2ot_dense_loss_scratch_bytes(int64_t, int64_t) { return kLossBlocks * sizeof(double); }
template <typename T>
static int launch_loss(int mode, const void* K, const void* M, const void* u, const void* v, long long n, long long m,
                       long long ldK, long long ldM, double* loss, void* scratch, cudaStream_t s) {
  const int nb = (int)std::max<long long>(1, std::min<long long>(n, kLossBlocks));
  if (mode == B2OT_MODE_KNOPP)
    loss_kernel<T, MODE_KNOPP><<<nb, 256, 0, s>>>((const T*)K, (const T*)M, (const T*)u, (const T*)v, n, m, ldK, ldM,
                                                  (double*)scratch);
  else
    loss_kernel<T, MODE_LOG><<<nb, 256, 0, s>>>((const T*)K, (const T*)M, (const T*)u, (const T*)v, n, m, ldK, ldM,
                                                (double*)scratch);
  B2OT_CHECK_LAUNCH();
  sum_blocks_kernel<<<1, 256, 0, s>>>((const double*)scratch, nb, loss);
  B2OT_CHECK_LAUNCH();
  return 0;
}
extern "C" int b2ot_dense_loss(int mode, int dtype, const void* K, const void* M, const void* u, const void* v,
                               int64_t n, int64_t m, int64_t ldK, int64_t ldM, double* loss_dev, void* scratch_dev,
                               void* stream) {
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || mode == B2OT_MODE_LOG, "b2ot_dense_loss: bad mode %d", mode);
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_dense_loss: bad dtype %d", dtype);
  return dtype == B2OT_F32
             ? launch_loss<float>(mode, K, M, u, v, n, m, ldK, ldM, loss_dev, scratch_dev, (cudaStream_t)stream)
             : launch_loss<double>(mode, K, M, u, v, n, m, ldK, ldM, loss_dev, scratch_dev, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// dense solver handle
// ------------------------------------------------------------------------------------------------
struct WsLayout {
  size_t off_state, off_blk, off_err, off_u, off_v, off_c, off_p, total;
  int err_cap;
};
static WsLayout ws_layout(long long n, long long m, int dtype, int max_iter, int check_every) {
  WsLayout L;
  const size_t sz = esize(dtype);
  const long long mp = padded_m(m);
  size_t o = 0;
  L.off_state = o;
  o += 256;
  L.off_blk = o;
  o += 1024 * sizeof(double);
  L.err_cap = max_iter / (check_every > 0 ? check_every : 1) + 2;
  L.off_err = o;
  o = align_up(o + (size_t)L.err_cap * sizeof(double), 256);
  L.off_u = o;
  o = align_up(o + 2 * (size_t)n * sz, 256);
  L.off_v = o;
  o = align_up(o + 3 * (size_t)mp * sz, 256);
  L.off_c = o;
  o = align_up(o + (size_t)mp * sz, 256);
  L.off_p = o;
  o = align_up(o + (size_t)NPMAX * mp * sz, 256);
  L.total = o;
  return L;
}
extern "C" size_t b2ot_dense_workspace_bytes(int64_t n, int64_t m, int dtype, int max_iter, int check_every) {
  return ws_layout(n, m, dtype, max_iter, check_every).total;
}

template <typename T, int VPT, int MODE, int NW, int OCC>
static int config_fused(b2ot_dense* h) {
  constexpr int R = rows_per_group(VPT, NW, OCC);
  constexpr int WCAP = NW * 32 * VPT * Vec<T>::N;
  const size_t stage_bytes = (size_t)R * WCAP * sizeof(T);
  int stages = (int)(((OCC == 2 ? 100 : 200) * 1024) / stage_bytes);
  if (stages > SMAX) stages = SMAX;
  if (stages < 2) stages = 2;
  const char* env = diag_env("B2OT_STAGES");
  if (env && atoi(env) >= 2 && atoi(env) <= SMAX && (size_t)atoi(env) * stage_bytes <= (size_t)(OCC == 2 ? 104 : 216) * 1024) stages = atoi(env);
  h->stages = stages;
  h->dyn_smem = stages * stage_bytes;
  auto kern = fused_sweep_kernel<T, VPT, MODE, NW, OCC>;
  B2OT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->dyn_smem));
  if (h->cs > 8) B2OT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)h->cs, 1, 1);
  cfg.blockDim = dim3(NW * 32, 1, 1);
  cfg.dynamicSmemBytes = h->dyn_smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)h->cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int maxcl = 0;
  if (int rc_occ = cached_max_clusters(kern, &cfg, h->cs, &maxcl)) return rc_occ;
  B2OT_REQUIRE(maxcl >= 1, "fused sweep: no cluster of size %d fits on this device", h->cs);
  const long long NG = (h->n + R - 1) / R;
  long long ncl = std::min<long long>(maxcl, std::max<long long>(NG, 1));
  const char* envc = diag_env("B2OT_CLUSTERS");
  if (envc && atoi(envc) >= 1) ncl = std::min<long long>(ncl, atoi(envc));
  if (ncl > NPMAX) ncl = NPMAX;
  h->ncl = (int)ncl;
  return 0;
}

template <typename T, int VPT, int MODE, int NW, int OCC>
static int launch_fused(b2ot_dense* h) {
  SweepParams p;
  p.K = h->K;
  p.ld = h->ld;
  p.n = h->n;
  p.m = h->m;
  p.mp = h->mp;
  p.a = h->a;
  p.uring = h->uring;
  p.vring = h->vring;
  p.partials = h->partials;
  p.st = h->st;
  p.W = h->W;
  p.stages = h->stages;
  p.use_hint = h->use_hint;
  p.fi = h->unbal ? h->fi1 : 1.0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(h->ncl * h->cs), 1, 1);
  cfg.blockDim = dim3(NW * 32, 1, 1);
  cfg.dynamicSmemBytes = h->dyn_smem;
  cfg.stream = h->stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)h->cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B2OT_CUDA(cudaLaunchKernelEx(&cfg, fused_sweep_kernel<T, VPT, MODE, NW, OCC>, p));
  count_launch();
  return 0;
}

// ---- ping-pong log-domain sweep (fp32): configuration and launch ----
template <int VPT, int OCC>
static int config_pp(b2ot_dense* h) {
  int stages = PP_SMAX;
  const size_t budget = OCC == 2 ? (size_t)(110 * 1024) : (size_t)(228 * 1024 - 4096);
  while (stages > 2 && pp_smem_bytes(VPT, stages) > budget) --stages;
  const char* env = diag_env("B2OT_PP_STAGES");
  if (env && atoi(env) >= 2 && atoi(env) <= stages) stages = atoi(env);
  h->stages = stages;
  h->dyn_smem = pp_smem_bytes(VPT, stages);
  auto kern = log_pp_sweep_kernel<VPT, OCC>;
  B2OT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->dyn_smem));
  if (h->cs > 8) B2OT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)h->cs, 1, 1);
  cfg.blockDim = dim3(PP_NH * PP_NTH, 1, 1);
  cfg.dynamicSmemBytes = h->dyn_smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)h->cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int maxcl = 0;
  if (int rc_occ = cached_max_clusters(kern, &cfg, h->cs, &maxcl)) return rc_occ;
  B2OT_REQUIRE(maxcl >= 1, "ping-pong sweep: no cluster of size %d fits on this device", h->cs);
  long long ncl = std::min<long long>(maxcl, std::max<long long>((h->n + PP_NH - 1) / PP_NH, 1));
  const char* envc = diag_env("B2OT_CLUSTERS");
  if (envc && atoi(envc) >= 1) ncl = std::min<long long>(ncl, atoi(envc));
  if (ncl > NPMAX) ncl = NPMAX;
  h->ncl = (int)ncl;
  return 0;
}
template <int VPT, int OCC>
static int launch_pp(b2ot_dense* h) {
  SweepParams p;
  p.K = h->K;
  p.ld = h->ld;
  p.n = h->n;
  p.m = h->m;
  p.mp = h->mp;
  p.a = h->a;
  p.uring = h->uring;
  p.vring = h->vring;
  p.partials = h->partials;
  p.st = h->st;
  p.W = h->W;
  p.stages = h->stages;
  p.use_hint = h->use_hint;
  p.fi = 1.0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(h->ncl * h->cs), 1, 1);
  cfg.blockDim = dim3(PP_NH * PP_NTH, 1, 1);
  cfg.dynamicSmemBytes = h->dyn_smem;
  cfg.stream = h->stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)h->cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B2OT_CUDA(cudaLaunchKernelEx(&cfg, log_pp_sweep_kernel<VPT, OCC>, p));
  count_launch();
  return 0;
}
static int dispatch_config_pp(b2ot_dense* h) {
  switch (h->pp_vpt) {
    case 4: return h->pp_occ == 2 ? config_pp<4, 2>(h) : config_pp<4, 1>(h);
    case 8: return h->pp_occ == 2 ? config_pp<8, 2>(h) : config_pp<8, 1>(h);
    case 16: return config_pp<16, 1>(h);
    default: break;
  }
  set_error("internal: bad ping-pong dispatch (vpt=%d)", h->pp_vpt);
  return -1;
}
static int dispatch_launch_pp(b2ot_dense* h) {
  switch (h->pp_vpt) {
    case 4: return h->pp_occ == 2 ? launch_pp<4, 2>(h) : launch_pp<4, 1>(h);
    case 8: return h->pp_occ == 2 ? launch_pp<8, 2>(h) : launch_pp<8, 1>(h);
    case 16: return launch_pp<16, 1>(h);
    default: break;
  }
  set_error("internal: bad ping-pong dispatch (vpt=%d)", h->pp_vpt);
  return -1;
}

#define B2OT_FUSED_CASES(FN, h, T, MODE)                          \
  if ((h)->nw == 8 && (h)->occ == 1) {                             \
    switch ((h)->vpt) {                                            \
      case 1: return FN<T, 1, MODE, 8, 1>(h);                      \
      case 2: return FN<T, 2, MODE, 8, 1>(h);                      \
      case 4: return FN<T, 4, MODE, 8, 1>(h);                      \
      case 8: return FN<T, 8, MODE, 8, 1>(h);                      \
      default: break;                                              \
    }                                                              \
  } else if ((h)->nw == 8 && (h)->occ == 2) {                      \
    switch ((h)->vpt) {                                            \
      case 1: return FN<T, 1, MODE, 8, 2>(h);                      \
      case 2: return FN<T, 2, MODE, 8, 2>(h);                      \
      case 4: return FN<T, 4, MODE, 8, 2>(h);                      \
      default: break;                                              \
    }                                                              \
  } else if ((h)->nw == 16) {                                      \
    switch ((h)->vpt) {                                            \
      case 1: return FN<T, 1, MODE, 16, 1>(h);                     \
      case 2: return FN<T, 2, MODE, 16, 1>(h);                     \
      case 4: return FN<T, 4, MODE, 16, 1>(h);                     \
      default: break;                                              \
    }                                                              \
  }

#define B2OT_DISPATCH_FUSED(FN, h)                                                   \
  do {                                                                               \
    if ((h)->dtype == B2OT_F32 && (h)->mode == B2OT_MODE_KNOPP && (h)->unbal) {      \
      B2OT_FUSED_CASES(FN, h, float, MODE_UNBAL)                                     \
    } else if ((h)->mode == B2OT_MODE_KNOPP && (h)->unbal) {                         \
      B2OT_FUSED_CASES(FN, h, double, MODE_UNBAL)                                    \
    } else if ((h)->dtype == B2OT_F32 && (h)->mode == B2OT_MODE_KNOPP) {             \
      B2OT_FUSED_CASES(FN, h, float, MODE_KNOPP)                                     \
    } else if ((h)->dtype == B2OT_F32) {                                             \
      B2OT_FUSED_CASES(FN, h, float, MODE_LOG)                                       \
    } else if ((h)->mode == B2OT_MODE_KNOPP) {                                       \
      B2OT_FUSED_CASES(FN, h, double, MODE_KNOPP)                                    \
    } else {                                                                         \
      B2OT_FUSED_CASES(FN, h, double, MODE_LOG)                                      \
    }                                                                                \
    set_error("internal: bad fused dispatch (nw=%d vpt=%d)", (h)->nw, (h)->vpt);     \
    return -1;                                                                       \
  } while (0)

static int dispatch_config_fused(b2ot_dense* h) {
  if (h->pp_vpt > 0) return dispatch_config_pp(h);
  B2OT_DISPATCH_FUSED(config_fused, h);
}
static int dispatch_launch_fused(b2ot_dense* h) {
  if (h->pp_vpt > 0) return dispatch_launch_pp(h);
  B2OT_DISPATCH_FUSED(launch_fused, h);
}

// Decide whether the fused sweep applies and pick (cluster size, vectors per thread, slice width).
static bool plan_fused(b2ot_dense* h) {
  const int vec = h->dtype == B2OT_F64 ? 2 : 4;
  const size_t sz = esize(h->dtype);
  if (h->m % vec != 0 || h->ld % vec != 0) return false;
  if (((uintptr_t)h->K) % 16 != 0) return false;
  if (h->n < 1 || h->m < vec) return false;
  // CTA shape: 16 warps for the MUFU-heavy log domain, 8 warps for the scaling domain
  h->nw = 8;
  h->occ = 1;
  const char* envw = diag_env("B2OT_WARPS");
  if (envw && (atoi(envw) == 8 || atoi(envw) == 16)) h->nw = atoi(envw);
  const char* envo = diag_env("B2OT_OCC");
  if (envo && (atoi(envo) == 1 || atoi(envo) == 2)) h->occ = atoi(envo);
  if (h->nw == 16) h->occ = 1;
  const int vptmax = (h->nw == 8 && h->occ == 1) ? 8 : 4;
  const long long nct = 32LL * h->nw;
  const long long wcapmax = nct * vptmax * vec;
  int max_cs = (h->occ == 2) ? 16 : 8;
  const char* env = diag_env("B2OT_MAX_CLUSTER");
  if (env && atoi(env) >= 1 && atoi(env) <= CSMAX) max_cs = atoi(env);
  if (h->m <= wcapmax) {
    h->cs = 1;
    h->vpt = vptmax;
    for (int vpt = 1; vpt <= vptmax; vpt *= 2) {
      if (nct * vpt * vec >= h->m) {
        h->vpt = vpt;
        break;
      }
    }
    h->W = (int)h->m;
  } else {
    int cs = 2;
    while (cs <= max_cs && (long long)cs * wcapmax < h->m) cs *= 2;
    if (cs > max_cs) return false;
    h->cs = cs;
    h->vpt = vptmax;
    long long w = (h->m + cs - 1) / cs;
    w = (w + vec - 1) / vec * vec;
    h->W = (int)w;
  }
  // K streams through L2 once per sweep: mark it evict-first only when it cannot stay resident
  h->use_hint = ((double)h->n * (double)h->ld * (double)sz > 80.0 * 1024 * 1024) ? 1 : 0;
  const char* envh = diag_env("B2OT_L2_HINT");
  if (envh) h->use_hint = atoi(envh) ? 1 : 0;
  // fp32 log domain: ping-pong schedule (two independent half-CTAs) when the slice is wide enough (B2OT_LOGPP=0: off)
  h->pp_vpt = 0;
  h->pp_occ = 1;
  const char* envp = diag_env("B2OT_LOGPP");
  const bool pp_on = !(envp && envp[0] == '0');
  const bool pp_cand = pp_on && h->mode == B2OT_MODE_LOG && h->dtype == B2OT_F32 && h->nw == 8 && h->occ == 1;
  // Slice width of the ping-pong kernel.  Default: the narrowest cluster that gives slices of <= 4096 columns
  // (8 vectors per thread, 126 registers), so that TWO CTAs = four independent halves fit on an SM
  // (16384^2: 199 -> 187 us per sweep, 65536^2: 3.79 -> 3.48 ms; profiles/r01c_logpp_occ2.log).
  // B2OT_PP_OCC=1 restores one CTA per SM with the widest slices; B2OT_PP_CLUSTER forces a cluster size.
  const char* envo2 = diag_env("B2OT_PP_OCC");
  const bool want_occ2 = !(envo2 && atoi(envo2) == 1);
  const char* envpc = diag_env("B2OT_PP_CLUSTER");
  int force_cs = 0;
  if (envpc && atoi(envpc) >= 1 && atoi(envpc) <= CSMAX && (atoi(envpc) & (atoi(envpc) - 1)) == 0) force_cs = atoi(envpc);
  if (pp_cand && (force_cs || (want_occ2 && h->m > pp_wcap(8)))) {
    int cs = force_cs;
    if (!cs) {
      cs = 1;
      while (cs < CSMAX && (h->m + cs - 1) / cs > pp_wcap(8)) cs *= 2;
    }
    long long w = (h->m + cs - 1) / cs;
    w = (w + vec - 1) / vec * vec;
    if (w >= 1024 && w <= pp_wcap(16) && (force_cs || w <= pp_wcap(8))) {
      h->cs = cs;
      h->W = (int)w;
    }
  }
  if (pp_cand && h->W >= 1024 && h->W <= pp_wcap(16)) {
    h->pp_vpt = h->W <= pp_wcap(4) ? 4 : (h->W <= pp_wcap(8) ? 8 : 16);
    if (want_occ2 && h->pp_vpt <= 8 && (h->m > pp_wcap(8) || force_cs)) h->pp_occ = 2;  // m <= 4096: one CTA with 4 stages is faster
  }
  return true;
}

static const void* resident_fn(const b2ot_dense* h) {
  if (h->mode == B2OT_MODE_KNOPP)
    return h->dtype == B2OT_F32 ? (const void*)resident_solve_kernel<float, MODE_KNOPP> : (const void*)resident_solve_kernel<double, MODE_KNOPP>;
  return h->dtype == B2OT_F32 ? (const void*)resident_solve_kernel<float, MODE_LOG> : (const void*)resident_solve_kernel<double, MODE_LOG>;
}
// Resident solver: eligible when the grid's shared memory holds K twice (a block of rows and a block of columns per
// CTA, at most one CTA per SM).  Scaling domain only.
static bool plan_resident(b2ot_dense* h, int sm_count) {
  h->res_ok = 0;
  if (const char* e = diag_env("B2OT_RESIDENT"))
    if (e[0] == '0') return false;
  if (sm_count > RES_MAXG) sm_count = RES_MAXG;
  const long long grid = std::min<long long>(sm_count, std::max(h->n, h->m));
  const long long rows = (h->n + grid - 1) / grid, cols = (h->m + grid - 1) / grid;
  const size_t sz = esize(h->dtype);
  const size_t smem = ((size_t)rows * h->m + (size_t)cols * h->n + 2 * h->m + h->n + 3 * (rows + cols) + 72) * sz + 64;
  if (smem > 200 * 1024) return false;
  // the exchange words live in the partial-sum scratch of the streaming solver (unused on this path): as many copies
  // as fit, at most 8
  int rep = 2;  // measured at n = m = 1000 fp64: 1 / 2 / 4 / 8 copies = 8.5 / 7.5 / 8.7 / 9.0 us per iteration
  while (rep > 1 && res_xbuf_bytes(h->n, h->m, (int)(sz / 4), rep) > (size_t)NPMAX * h->mp * sz) rep >>= 1;
  if (res_xbuf_bytes(h->n, h->m, (int)(sz / 4), rep) > (size_t)NPMAX * h->mp * sz) return false;
  if (const char* e = diag_env("B2OT_RES_REPLICAS")) rep = std::max(1, std::min(rep, atoi(e)));
  h->res_rep = rep;
  int coop = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
  if (!coop) return false;
  const void* fn = resident_fn(h);
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024 + 64) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, RES_THREADS, smem) != cudaSuccess || per_sm < 1) {
    cudaGetLastError();
    return false;
  }
  h->res_rows = (int)rows;
  h->res_cols = (int)cols;
  h->res_grid = (int)grid;
  h->res_smem = smem;
  h->res_ok = 1;
  return true;
}
// the resident path serves the plain balanced scaling iteration on one device
static inline bool resident_active(const b2ot_dense* h) {
  return h->res_ok && !h->unbal && h->tau <= 0.0 && h->ar == nullptr && h->comm == nullptr;
}
template <typename T>
static int launch_resident(b2ot_dense* h, int n_iters) {
  if (n_iters <= 0) return 0;
  ResParams p;
  memset(&p, 0, sizeof(p));
  p.K = h->K;  p.ld = h->ld;  p.n = h->n;  p.m = h->m;  p.mp = h->mp;
  p.a = h->a;  p.b = h->b;  p.logb = h->logb;
  p.uring = h->uring;  p.vring = h->vring;  p.cbuf = h->cbuf;  p.xbuf = h->partials;
  p.st = h->st;  p.err_hist = h->err_hist;
  p.max_iter = h->max_iter;  p.check_every = h->check_every;  p.stop_thr = h->stop_thr;
  p.n_iters = n_iters;
  p.rows_per_cta = h->res_rows;
  p.cols_per_cta = h->res_cols;
  p.replicas = h->res_rep;
#ifdef B2OT_DIAG
  if (diag_env("B2OT_RES_TRACE")) p.trace = (long long*)h->blk_err;  // 64 iterations x 8 stamps = the 512 leading doubles' bytes... (4 KB of the 8 KB block)
#endif
  void* args[] = {(void*)&p};
  const cudaError_t e = cudaLaunchCooperativeKernel(resident_fn(h), dim3((unsigned)h->res_grid), dim3(RES_THREADS), args,
                                                    h->res_smem, h->stream);
  if (e == cudaErrorCooperativeLaunchTooLarge || e == cudaErrorLaunchOutOfResources) {
    // the grid cannot be co-resident right now (SMs taken away by MPS / a partitioned device): the streaming solver
    // continues the same solve from the same state
    cudaGetLastError();
    h->res_ok = 0;
    return 1;
  }
  B2OT_CUDA(e);
  count_launch();
  return 0;
}

extern "C" int b2ot_dense_create(b2ot_dense** handle, int mode, int dtype, int64_t n, int64_t m, int64_t ld,
                                 const void* K, const void* a, const void* b, const void* logb, void* workspace,
                                 size_t workspace_bytes, int max_iter, int check_every, double stop_thr, int variant,
                                 void* stream) {
  B2OT_REQUIRE(handle != nullptr, "b2ot_dense_create: null handle pointer");
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || mode == B2OT_MODE_LOG, "b2ot_dense_create: bad mode %d", mode);
  B2OT_REQUIRE(variant >= B2OT_VARIANT_AUTO && variant <= B2OT_VARIANT_RESIDENT, "b2ot_dense_create: bad variant %d", variant);
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_dense_create: bad dtype %d", dtype);
  B2OT_REQUIRE(n >= 1 && m >= 1 && ld >= m, "b2ot_dense_create: bad shape n=%lld m=%lld ld=%lld", (long long)n,
               (long long)m, (long long)ld);
  B2OT_REQUIRE(K && a && b && workspace, "b2ot_dense_create: null device pointer");
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || logb != nullptr, "b2ot_dense_create: log mode needs logb");
  B2OT_REQUIRE(max_iter >= 1 && check_every >= 1, "b2ot_dense_create: max_iter and check_every must be >= 1");
  const WsLayout L = ws_layout(n, m, dtype, max_iter, check_every);
  B2OT_REQUIRE(workspace_bytes >= L.total, "b2ot_dense_create: workspace too small (%zu < %zu)", workspace_bytes,
               L.total);
  B2OT_REQUIRE(((uintptr_t)workspace) % 256 == 0, "b2ot_dense_create: workspace must be 256-byte aligned");

  b2ot_dense* h = new (std::nothrow) b2ot_dense();
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_create: out of host memory");
  memset(h, 0, sizeof(*h));
  h->mode = mode;
  h->dtype = dtype;
  h->n = n;
  h->m = m;
  h->ld = ld;
  h->mp = padded_m(m);
  h->K = K;
  h->a = a;
  h->b = b;
  h->logb = logb;
  h->max_iter = max_iter;
  h->check_every = check_every;
  h->stop_thr = stop_thr;
  h->stream = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  h->st = (SolveState*)(ws + L.off_state);
  h->blk_err = (double*)(ws + L.off_blk);
  h->err_hist = (double*)(ws + L.off_err);
  h->err_cap = L.err_cap;
  h->uring = ws + L.off_u;
  h->vring = ws + L.off_v;
  h->cbuf = ws + L.off_c;
  h->partials = ws + L.off_p;

  // column-kernel row split: enough blocks to fill the machine, bounded by the scratch rows
  int sm = 148;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev);
  const long long colblocks = (m + 255) / 256;
  long long ns = (4LL * sm + colblocks - 1) / colblocks;
  if (ns < 1) ns = 1;
  if (ns > n) ns = n;
  if (ns > 64) ns = 64;
  h->nsplit = (int)ns;
  h->nsplit_lse = (int)ns;
  h->fin_blocks = (int)std::min<long long>(colblocks, 1024);

  if (variant == B2OT_VARIANT_AUTO || variant == B2OT_VARIANT_RESIDENT) plan_resident(h, sm);
  if (variant == B2OT_VARIANT_RESIDENT) {
    if (!h->res_ok) {
      delete h;
      set_error("b2ot_dense_create: the resident solver needs K twice within the grid's shared memory (a row block and a "
                "column block per CTA, <= 200 KB), got n=%lld m=%lld", (long long)n, (long long)m);
      return -1;
    }
    variant = B2OT_VARIANT_AUTO;  // the streaming fallback behind it (unbalanced / stabilized / sharded use)
  }
  bool fused_ok = plan_fused(h);
  if (variant == B2OT_VARIANT_FUSED && !fused_ok) {
    delete h;
    set_error("b2ot_dense_create: fused sweep needs 16-byte aligned rows (m, ld multiples of %d), got m=%lld ld=%lld",
              dtype == B2OT_F64 ? 2 : 4, (long long)m, (long long)ld);
    return -1;
  }
  h->variant = (variant == B2OT_VARIANT_SPLIT || !fused_ok) ? B2OT_VARIANT_SPLIT : B2OT_VARIANT_FUSED;
  if (h->variant == B2OT_VARIANT_FUSED) {
    int rc = dispatch_config_fused(h);
    if (rc != 0) {
      if (variant == B2OT_VARIANT_FUSED) {
        delete h;
        return rc;
      }
      h->variant = B2OT_VARIANT_SPLIT;
    }
  }
  h->h_state = (SolveState*)pinned_acquire(sizeof(SolveState));
  if (h->h_state == nullptr) {
    delete h;
    set_error("b2ot_dense_create: cudaMallocHost failed");
    return -2;
  }
  *handle = h;
  return 0;
}

extern "C" int b2ot_dense_set_unbalanced(b2ot_dense* h, double fi_1, double fi_2) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_set_unbalanced: null handle");
  B2OT_REQUIRE(h->mode == B2OT_MODE_KNOPP, "b2ot_dense_set_unbalanced: the handle must be in the scaling domain (KNOPP)");
  B2OT_REQUIRE(h->check_every == 1, "b2ot_dense_set_unbalanced: the unbalanced error is evaluated every iteration; "
                                    "create the handle with check_every = 1 (got %d)", h->check_every);
  B2OT_REQUIRE(fi_1 > 0.0 && fi_1 <= 1.0 && fi_2 > 0.0 && fi_2 <= 1.0,
               "b2ot_dense_set_unbalanced: exponents must be in (0, 1], got %g %g", fi_1, fi_2);
  drop_graph(h);
  h->unbal = 1;
  h->fi1 = fi_1;
  h->fi2 = fi_2;
  // the unbalanced sweep is its own kernel instantiation: set its shared-memory attribute / cluster occupancy
  if (h->variant == B2OT_VARIANT_FUSED) return dispatch_config_fused(h);
  return 0;
}

extern "C" int b2ot_dense_set_allreduce(b2ot_dense* h, b2ot_allreduce_fn fn, void* ctx) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_set_allreduce: null handle");
  h->ar = fn;
  h->ar_ctx = ctx;
  return 0;
}

extern "C" int b2ot_dense_set_comm(b2ot_dense* h, b2ot_comm* comm) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_set_comm: null handle");
  if (comm != nullptr) {
    const size_t need = (size_t)(h->m + 1) * esize(h->dtype);
    B2OT_REQUIRE(comm->view.slot_bytes >= need, "b2ot_dense_set_comm: exchange slots hold %llu bytes, %zu needed",
                 comm->view.slot_bytes, need);
  }
  h->comm = comm;
  h->comm_ll = 0;
  return 0;
}

// Tagged-word exchange (comm.cuh) instead of put kernel + flags: the column sums travel as {value, sequence} words
// inside the finalize kernel.  Needs slots with room for the word region behind the plain one:
// slot_bytes >= roundup64(8 (m + 1)) + 2 (m + 1) sizeof(T).
extern "C" int b2ot_dense_set_comm_ll(b2ot_dense* h, int on) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_set_comm_ll: null handle");
  if (!on) {
    h->comm_ll = 0;
    return 0;
  }
  B2OT_REQUIRE(h->comm != nullptr, "b2ot_dense_set_comm_ll: no communicator set");
  const size_t off = ((size_t)(h->m + 1) * 8 + 63) / 64 * 64;
  const size_t need_ll = off + (size_t)(h->m + 1) * 2 * esize(h->dtype);
  B2OT_REQUIRE(h->comm->view.slot_bytes >= need_ll, "b2ot_dense_set_comm_ll: exchange slots hold %llu bytes, %zu needed",
               h->comm->view.slot_bytes, need_ll);
  h->comm_ll = 1;
  h->comm_ll_off = off;
  return 0;
}

extern "C" int b2ot_dense_info(const b2ot_dense* h, int* variant, int* cluster_size, int* n_clusters, int* stages,
                               int* vecs_per_thread) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_info: null handle");
  if (resident_active(h)) {  // K in the grid's shared memory: n_clusters = CTAs, stages = rows per CTA
    if (variant) *variant = B2OT_VARIANT_RESIDENT;
    if (cluster_size) *cluster_size = 1;
    if (n_clusters) *n_clusters = h->res_grid;
    if (stages) *stages = h->res_rows;
    if (vecs_per_thread) *vecs_per_thread = 0;
    return 0;
  }
  if (variant) *variant = h->variant;
  if (cluster_size) *cluster_size = h->variant == B2OT_VARIANT_FUSED ? h->cs : 0;
  if (n_clusters) *n_clusters = h->variant == B2OT_VARIANT_FUSED ? h->ncl : h->nsplit;
  if (stages) *stages = h->variant == B2OT_VARIANT_FUSED ? h->stages : 0;
  // ping-pong log-domain sweep: reported as 100 + vectors per thread of a half-CTA
  if (vecs_per_thread) *vecs_per_thread = h->variant == B2OT_VARIANT_FUSED ? (h->pp_vpt > 0 ? 100 + h->pp_vpt : h->vpt) : 0;
  return 0;
}

extern "C" int b2ot_dense_destroy(b2ot_dense* h) {
  if (h == nullptr) return 0;
  pinned_release(h->h_state);
  if (h->gexec) cudaGraphExecDestroy(h->gexec);
  delete h;
  return 0;
}

// ---- launch helpers ---------------------------------------------------------------------------
static SplitParams split_params(const b2ot_dense* h, int u_slot_xor, int nsplit, int expect_status) {
  SplitParams p;
  p.K = h->K;
  p.ld = h->ld;
  p.n = h->n;
  p.m = h->m;
  p.mp = h->mp;
  p.a = h->a;
  p.uring = h->uring;
  p.vring = h->vring;
  p.partials = h->partials;
  p.st = h->st;
  p.u_slot_xor = u_slot_xor;
  p.nsplit = nsplit;
  p.expect_status = expect_status;
  p.fi = h->unbal ? h->fi1 : 1.0;
  return p;
}

template <typename T>
static int launch_cols(const b2ot_dense* h, int u_slot_xor) {
  const SplitParams p = split_params(h, u_slot_xor, h->nsplit, ST_RUNNING);
  dim3 grid((unsigned)((h->m + 255) / 256), (unsigned)h->nsplit);
  if (h->mode == B2OT_MODE_KNOPP)
    cols_kernel<T, MODE_KNOPP><<<grid, 256, 0, h->stream>>>(p);
  else
    cols_kernel<T, MODE_LOG><<<grid, 256, 0, h->stream>>>(p);
  B2OT_CHECK_LAUNCH();
  return 0;
}
template <typename T>
static int launch_rows(const b2ot_dense* h) {
  const SplitParams p = split_params(h, 0, h->nsplit, ST_RUNNING);
  const unsigned grid = (unsigned)((h->n + 7) / 8);
  if (h->mode == B2OT_MODE_KNOPP)
    rows_kernel<T, MODE_KNOPP><<<grid, 256, 0, h->stream>>>(p);
  else
    rows_kernel<T, MODE_LOG><<<grid, 256, 0, h->stream>>>(p);
  B2OT_CHECK_LAUNCH();
  return 0;
}

static int call_allreduce(const b2ot_dense* h, void* buf, long long count, int op) {
  if (h->comm) return comm_allreduce(h->comm, buf, count, h->dtype, op, &h->st->status, h->stream);
  const int rc = h->ar(buf, count, h->dtype, op, (void*)h->stream, h->ar_ctx);
  B2OT_REQUIRE(rc == 0, "all-reduce hook failed with code %d", rc);
  return 0;
}

template <typename T>
static int launch_finalize(const b2ot_dense* h, int nsrc, int init, int gate_rebuild = 0) {
  FinParams p;
  memset(&p, 0, sizeof(p));
  const bool multi = h->ar != nullptr || h->comm != nullptr;
  if (h->comm) {
    // fused exchange: "put" reduces this rank's partial rows straight into every peer's slot, the finalize
    // kernel below is the "take" (waits on the flags, sums the slots in rank order)
    b2ot_comm* c = h->comm;
    const unsigned long long seq = ++c->seq;
    if (h->comm_ll && !(h->tau > 0.0)) {
      // tagged-word exchange inside the finalize kernel: no put kernel, no flags (comm.cuh, dense_kernels.cuh)
      p.use_comm = 2;
      p.seq = seq;
      p.comm = c->view;
      p.lsrc = h->partials;
      p.lsrc_ld = h->mp;
      p.lnsrc = nsrc;
      p.uflag_src = &h->st->uflag;
      p.ll_off = h->comm_ll_off;
      p.nsrc = c->view.nranks;
    } else {
    // one 16-byte column vector per thread (comm_put_kernel): 64 blocks at m = 65536 fp32
    const long long vn = 16 / (long long)sizeof(T);
    const unsigned g = (unsigned)std::max<long long>(1, std::min<long long>((h->m / vn + 1 + 255) / 256, 148));
    B2OT_CUDA(launch_pdl(comm_put_kernel<T>, dim3(g), dim3(256), 0, h->stream, c->view, (const T*)h->partials,
                         (long long)h->mp, nsrc, (long long)h->m, (const int*)&h->st->uflag,
                         (const int*)&h->st->status, (int)ST_RUNNING, seq));
    count_launch();
    p.src = c->view.data[c->view.rank] + (size_t)((seq & 1ull) * c->view.nranks) * c->view.slot_bytes;
    p.src_ld = (long long)(c->view.slot_bytes / sizeof(T));
    p.nsrc = c->view.nranks;
    p.extra_flag = (const char*)p.src + (size_t)h->m * sizeof(T);
    p.extra_flag_ld = p.src_ld;
    p.extra_flag_n = c->view.nranks;
    p.use_comm = 1;
    p.seq = seq;
    p.comm = c->view;
    }
  } else if (multi) {
    const unsigned g = (unsigned)((h->m + 1 + 255) / 256);
    reduce_partials_kernel<T><<<g, 256, 0, h->stream>>>(h->partials, h->mp, nsrc, h->cbuf, h->m, h->st, ST_RUNNING);
    B2OT_CHECK_LAUNCH();
    int rc = call_allreduce(h, h->cbuf, h->m + 1, 0);
    if (rc) return rc;
    p.src = h->cbuf;
    p.src_ld = h->mp;
    p.nsrc = 1;
    p.extra_flag = (const char*)h->cbuf + (size_t)h->m * sizeof(T);
    p.extra_flag_ld = 0;
    p.extra_flag_n = 1;
  } else {
    p.src = h->partials;
    p.src_ld = h->mp;
    p.nsrc = nsrc;
    p.extra_flag = nullptr;
  }
  p.b = h->b;
  p.logb = h->logb;
  p.vring = h->vring;
  p.cbuf = h->cbuf;
  p.m = h->m;
  p.mp = h->mp;
  p.st = h->st;
  p.blk_err = h->blk_err;
  p.err_hist = h->err_hist;
  p.max_iter = h->max_iter;
  p.check_every = h->check_every;
  p.stop_thr = h->stop_thr;
  p.init = init;
  p.tau = h->tau;
  p.uring = h->uring;
  p.n = h->n;
  p.gate_rebuild = gate_rebuild;
  int blocks = h->tau > 0.0 ? std::min(h->fin_blocks, 512) : h->fin_blocks;  // second half of blk_err: scaled error
  static const int qsplit_env = [] {
    const char* e = diag_env("B2OT_FIN_QSPLIT");  // diagnosis: 0 = never, 1 = rule below (default), 2 = m <= 32768, 3 = m <= 4096
    return e ? atoi(e) : 1;
  }();
  const long long qsplit_m = (qsplit_env == 2) ? 32768 : (qsplit_env == 3 ? 4096 : ((h->mode == B2OT_MODE_LOG && h->pp_vpt > 0) ? 4096 : 32768));
  if (qsplit_env && !p.use_comm && p.nsrc >= 16 && h->m <= qsplit_m) {  // small problem: spread the partial-row walk over 8 warps
    p.qsplit = 1;
    blocks = (int)std::min<long long>((h->m + 31) / 32, h->tau > 0.0 ? 512 : 1024);
  }
  if (h->mode == B2OT_MODE_KNOPP && h->tau > 0.0)
    B2OT_CUDA(launch_pdl(finalize_kernel<T, MODE_KNOPP, true>, dim3(blocks), dim3(256), 0, h->stream, p));
  else if (p.use_comm == 2 && h->mode == B2OT_MODE_KNOPP)
    B2OT_CUDA(launch_pdl(finalize_ll_kernel<T, MODE_KNOPP>, dim3(blocks), dim3(256), 0, h->stream, p));
  else if (p.use_comm == 2)
    B2OT_CUDA(launch_pdl(finalize_ll_kernel<T, MODE_LOG>, dim3(blocks), dim3(256), 0, h->stream, p));
  else if (h->mode == B2OT_MODE_KNOPP)
    B2OT_CUDA(launch_pdl(finalize_kernel<T, MODE_KNOPP>, dim3(blocks), dim3(256), 0, h->stream, p));
  else
    B2OT_CUDA(launch_pdl(finalize_kernel<T, MODE_LOG>, dim3(blocks), dim3(256), 0, h->stream, p));
  count_launch();
  return 0;
}

// unbalanced iteration: v = (b / K^T u)^fi2, guard, error, stop decision.  Row-sharded runs (all-reduce hook or peer
// communicator): the column sums + the "u not finite" flag are collapsed and SUM-all-reduced like the balanced
// iteration's (reduce_partials_kernel), the three u statistics of the error are MAX-all-reduced (the error couples all
// rows, ot/unbalanced/_sinkhorn.py:774-781); every rank then takes the v step and the stop decision on identical data.
template <typename T>
static int launch_finalize_unbal(const b2ot_dense* h, int nsrc) {
  FinUnbalParams p;
  memset(&p, 0, sizeof(p));
  p.src = h->partials;
  p.src_ld = h->mp;
  p.nsrc = nsrc;
  if (h->ar != nullptr || h->comm != nullptr) {
    const unsigned g = (unsigned)((h->m + 1 + 255) / 256);
    reduce_partials_kernel<T><<<g, 256, 0, h->stream>>>(h->partials, h->mp, nsrc, h->cbuf, h->m, h->st, ST_RUNNING);
    B2OT_CHECK_LAUNCH();
    int rc = call_allreduce(h, h->cbuf, h->m + 1, B2OT_OP_SUM);
    if (rc) return rc;
    void* ustats = (void*)(h->blk_err + 1000);  // behind the 6 x 160 block maxima of finalize_unbal_kernel
    unbal_ustats_kernel<T><<<1, 1024, 0, h->stream>>>(h->uring, h->n, h->st, ustats);
    B2OT_CHECK_LAUNCH();
    rc = call_allreduce(h, ustats, 3, B2OT_OP_MAX);
    if (rc) return rc;
    p.src = h->cbuf;
    p.nsrc = 1;
    p.ustats = ustats;
    p.extra_flag = (const char*)h->cbuf + (size_t)h->m * sizeof(T);
  }
  p.b = h->b;
  p.vring = h->vring;
  p.uring = h->uring;
  p.cbuf = h->cbuf;
  p.n = h->n;
  p.m = h->m;
  p.mp = h->mp;
  p.st = h->st;
  p.blk = h->blk_err;
  p.err_hist = h->err_hist;
  p.max_iter = h->max_iter;
  p.stop_thr = h->stop_thr;
  p.fi = h->fi2;
  int blocks = std::min(h->fin_blocks, 160);  // 6 x blocks doubles of block scratch (1024 available)
  if (nsrc >= 16 && h->m <= 32768) {
    p.qsplit = 1;
    blocks = (int)std::min<long long>((h->m + 31) / 32, 160);
  }
  B2OT_CUDA(launch_pdl(finalize_unbal_kernel<T>, dim3(blocks), dim3(256), 0, h->stream, p));
  count_launch();
  return 0;
}

// exact column log-sum-exp step: v_next = logb - LSE_i(K_ij + u_i)
template <typename T>
static int launch_exact_cols(const b2ot_dense* h, int expect_status, int advance_iter) {
  const int ns = h->nsplit_lse;
  const SplitParams sp = split_params(h, 0, ns, expect_status);
  dim3 grid((unsigned)((h->m + 255) / 256), (unsigned)ns);
  cols_lse_kernel<T><<<grid, 256, 0, h->stream>>>(sp);
  B2OT_CHECK_LAUNCH();
  LseParams lp;
  lp.partials = h->partials;
  lp.m = h->m;
  lp.mp = h->mp;
  lp.nsplit = ns;
  lp.logb = h->logb;
  lp.vring = h->vring;
  lp.st = h->st;
  lp.expect_status = expect_status;
  lp.advance_iter = advance_iter;
  lp.max_iter = h->max_iter;
  const unsigned g1 = (unsigned)((h->m + 255) / 256);
  lse_combine_kernel<T><<<g1, 256, 0, h->stream>>>(lp);
  B2OT_CHECK_LAUNCH();
  char* P = (char*)h->partials;
  if (h->ar || h->comm) {
    int rc = call_allreduce(h, P + (size_t)(2 * ns + 1) * h->mp * sizeof(T), h->m, 1);
    if (rc) return rc;
  }
  lse_rescale_kernel<T><<<g1, 256, 0, h->stream>>>(lp);
  B2OT_CHECK_LAUNCH();
  if (h->ar || h->comm) {
    int rc = call_allreduce(h, P + (size_t)(2 * ns + 2) * h->mp * sizeof(T), h->m, 0);
    if (rc) return rc;
  }
  lse_finish_kernel<T><<<g1, 256, 0, h->stream>>>(lp);
  B2OT_CHECK_LAUNCH();
  lse_advance_kernel<<<1, 1, 0, h->stream>>>(h->st, expect_status, advance_iter);
  B2OT_CHECK_LAUNCH();
  return 0;
}

template <typename T>
static int dense_init_t(b2ot_dense* h, const void* u0, const void* v0) {
  B2OT_CUDA(cudaMemsetAsync(h->st, 0, 256, h->stream));
  if (h->res_ok)  // exchange words of the resident solver: epoch 0 everywhere
    B2OT_CUDA(cudaMemsetAsync(h->partials, 0, res_xbuf_bytes(h->n, h->m, (int)(sizeof(T) / 4), h->res_rep), h->stream));
  B2OT_CUDA(cudaMemcpyAsync(h->uring, u0, (size_t)h->n * sizeof(T), cudaMemcpyDeviceToDevice, h->stream));
  B2OT_CUDA(cudaMemcpyAsync(h->vring, v0, (size_t)h->m * sizeof(T), cudaMemcpyDeviceToDevice, h->stream));
  if (h->unbal) return 0;  // u-then-v order: the first sweep starts from v0, no seeding pass
  if (h->mode == B2OT_MODE_KNOPP) {
    int rc = launch_cols<T>(h, 0);  // c0 = K^T u0
    if (rc) return rc;
    return launch_finalize<T>(h, h->nsplit, 1);
  }
  return launch_exact_cols<T>(h, ST_RUNNING, 0);  // v0' = logb - LSE_i(Mr + u0)
}

// stabilized: the gated absorption chain that follows every finalize (dense_kernels.cuh, absorb_vec_kernel)
template <typename T>
static int launch_absorb(const b2ot_dense* h) {
  const long long mx = std::max(h->n, h->m);
  const unsigned gv = (unsigned)std::min<long long>((mx + 255) / 256, 592);
  absorb_vec_kernel<T><<<gv, 256, 0, h->stream>>>(h->st, h->uring, h->vring, h->n, h->m, h->mp, (double*)h->stab_alpha,
                                                   (double*)h->stab_beta);
  B2OT_CHECK_LAUNCH();
  count_launch();
  constexpr int VEC = Vec<T>::N;
  const bool vec = h->m % VEC == 0 && h->ld % VEC == 0 && h->stab_ld % VEC == 0 &&
                   ((uintptr_t)h->K % 16 == 0) && ((uintptr_t)h->stab_mr % 16 == 0);
  const long long cols = vec ? h->m / VEC : h->m;
  const unsigned gx = (unsigned)((cols + 255) / 256);
  long long gy = std::max<long long>(1, 2368 / gx);
  gy = std::min<long long>(std::min<long long>(gy, NPMAX), h->n);
  if (vec)
    stab_rebuild_kernel<T, true><<<dim3(gx, (unsigned)gy), 256, 0, h->stream>>>(
        (const T*)h->stab_mr, h->stab_ld, (T*)h->K, h->ld, (const double*)h->stab_alpha, (const double*)h->stab_beta, h->n, h->m,
        h->mp, (T*)h->partials, h->st);
  else
    stab_rebuild_kernel<T, false><<<dim3(gx, (unsigned)gy), 256, 0, h->stream>>>(
        (const T*)h->stab_mr, h->stab_ld, (T*)h->K, h->ld, (const double*)h->stab_alpha, (const double*)h->stab_beta, h->n, h->m,
        h->mp, (T*)h->partials, h->st);
  B2OT_CHECK_LAUNCH();
  count_launch();
  return launch_finalize<T>(h, (int)gy, 1, 1);
}

template <typename T>
static int enqueue_iterations(b2ot_dense* h, int n_iters) {
  for (int it = 0; it < n_iters; ++it) {
    int nsrc;
    if (h->variant == B2OT_VARIANT_FUSED) {
      int rc = dispatch_launch_fused(h);
      if (rc) return rc;
      nsrc = h->ncl;
    } else {
      int rc = launch_rows<T>(h);
      if (rc) return rc;
      rc = launch_cols<T>(h, 1);
      if (rc) return rc;
      nsrc = h->nsplit;
    }
    int rc = h->unbal ? launch_finalize_unbal<T>(h, nsrc) : launch_finalize<T>(h, nsrc, 0);
    if (rc) return rc;
    if (h->tau > 0.0) {
      rc = launch_absorb<T>(h);
      if (rc) return rc;
    }
  }
  return 0;
}

// CUDA graph of kGraphIters iterations (sweep + finalize pairs, programmatic-dependent-launch edges included):
// every kernel parameter of a solve is fixed at creation and the whole state lives in device memory, so the same
// graph replays for the entire solve.  Small problems are launch-bound (1000 x 1000 fp64: 21.5 us per iteration with
// two cudaLaunchKernelEx calls, the kernels themselves need a fraction of that); replaying a graph costs one launch
// per kGraphIters iterations.  Single device only (the NCCL hook is host code, the peer exchange carries a sequence
// number in its parameters).  B2OT_GRAPH=0 turns it off; any capture error falls back to plain launches.
static const int kGraphIters = 10;
static bool graphs_enabled() {
  static const bool on = [] {
    const char* e = diag_env("B2OT_GRAPH");
    return !(e && e[0] == '0');
  }();
  return on;
}
static void drop_graph(b2ot_dense* h) {
  if (h->gexec) cudaGraphExecDestroy(h->gexec);
  h->gexec = nullptr;
  h->graph_state = 0;
}
template <typename T>
static void build_graph(b2ot_dense* h) {
  h->graph_state = -1;
  cudaStream_t cap = nullptr;
  if (cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking) != cudaSuccess) {
    cudaGetLastError();
    return;
  }
  const cudaStream_t user = h->stream;
  const unsigned long long before = launch_counter().load();
  cudaGraph_t graph = nullptr;
  bool ok = cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
  if (ok) {
    h->stream = cap;
    const int rc = enqueue_iterations<T>(h, kGraphIters);
    h->stream = user;
    const cudaError_t e = cudaStreamEndCapture(cap, &graph);
    ok = (rc == 0) && (e == cudaSuccess) && graph != nullptr;
  }
  const unsigned long long after = launch_counter().load();
  launch_counter().fetch_sub(after - before);  // nothing was launched while capturing
  if (ok && cudaGraphInstantiate(&h->gexec, graph, 0) == cudaSuccess) {
    h->graph_state = 1;
    h->graph_launches = (int)(after - before);
  } else {
    h->gexec = nullptr;
    cudaGetLastError();  // clear the sticky-free error of the failed capture; plain launches take over
  }
  if (graph) cudaGraphDestroy(graph);
  cudaStreamDestroy(cap);
}

template <typename T>
static int dense_iterate_t(b2ot_dense* h, int n_iters) {
  if (resident_active(h)) {
    const int rc = launch_resident<T>(h, n_iters);
    if (rc <= 0) return rc;  // 1: not launchable, go on with the streaming solver
  }
  if (graphs_enabled() && h->ar == nullptr && h->comm == nullptr && n_iters >= kGraphIters) {
    if (h->graph_state == 0) build_graph<T>(h);
    while (h->graph_state == 1 && n_iters >= kGraphIters) {
      B2OT_CUDA(cudaGraphLaunch(h->gexec, h->stream));
      count_launch(h->graph_launches);
      n_iters -= kGraphIters;
    }
  }
  return enqueue_iterations<T>(h, n_iters);
}

extern "C" int b2ot_dense_init(b2ot_dense* h, const void* u0, const void* v0) {
  B2OT_REQUIRE(h && u0 && v0, "b2ot_dense_init: null argument");
  return h->dtype == B2OT_F32 ? dense_init_t<float>(h, u0, v0) : dense_init_t<double>(h, u0, v0);
}

extern "C" int b2ot_dense_iterate(b2ot_dense* h, int n_iters) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_iterate: null handle");
  return h->dtype == B2OT_F32 ? dense_iterate_t<float>(h, n_iters) : dense_iterate_t<double>(h, n_iters);
}

// Run n_iters iterations with CUDA events bracketing every sweep launch (events are recorded on the
// launch stream).  Reports the average duration of the sweep kernel(s) alone and of a whole iteration.
extern "C" int b2ot_dense_profile(b2ot_dense* h, int n_iters, float* sweep_ms_avg, float* iter_ms_avg) {
  B2OT_REQUIRE(h != nullptr && n_iters >= 1 && n_iters <= 4096, "b2ot_dense_profile: bad arguments");
  if (resident_active(h)) {  // one launch runs all n_iters iterations: sweep and iteration time coincide
    cudaEvent_t e0, e1;
    B2OT_CUDA(cudaEventCreate(&e0));
    B2OT_CUDA(cudaEventCreate(&e1));
    B2OT_CUDA(cudaEventRecord(e0, h->stream));
    int rc = h->dtype == B2OT_F32 ? launch_resident<float>(h, n_iters) : launch_resident<double>(h, n_iters);
    if (rc > 0) {
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
      return b2ot_dense_profile(h, n_iters, sweep_ms_avg, iter_ms_avg);  // streaming path from here on
    }
    if (rc == 0) {
      B2OT_CUDA(cudaEventRecord(e1, h->stream));
      B2OT_CUDA(cudaStreamSynchronize(h->stream));
      float ms = 0.f;
      B2OT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      if (sweep_ms_avg) *sweep_ms_avg = ms / n_iters;
      if (iter_ms_avg) *iter_ms_avg = ms / n_iters;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return rc;
  }
  std::vector<cudaEvent_t> ev(2 * (size_t)n_iters + 1);
  for (auto& e : ev) B2OT_CUDA(cudaEventCreate(&e));
  int rc = 0;
  for (int it = 0; it < n_iters && rc == 0; ++it) {
    B2OT_CUDA(cudaEventRecord(ev[2 * it], h->stream));
    int nsrc = 0;
    if (h->variant == B2OT_VARIANT_FUSED) {
      rc = dispatch_launch_fused(h);
      nsrc = h->ncl;
    } else {
      rc = h->dtype == B2OT_F32 ? launch_rows<float>(h) : launch_rows<double>(h);
      if (rc == 0) rc = h->dtype == B2OT_F32 ? launch_cols<float>(h, 1) : launch_cols<double>(h, 1);
      nsrc = h->nsplit;
    }
    if (rc) break;
    B2OT_CUDA(cudaEventRecord(ev[2 * it + 1], h->stream));
    if (h->unbal)
      rc = h->dtype == B2OT_F32 ? launch_finalize_unbal<float>(h, nsrc) : launch_finalize_unbal<double>(h, nsrc);
    else
      rc = h->dtype == B2OT_F32 ? launch_finalize<float>(h, nsrc, 0) : launch_finalize<double>(h, nsrc, 0);
    if (rc == 0 && h->tau > 0.0) rc = h->dtype == B2OT_F32 ? launch_absorb<float>(h) : launch_absorb<double>(h);
  }
  if (rc == 0) {
    B2OT_CUDA(cudaEventRecord(ev[2 * n_iters], h->stream));
    B2OT_CUDA(cudaStreamSynchronize(h->stream));
    double sweep = 0.0;
    for (int it = 0; it < n_iters; ++it) {
      float ms = 0.f;
      B2OT_CUDA(cudaEventElapsedTime(&ms, ev[2 * it], ev[2 * it + 1]));
      sweep += ms;
    }
    float total = 0.f;
    B2OT_CUDA(cudaEventElapsedTime(&total, ev[0], ev[2 * n_iters]));
    if (sweep_ms_avg) *sweep_ms_avg = (float)(sweep / n_iters);
    if (iter_ms_avg) *iter_ms_avg = total / n_iters;
  }
  for (auto& e : ev) cudaEventDestroy(e);
  return rc;
}

extern "C" int b2ot_dense_poll(b2ot_dense* h, int* status, int* iter) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_poll: null handle");
  for (int guard = 0; guard < 4; ++guard) {
    B2OT_CUDA(cudaMemcpyAsync(h->h_state, h->st, sizeof(SolveState), cudaMemcpyDeviceToHost, h->stream));
    B2OT_CUDA(cudaStreamSynchronize(h->stream));
    if (h->h_state->status != ST_PAUSE_EXACT) break;
    // log domain: a row-frame column sum underflowed; redo this v update with the exact column pass
    int rc = h->dtype == B2OT_F32 ? launch_exact_cols<float>(h, ST_PAUSE_EXACT, 1)
                                  : launch_exact_cols<double>(h, ST_PAUSE_EXACT, 1);
    if (rc) return rc;
  }
  if (status) *status = h->h_state->status;
  if (iter) *iter = h->h_state->iter;
  return 0;
}

static int run_until_stopped(b2ot_dense* h, int it, int poll_every, int* status, int* niter, int* n_err) {
  int rc = 0;
  int st = ST_RUNNING;
  while (st == ST_RUNNING) {
    int chunk = h->max_iter - it;
    if (poll_every > 0 && chunk > poll_every) chunk = poll_every;
    if (chunk < 1) chunk = 1;
    rc = b2ot_dense_iterate(h, chunk);
    if (rc) return rc;
    rc = b2ot_dense_poll(h, &st, &it);
    if (rc) return rc;
  }
  if (status) *status = st;
  if (niter) *niter = h->h_state->niter;
  if (n_err) *n_err = h->h_state->n_err;
  return 0;
}

extern "C" int b2ot_dense_solve(b2ot_dense* h, const void* u0, const void* v0, int poll_every, int* status,
                                int* niter, int* n_err) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_solve: null handle");
  int rc = b2ot_dense_init(h, u0, v0);
  if (rc) return rc;
  return run_until_stopped(h, 0, poll_every, status, niter, n_err);
}

extern "C" int b2ot_dense_set_stabilized(b2ot_dense* h, double tau, const void* Mr, int64_t ldMr, void* alpha_r,
                                         void* beta_r) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_set_stabilized: null handle");
  B2OT_REQUIRE(h->mode == B2OT_MODE_KNOPP && !h->unbal, "b2ot_dense_set_stabilized: balanced scaling-domain handles only");
  B2OT_REQUIRE(h->ar == nullptr && h->comm == nullptr, "b2ot_dense_set_stabilized: single device only");
  B2OT_REQUIRE(tau > 0.0, "b2ot_dense_set_stabilized: tau must be positive");
  B2OT_REQUIRE(Mr && alpha_r && beta_r && ldMr >= h->m, "b2ot_dense_set_stabilized: bad Mr / potentials");
  drop_graph(h);  // the iteration stream changes (finalize parameters, absorption chain)
  h->tau = tau;
  h->stab_mr = Mr;
  h->stab_ld = ldMr;
  h->stab_alpha = alpha_r;
  h->stab_beta = beta_r;
  return 0;
}

extern "C" int b2ot_dense_absorb_count(b2ot_dense* h, int* count) {
  B2OT_REQUIRE(h != nullptr && count != nullptr, "b2ot_dense_absorb_count: null argument");
  B2OT_CUDA(cudaMemcpyAsync(h->h_state, h->st, sizeof(SolveState), cudaMemcpyDeviceToHost, h->stream));
  B2OT_CUDA(cudaStreamSynchronize(h->stream));
  *count = h->h_state->n_absorb;
  return 0;
}

extern "C" int b2ot_dense_export(b2ot_dense* h, void* u_out, void* v_out, double* err_host, int err_capacity,
                                 int* n_err) {
  B2OT_REQUIRE(h != nullptr, "b2ot_dense_export: null handle");
  const long long mx = h->n > h->m ? h->n : h->m;
  const unsigned grid = (unsigned)((mx + 255) / 256);
  if (h->dtype == B2OT_F32)
    export_kernel<float><<<grid, 256, 0, h->stream>>>(h->uring, h->vring, h->n, h->m, h->mp, h->st, u_out, v_out);
  else
    export_kernel<double><<<grid, 256, 0, h->stream>>>(h->uring, h->vring, h->n, h->m, h->mp, h->st, u_out, v_out);
  B2OT_CHECK_LAUNCH();
  B2OT_CUDA(cudaMemcpyAsync(h->h_state, h->st, sizeof(SolveState), cudaMemcpyDeviceToHost, h->stream));
  B2OT_CUDA(cudaStreamSynchronize(h->stream));
  int ne = h->h_state->n_err;
  if (ne > h->err_cap) ne = h->err_cap;
  if (n_err) *n_err = ne;
  if (err_host != nullptr && err_capacity > 0 && ne > 0) {
    const int k = ne < err_capacity ? ne : err_capacity;
    B2OT_CUDA(cudaMemcpyAsync(err_host, h->err_hist, (size_t)k * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    B2OT_CUDA(cudaStreamSynchronize(h->stream));
  }
  return 0;
}
