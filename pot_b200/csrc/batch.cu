// Host driver + C ABI of the batched on-chip Sinkhorn solver (see include/b2ot.h, batch_kernels.cuh).
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "../../include/b2ot.h"
#include "batch_kernels.cuh"
#include "batch_hyb_kernels.cuh"
#include "host_util.h"

using namespace b2ot;

namespace {

constexpr size_t kSmemLimit = 227 * 1024 - 1024;  // dynamic shared memory we allow a CTA to ask for

struct BatchCfg {
  int cpl, cs, rpc;
  size_t smem;
};

static inline size_t esize(int dtype) { return dtype == B2OT_F64 ? 8 : 4; }

// columns per lane (power of two, <= 32) and the smallest cluster whose CTAs can hold their rows
static bool batch_config(int mode, int dtype, long long n, long long m, BatchCfg* c) {
  if (n < 1 || m < 1 || m > 1024 || n > (1 << 20)) return false;
  int cpl = 1;
  while (cpl * 32 < m) cpl *= 2;
  for (int cs = 1; cs <= BATCH_CSMAX; cs *= 2) {
    if (cs > n && cs > 1) break;
    const int rpc = (int)((n + cs - 1) / cs);
    const BatchSmem L = batch_smem_layout(rpc, (int)m, cpl * 32, cs, (int)esize(dtype), mode);
    if (L.total <= kSmemLimit) {
      c->cpl = cpl;
      c->cs = cs;
      c->rpc = rpc;
      c->smem = L.total;
      return true;
    }
  }
  return false;
}

template <typename T, int MODE, int CPL, int NTRIP = 0>
static int launch_one(const BatchParams& p, const BatchCfg& c, cudaStream_t stream) {
  auto kern = batch_solve_kernel<T, MODE, CPL, NTRIP>;
  static thread_local size_t configured = 0;
  if (c.smem > configured) {
    B2OT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem));
    configured = c.smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(p.B * c.cs), 1, 1);
  cfg.blockDim = dim3(batch_threads(CPL), 1, 1);
  cfg.dynamicSmemBytes = c.smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)c.cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B2OT_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
  count_launch();
  return 0;
}

// Register-resident fast path (batch_kernels.cuh, NTRIP): full tiles only -- m == 32 cpl, n == cs * rpc and every
// CTA's rows are a whole number (1 or 2) of trips of warps x rows-per-group.  B2OT_BATCH_REGTILE=0 turns it off.
template <typename T>
static int fast_trips(const BatchParams& p, const BatchCfg& c) {
  static const bool on = [] {
    const char* e = diag_env("B2OT_BATCH_REGTILE");
    return !(e && e[0] == '0');
  }();
  if (!on || (c.cpl != 4 && c.cpl != 8)) return 0;
  if (p.m != c.cpl * 32 || (long long)c.cs * c.rpc != p.n) return 0;
  const int nw = batch_threads(c.cpl) / 32;
  const int rg = batch_rows_per_group(c.cpl * (int)(sizeof(T) / 4));
  if (c.rpc % (nw * rg) != 0) return 0;
  const int nt = c.rpc / (nw * rg);
  if (nt < 1 || nt > 2 || nt * rg * c.cpl * (int)(sizeof(T) / 4) > 64) return 0;
  return nt;
}

template <typename T, int MODE>
static int launch_cpl(const BatchParams& p, const BatchCfg& c, cudaStream_t s) {
  const int nt = fast_trips<T>(p, c);
  if (nt == 1 && c.cpl == 4) return launch_one<T, MODE, 4, 1>(p, c, s);
  if (nt == 2 && c.cpl == 4) return launch_one<T, MODE, 4, 2>(p, c, s);
  if (nt == 1 && c.cpl == 8) return launch_one<T, MODE, 8, 1>(p, c, s);
  if (nt == 2 && c.cpl == 8) return launch_one<T, MODE, 8, 2>(p, c, s);
  switch (c.cpl) {
    case 1: return launch_one<T, MODE, 1>(p, c, s);
    case 2: return launch_one<T, MODE, 2>(p, c, s);
    case 4: return launch_one<T, MODE, 4>(p, c, s);
    case 8: return launch_one<T, MODE, 8>(p, c, s);
    case 16: return launch_one<T, MODE, 16>(p, c, s);
    case 32: return launch_one<T, MODE, 32>(p, c, s);
    default: break;
  }
  set_error("internal: bad batch dispatch (cpl=%d)", c.cpl);
  return -1;
}

// One whole 256-column fp32 problem per CTA, half in registers, half in shared memory (batch_hyb_kernels.cuh).
// B2OT_BATCH_HYB=0 (diagnosis build) falls back to the cluster kernel.
static bool hyb_enabled() {
  static const bool on = [] {
    const char* e = diag_env("B2OT_BATCH_HYB");
    return !(e && e[0] == '0');
  }();
  return on;
}
template <int MODE, int NS>
static int launch_hyb_one(const BatchParams& p, cudaStream_t stream) {
  auto kern = batch_hyb_kernel<MODE, NS>;
  const size_t smem = hyb_smem_bytes(128 + 64 * NS);
  static thread_local bool configured = false;
  if (!configured) {
    B2OT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  kern<<<(unsigned)p.B, HYB_NT, smem, stream>>>(p);
  B2OT_CHECK_LAUNCH();
  return 0;
}
static int launch_hyb(int mode, const BatchParams& p, cudaStream_t s) {
  if (p.n == 256)
    return mode == B2OT_MODE_KNOPP ? launch_hyb_one<MODE_KNOPP, 2>(p, s) : launch_hyb_one<MODE_LOG, 2>(p, s);
  return mode == B2OT_MODE_KNOPP ? launch_hyb_one<MODE_KNOPP, 1>(p, s) : launch_hyb_one<MODE_LOG, 1>(p, s);
}

static int launch(int mode, int dtype, const BatchParams& p, const BatchCfg& c, cudaStream_t s) {
  if (hyb_enabled() && hyb_eligible(dtype == B2OT_F32, p.n, p.m) && p.B < (1ll << 31)) return launch_hyb(mode, p, s);
  if (dtype == B2OT_F32)
    return mode == B2OT_MODE_KNOPP ? launch_cpl<float, MODE_KNOPP>(p, c, s) : launch_cpl<float, MODE_LOG>(p, c, s);
  return mode == B2OT_MODE_KNOPP ? launch_cpl<double, MODE_KNOPP>(p, c, s) : launch_cpl<double, MODE_LOG>(p, c, s);
}

}  // namespace

extern "C" size_t b2ot_batch_workspace_bytes(int64_t B) {
  return 256 + align_up((size_t)(B > 0 ? B : 1) * sizeof(double), 256);
}

extern "C" int b2ot_batch_config(int mode, int dtype, int64_t n, int64_t m, int* cluster_size, int* smem_bytes,
                                 int* cols_per_lane) {
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || mode == B2OT_MODE_LOG, "b2ot_batch_config: bad mode %d", mode);
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_batch_config: bad dtype %d", dtype);
  BatchCfg c;
  if (!batch_config(mode, dtype, n, m, &c)) {
    set_error("b2ot_batch: a %lld x %lld problem does not fit the on-chip solver (m <= 1024 and n*m*%d bytes <= ~1.6 MB)",
              (long long)n, (long long)m, (int)esize(dtype));
    return -4;
  }
  if (cluster_size) *cluster_size = c.cs;
  if (smem_bytes) *smem_bytes = (int)c.smem;
  if (cols_per_lane) *cols_per_lane = c.cpl;
  return 0;
}

extern "C" int b2ot_batch_solve(int mode, int dtype, int64_t B, int64_t n, int64_t m, const void* X, int transform,
                                double reg, const void* Mval, const void* a, const void* b, void* pa, void* pb,
                                void* T_out, void* logT_out, double* value_linear, double* value_reg, int reg_type,
                                int max_iter, double tol, void* workspace, size_t workspace_bytes, int* n_iters,
                                int flags, void* stream) {
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || mode == B2OT_MODE_LOG, "b2ot_batch_solve: bad mode %d", mode);
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_batch_solve: bad dtype %d", dtype);
  B2OT_REQUIRE(B >= 1 && n >= 1 && m >= 1, "b2ot_batch_solve: bad shape B=%lld n=%lld m=%lld", (long long)B,
               (long long)n, (long long)m);
  B2OT_REQUIRE(X && a && b && pa && pb && workspace, "b2ot_batch_solve: null device pointer");
  B2OT_REQUIRE(transform >= 0 && transform <= 2, "b2ot_batch_solve: bad transform %d", transform);
  B2OT_REQUIRE(transform == 0 || reg > 0, "b2ot_batch_solve: reg must be > 0");
  B2OT_REQUIRE(max_iter >= 0, "b2ot_batch_solve: max_iter must be >= 0");
  B2OT_REQUIRE(reg_type >= 0 && reg_type <= 2, "b2ot_batch_solve: bad reg_type %d", reg_type);
  B2OT_REQUIRE(workspace_bytes >= b2ot_batch_workspace_bytes(B), "b2ot_batch_solve: workspace too small");
  B2OT_REQUIRE(((uintptr_t)workspace) % 256 == 0, "b2ot_batch_solve: workspace must be 256-byte aligned");
  BatchCfg c;
  if (!batch_config(mode, dtype, n, m, &c)) {
    set_error("b2ot_batch_solve: a %lld x %lld problem does not fit the on-chip solver", (long long)n, (long long)m);
    return -4;
  }
  cudaStream_t s = (cudaStream_t)stream;
  BatchParams p;
  memset(&p, 0, sizeof(p));
  p.X = X;
  p.Mval = Mval;
  p.a = a;
  p.b = b;
  p.pa = pa;
  p.pb = pb;
  p.T_out = T_out;
  p.logT_out = (mode == B2OT_MODE_LOG) ? logT_out : nullptr;
  p.value_linear = value_linear;
  p.value_reg = value_reg;
  p.st = (BatchState*)workspace;
  p.err_b = (double*)((char*)workspace + 256);
  p.B = B;
  p.n = (int)n;
  p.m = (int)m;
  p.rpc = c.rpc;
  p.transform = transform;
  p.reg = reg;
  p.reg_type = reg_type;
  p.tol = tol;
  p.force_exact = (flags & B2OT_BATCH_FORCE_EXACT) ? 1 : 0;
  B2OT_CUDA(cudaMemsetAsync(workspace, 0, 256, s));

  const bool want_epilogue = T_out || logT_out || value_linear || value_reg;
  int rc = 0;
  if (!(tol > 0.0) || max_iter == 0) {
    // the stop test can never fire (err < tol with tol <= 0): the whole solve is ONE launch, on chip
    p.it_begin = 0;
    p.it_end = max_iter;
    p.load_state = 0;
    p.check = 0;
    p.epilogue = want_epilogue ? 1 : 0;
    p.expect_running = 0;
    rc = launch(mode, dtype, p, c, s);
    if (rc) return rc;
    if (n_iters) *n_iters = max_iter - 1;
    return 0;
  }
  // launches end on the reference's check iterations (it % 10 == 0); state travels through pa / pb
  BatchState host_state;
  int it = 0, launches = 0;
  bool stopped = false;
  while (it < max_iter && !stopped) {
    int end = (it % 10 == 0) ? it + 1 : ((it + 9) / 10) * 10 + 1;
    if (end > max_iter) end = max_iter;
    p.it_begin = it;
    p.it_end = end;
    p.load_state = (it > 0) ? 1 : 0;
    p.check = ((end - 1) % 10 == 0) ? 1 : 0;
    p.epilogue = 0;
    p.expect_running = 1;
    rc = launch(mode, dtype, p, c, s);
    if (rc) return rc;
    it = end;
    if (++launches % 16 == 0) {  // bounded look-ahead: read the sticky status word
      B2OT_CUDA(cudaMemcpyAsync(&host_state, p.st, sizeof(BatchState), cudaMemcpyDeviceToHost, s));
      B2OT_CUDA(cudaStreamSynchronize(s));
      stopped = host_state.status != ST_RUNNING;
    }
  }
  if (want_epilogue) {
    p.it_begin = p.it_end = 0;
    p.load_state = 1;
    p.check = 0;
    p.epilogue = 1;
    p.expect_running = 0;
    rc = launch(mode, dtype, p, c, s);
    if (rc) return rc;
  }
  if (n_iters) {
    B2OT_CUDA(cudaMemcpyAsync(&host_state, p.st, sizeof(BatchState), cudaMemcpyDeviceToHost, s));
    B2OT_CUDA(cudaStreamSynchronize(s));
    *n_iters = host_state.status == ST_CONVERGED ? host_state.n_iters : max_iter - 1;
  }
  return 0;
}
