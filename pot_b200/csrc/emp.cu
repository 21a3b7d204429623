// Host driver + C ABI of the online-cost (empirical) Sinkhorn solver, ot.dist and the lazy plan blocks
// (see include/b2ot.h, emp_kernels.cuh).
#include <algorithm>
#include <vector>
#include <cstring>
#include <new>

#include "../../include/b2ot.h"
#include "comm_host.h"
#include "emp_kernels.cuh"
#include "emp_tc_kernels.cuh"
#include "host_util.h"

using namespace b2ot;

struct b2ot_emp {
  int dtype, d, metric;
  long long n, m, ldxs, ldxt;
  const void *Xs, *Xt, *a, *loga, *logb;
  double reg, stop_thr;
  int max_iter;
  cudaStream_t stream;
  // workspace
  EmpState* st;
  double *blk_err, *err_hist, *err_xchg;
  int err_cap;
  void *fring, *g, *xx, *yy, *cbuf, *pmaxA, *psumA, *pmaxB, *psumB;
  int nsplitA, nsplitB, fin_blocks_f, fin_blocks_g;
  int use_tc, kc;          // tensor-core sweep: packed (split + pre-swizzled) clouds
  int tc_kind;             // 0 = 3xTF32, 1 = 3xFP16 (scaled split)
  float* tc_scales;        // 3xFP16: [s_x, s_y, 1 / (s_x s_y)] followed by the two abs-max words (device)
  void *packs, *packt;
  float* tc_centre;        // common centre of the clouds (TC_MAX_D floats) followed by the partial sums of its reduction
  b2ot_allreduce_fn ar;
  void* ar_ctx;
  b2ot_comm* comm;  // NVLink peer-memory exchange (takes precedence over the hook)
  EmpState* h_state;
};

namespace {

inline size_t esize(int dtype) { return dtype == B2OT_F64 ? 8 : 4; }

int sm_count() {
  int dev = 0, sm = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev);
  return sm;
}

// split of the reduced axis so that the tile grid fills the machine
int pick_nsplit(long long rows, long long cols) {
  const long long row_tiles = (rows + EMP_BM - 1) / EMP_BM;
  const long long col_tiles = (cols + EMP_BN - 1) / EMP_BN;
  long long ns = (4LL * 148 + row_tiles - 1) / row_tiles;
  ns = std::max<long long>(1, std::min<long long>(ns, std::min<long long>(64, col_tiles)));
  return (int)ns;
}

// tensor-core sweep applies to fp32, squared euclidean cost, d <= 64; AUTO takes it when the problem is big enough
// for 128 x 128 tiles to make sense
bool resolve_tc(int dtype, int d, int metric, int variant, long long n, long long m) {
  const bool eligible = dtype == B2OT_F32 && metric == B2OT_METRIC_SQEUCLIDEAN && d <= TC_MAX_D;
  if (variant == B2OT_EMP_TC) return eligible;
  if (variant == B2OT_EMP_SIMT) return false;
  return eligible && std::max(n, m) >= 256;
}
int pick_nsplit_tc(long long rows, long long cols) {
  // CTAs = row tiles x splits of the reduced axis, one CTA per SM at a time: minimise waves x (tiles per CTA + fill /
  // drain), keeping at least 8 column tiles per CTA so the row operand load stays amortised
  const long long row_tiles = (rows + TC_BM - 1) / TC_BM;
  const long long col_tiles = (cols + TC_BN - 1) / TC_BN;
  const long long sms = sm_count();
  const long long ns_max = std::max<long long>(1, std::min<long long>(64, col_tiles / 8));
  long long best = 1;
  double best_cost = 1e300;
  for (long long ns = 1; ns <= ns_max; ++ns) {
    const long long per = (col_tiles + ns - 1) / ns;          // column tiles per CTA (the longest one)
    const long long ctas = row_tiles * ns;
    const long long waves = (ctas + sms - 1) / sms;
    const double cost = (double)waves * (double)(per + 2);    // + ~2 tile times of pipeline fill / drain per CTA
    if (cost < best_cost * 0.999) {
      best_cost = cost;
      best = ns;
    }
  }
  return (int)best;
}

struct EmpWs {
  size_t off_state, off_blk, off_err, off_xchg, off_f, off_g, off_xx, off_yy, off_c, off_pa, off_pb, off_centre,
      off_packs, off_packt, total;
  int err_cap, nsA, nsB, use_tc, kc;
};
EmpWs emp_ws(long long n, long long m, int d, int dtype, int metric, int variant, int max_iter) {
  EmpWs L;
  const size_t sz = esize(dtype);
  L.use_tc = resolve_tc(dtype, d, metric, variant, n, m) ? 1 : 0;
  L.kc = (d + TC_KCH - 1) / TC_KCH;
  L.nsA = L.use_tc ? pick_nsplit_tc(n, m) : pick_nsplit(n, m);
  L.nsB = L.use_tc ? pick_nsplit_tc(m, n) : pick_nsplit(m, n);
  size_t o = 0;
  L.off_state = o;  o += 256;
  L.off_blk = o;    o += 1024 * sizeof(double);
  L.err_cap = max_iter / 10 + 2;
  L.off_err = o;    o = align_up(o + (size_t)L.err_cap * sizeof(double), 256);
  L.off_xchg = o;   o += 256;
  L.off_f = o;      o = align_up(o + 2 * (size_t)n * sz, 256);
  L.off_g = o;      o = align_up(o + (size_t)m * sz, 256);
  L.off_xx = o;     o = align_up(o + (size_t)n * sz, 256);
  L.off_yy = o;     o = align_up(o + (size_t)m * sz, 256);
  L.off_c = o;      o = align_up(o + 3 * (size_t)m * sz, 256);
  L.off_pa = o;     o = align_up(o + 2 * (size_t)L.nsA * n * sz, 256);
  L.off_pb = o;     o = align_up(o + 2 * (size_t)L.nsB * m * sz, 256);
  L.off_centre = o;  // tensor-core path: centre (TC_MAX_D floats) + per-block partial sums (doubles)
  o += L.use_tc ? 256 + (size_t)TC_CENTRE_BLOCKS * TC_MAX_D * sizeof(double) : 0;
  o = align_up(o, 1024);
  L.off_packs = o;  o += L.use_tc ? tc_pack_bytes(n, L.kc) : 0;
  L.off_packt = o;  o += L.use_tc ? tc_pack_bytes(m, L.kc) : 0;
  L.total = o;
  return L;
}

template <typename T>
int launch_sqnorm(const void* X, long long n, int d, long long ld, void* out, cudaStream_t s) {
  if (n <= 0) return 0;
  sqnorm_kernel<T><<<(unsigned)((n + 7) / 8), 256, 0, s>>>((const T*)X, n, d, ld, (T*)out);
  B2OT_CHECK_LAUNCH();
  return 0;
}

template <typename T, int EPI>
int launch_tile(const EmpTileParams& p, cudaStream_t s) {
  if (p.nx <= 0 || p.ny <= 0) return 0;
  dim3 grid((unsigned)((p.nx + EMP_BM - 1) / EMP_BM),
            (unsigned)((EPI == EPI_LSE || EPI == EPI_LOSS) ? p.nsplit : (p.ny + EMP_BN - 1) / EMP_BN));
  emp_tile_kernel<T, EPI><<<grid, EMP_NT, 0, s>>>(p);
  B2OT_CHECK_LAUNCH();
  return 0;
}

// pass A (rows = source points, reduce over targets with g) or pass B (rows = targets, reduce over sources with f)
int launch_tc_pass(const b2ot_emp* h, bool passA) {
  EmpTcParams p;
  memset(&p, 0, sizeof(p));
  if (passA) {
    p.Apack = (const float*)h->packs;  p.nx = h->n;  p.xx = (const float*)h->xx;
    p.Bpack = (const float*)h->packt;  p.ny = h->m;  p.yy = (const float*)h->yy;
    p.poty = (const float*)h->g;
    p.pmax = (float*)h->pmaxA;  p.psum = (float*)h->psumA;  p.nsplit = h->nsplitA;
  } else {
    p.Apack = (const float*)h->packt;  p.nx = h->m;  p.xx = (const float*)h->yy;
    p.Bpack = (const float*)h->packs;  p.ny = h->n;  p.yy = (const float*)h->xx;
    p.poty = (const float*)h->fring;
    p.pot_sel = &h->st->sel_f;
    p.pot_stride = h->n;
    p.pmax = (float*)h->pmaxB;  p.psum = (float*)h->psumB;  p.nsplit = h->nsplitB;
  }
  p.kc = h->kc;
  p.kind = h->tc_kind;
  p.dotscale = h->tc_kind ? h->tc_scales + 2 : nullptr;
  p.k2 = (float)(1.4426950408889634 / h->reg);
  p.status = &h->st->status;
  p.expect_status = ST_RUNNING;
#ifdef B2OT_DIAG
  if (const char* e = diag_env("B2OT_EMP_DBG")) p.dbg = atoi(e);
  if (const char* e = diag_env("B2OT_EMP_TRACE")) p.trace = (long long*)strtoull(e, nullptr, 0);  // device address
#endif

  // cluster along the row tiles (multicast of the column operand); B2OT_EMP_CLUSTER = 1, 2 or 4
  int cs = 1;
  if (const char* e = diag_env("B2OT_EMP_CLUSTER")) {
    const int v = atoi(e);
    if (v == 1 || v == 2 || v == 4) cs = v;
  }
  const long long row_tiles = (p.nx + TC_BM - 1) / TC_BM;
  if (row_tiles < cs) cs = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)((row_tiles + cs - 1) / cs * cs), (unsigned)p.nsplit);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = tc_smem_bytes(2);
  cfg.stream = h->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B2OT_CUDA(cudaLaunchKernelEx(&cfg, emp_tc_lse_kernel, p));
  count_launch();
  return 0;
}

template <typename T>
int launch_lse_pass(const b2ot_emp* h, bool passA) {
  if (h->use_tc) return launch_tc_pass(h, passA);
  EmpTileParams p;
  memset(&p, 0, sizeof(p));
  if (passA) {
    p.X = h->Xs;  p.ldx = h->ldxs;  p.nx = h->n;  p.xx = h->xx;
    p.Y = h->Xt;  p.ldy = h->ldxt;  p.ny = h->m;  p.yy = h->yy;
    p.poty = h->g;
    p.pmax = h->pmaxA;  p.psum = h->psumA;  p.nsplit = h->nsplitA;
  } else {
    p.X = h->Xt;  p.ldx = h->ldxt;  p.nx = h->m;  p.xx = h->yy;
    p.Y = h->Xs;  p.ldy = h->ldxs;  p.ny = h->n;  p.yy = h->xx;
    p.poty = h->fring;  // f lives in a 2-ring; the slot emp_fin_f just committed is selected on the device
    p.pot_sel = &h->st->sel_f;
    p.pot_stride = h->n;
    p.pmax = h->pmaxB;  p.psum = h->psumB;  p.nsplit = h->nsplitB;
  }
  p.d = h->d;
  p.metric = h->metric;
  p.reg = h->reg;
  p.status = &h->st->status;
  p.expect_status = ST_RUNNING;
  return launch_tile<T, EPI_LSE>(p, h->stream);
}

int call_allreduce(const b2ot_emp* h, void* buf, long long count, int dtype, int op) {
  if (h->comm) return comm_allreduce(h->comm, buf, count, dtype, op, &h->st->status, h->stream);
  const int rc = h->ar(buf, count, dtype, op, (void*)h->stream, h->ar_ctx);
  B2OT_REQUIRE(rc == 0, "all-reduce hook failed with code %d", rc);
  return 0;
}

EmpFinParams fin_params(const b2ot_emp* h) {
  EmpFinParams p;
  memset(&p, 0, sizeof(p));
  p.fring = h->fring;
  p.g = h->g;
  p.cbuf = h->cbuf;
  p.st = h->st;
  p.blk_err = h->blk_err;
  p.err_hist = h->err_hist;
  p.max_iter = h->max_iter;
  p.stop_thr = h->stop_thr;
  return p;
}

// f update (+ the pending marginal check and the stop decision)
template <typename T>
int launch_fin_f(const b2ot_emp* h) {
  EmpFinParams p = fin_params(h);
  p.pmax = h->pmaxA;
  p.psum = h->psumA;
  p.nsplit = h->nsplitA;
  p.len = h->n;
  p.logw = h->loga;
  p.w = h->a;
  const bool multi = h->ar != nullptr || h->comm != nullptr;
  p.err_xchg = multi ? h->err_xchg : nullptr;
  emp_fin_f_kernel<T><<<h->fin_blocks_f, 256, 0, h->stream>>>(p);
  B2OT_CHECK_LAUNCH();
  if (multi) {
    int rc = call_allreduce(h, h->err_xchg, 1, B2OT_F64, B2OT_OP_SUM);
    if (rc) return rc;
    emp_decide_kernel<<<1, 1, 0, h->stream>>>(h->st, h->err_xchg, h->err_hist, h->max_iter, h->stop_thr);
    B2OT_CHECK_LAUNCH();
  }
  return 0;
}

// g update: log b - LSE over all source rows (combined across ranks when row-sharded)
template <typename T>
int launch_fin_g(const b2ot_emp* h) {
  EmpFinParams p = fin_params(h);
  p.pmax = h->pmaxB;
  p.psum = h->psumB;
  p.nsplit = h->nsplitB;
  p.len = h->m;
  p.logw = h->logb;
  if (!h->ar && !h->comm) {
    p.stage = 0;
    emp_fin_g_kernel<T><<<h->fin_blocks_g, 256, 0, h->stream>>>(p);
    B2OT_CHECK_LAUNCH();
    return 0;
  }
  char* cb = (char*)h->cbuf;
  const size_t sz = sizeof(T);
  p.stage = 1;
  emp_fin_g_kernel<T><<<h->fin_blocks_g, 256, 0, h->stream>>>(p);
  B2OT_CHECK_LAUNCH();
  int rc = call_allreduce(h, cb + (size_t)h->m * sz, h->m, h->dtype, B2OT_OP_MAX);
  if (rc) return rc;
  p.stage = 2;
  emp_fin_g_kernel<T><<<h->fin_blocks_g, 256, 0, h->stream>>>(p);
  B2OT_CHECK_LAUNCH();
  rc = call_allreduce(h, cb + 2 * (size_t)h->m * sz, h->m, h->dtype, B2OT_OP_SUM);
  if (rc) return rc;
  p.stage = 3;
  emp_fin_g_kernel<T><<<h->fin_blocks_g, 256, 0, h->stream>>>(p);
  B2OT_CHECK_LAUNCH();
  return 0;
}

template <typename T>
int emp_iterate_t(b2ot_emp* h, int n_iters) {
  for (int it = 0; it < n_iters; ++it) {
    int rc = launch_lse_pass<T>(h, true);
    if (rc == 0) rc = launch_fin_f<T>(h);
    if (rc == 0) rc = launch_lse_pass<T>(h, false);
    if (rc == 0) rc = launch_fin_g<T>(h);
    if (rc) return rc;
  }
  return 0;
}

// the check of the LAST iteration ((max_iter % 10 == 0) only): one more pass A, no f commit
template <typename T>
int emp_tail_t(b2ot_emp* h) {
  int rc = launch_lse_pass<T>(h, true);
  if (rc == 0) rc = launch_fin_f<T>(h);
  return rc;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" size_t b2ot_emp_workspace_bytes(int64_t n, int64_t m, int d, int dtype, int metric, int variant,
                                           int max_iter) {
  return emp_ws(n, m, d, dtype, metric, variant, max_iter).total;
}

extern "C" int b2ot_emp_create(b2ot_emp** handle, int dtype, int64_t n, int64_t m, int d, const void* Xs, int64_t ldXs,
                               const void* Xt, int64_t ldXt, const void* a, const void* loga, const void* logb,
                               double reg, int metric, void* workspace, size_t workspace_bytes, int max_iter,
                               double stop_thr, int variant, void* stream) {
  B2OT_REQUIRE(handle != nullptr, "b2ot_emp_create: null handle pointer");
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_emp_create: bad dtype %d", dtype);
  B2OT_REQUIRE(n >= 1 && m >= 1 && d >= 1 && ldXs >= d && ldXt >= d, "b2ot_emp_create: bad shape n=%lld m=%lld d=%d",
               (long long)n, (long long)m, d);
  B2OT_REQUIRE(Xs && Xt && a && loga && logb && workspace, "b2ot_emp_create: null device pointer");
  B2OT_REQUIRE(reg > 0, "b2ot_emp_create: reg must be > 0");
  B2OT_REQUIRE(metric == B2OT_METRIC_SQEUCLIDEAN || metric == B2OT_METRIC_EUCLIDEAN, "b2ot_emp_create: bad metric %d",
               metric);
  B2OT_REQUIRE(max_iter >= 1, "b2ot_emp_create: max_iter must be >= 1");
  B2OT_REQUIRE(variant >= B2OT_EMP_AUTO && variant <= B2OT_EMP_TC, "b2ot_emp_create: bad variant %d", variant);
  const EmpWs L = emp_ws(n, m, d, dtype, metric, variant, max_iter);
  B2OT_REQUIRE(variant != B2OT_EMP_TC || L.use_tc,
               "b2ot_emp_create: the tensor-core sweep needs fp32, sqeuclidean and d <= %d (got dtype %d, metric %d, d %d)",
               TC_MAX_D, dtype, metric, d);
  B2OT_REQUIRE(workspace_bytes >= L.total, "b2ot_emp_create: workspace too small (%zu < %zu)", workspace_bytes,
               L.total);
  B2OT_REQUIRE(((uintptr_t)workspace) % 256 == 0, "b2ot_emp_create: workspace must be 256-byte aligned");
  b2ot_emp* h = new (std::nothrow) b2ot_emp();
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_create: out of host memory");
  memset(h, 0, sizeof(*h));
  h->dtype = dtype;
  h->d = d;
  h->metric = metric;
  h->n = n;
  h->m = m;
  h->ldxs = ldXs;
  h->ldxt = ldXt;
  h->Xs = Xs;
  h->Xt = Xt;
  h->a = a;
  h->loga = loga;
  h->logb = logb;
  h->reg = reg;
  h->stop_thr = stop_thr;
  h->max_iter = max_iter;
  h->stream = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  h->st = (EmpState*)(ws + L.off_state);
  h->blk_err = (double*)(ws + L.off_blk);
  h->err_hist = (double*)(ws + L.off_err);
  h->err_xchg = (double*)(ws + L.off_xchg);
  h->err_cap = L.err_cap;
  h->fring = ws + L.off_f;
  h->g = ws + L.off_g;
  h->xx = ws + L.off_xx;
  h->yy = ws + L.off_yy;
  h->cbuf = ws + L.off_c;
  const size_t sz = esize(dtype);
  h->nsplitA = L.nsA;
  h->nsplitB = L.nsB;
  h->pmaxA = ws + L.off_pa;
  h->psumA = ws + L.off_pa + (size_t)L.nsA * n * sz;
  h->pmaxB = ws + L.off_pb;
  h->psumB = ws + L.off_pb + (size_t)L.nsB * m * sz;
  h->use_tc = L.use_tc;
  h->tc_kind = 1;                              // 3xFP16 scaled split; b2ot_emp_set_operand switches to 3xTF32
  h->kc = (d + 63) / 64;                       // the workspace is sized for the (larger) TF32 packing
  h->tc_scales = (float*)((char*)(ws + L.off_xchg) + 64);
  h->packs = ws + L.off_packs;
  h->packt = ws + L.off_packt;
  h->tc_centre = (float*)(ws + L.off_centre);
  if (h->use_tc) {
    B2OT_CUDA(cudaFuncSetAttribute(emp_tc_lse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)tc_smem_bytes(2)));
  }
  h->fin_blocks_f = (int)std::max<long long>(1, std::min<long long>((n + 255) / 256, 1024));
  h->fin_blocks_g = (int)std::max<long long>(1, std::min<long long>((m + 255) / 256, 1024));
  (void)sm_count;
  static_assert(sizeof(EmpState) <= kPinnedBlock, "EmpState must fit a pinned pool block");
  h->h_state = (EmpState*)pinned_acquire(sizeof(EmpState));
  if (h->h_state == nullptr) {
    delete h;
    set_error("b2ot_emp_create: cudaMallocHost failed");
    return -2;
  }
  *handle = h;
  return 0;
}

extern "C" int b2ot_emp_set_operand(b2ot_emp* h, int operand) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_set_operand: null handle");
  B2OT_REQUIRE(operand == B2OT_EMP_OPERAND_F16 || operand == B2OT_EMP_OPERAND_TF32,
               "b2ot_emp_set_operand: bad operand kind %d", operand);
  h->tc_kind = operand == B2OT_EMP_OPERAND_TF32 ? 0 : 1;
  h->kc = h->tc_kind ? (h->d + 63) / 64 : (h->d + TC_KCH - 1) / TC_KCH;
  return 0;
}

extern "C" int b2ot_emp_set_allreduce(b2ot_emp* h, b2ot_allreduce_fn fn, void* ctx) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_set_allreduce: null handle");
  h->ar = fn;
  h->ar_ctx = ctx;
  return 0;
}

extern "C" int b2ot_emp_set_comm(b2ot_emp* h, b2ot_comm* comm) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_set_comm: null handle");
  if (comm != nullptr) {
    const size_t need = (size_t)h->m * esize(h->dtype);
    B2OT_REQUIRE(comm->view.slot_bytes >= need, "b2ot_emp_set_comm: exchange slots hold %llu bytes, %zu needed",
                 comm->view.slot_bytes, need);
  }
  h->comm = comm;
  return 0;
}

extern "C" int b2ot_emp_destroy(b2ot_emp* h) {
  if (h == nullptr) return 0;
  pinned_release(h->h_state);
  delete h;
  return 0;
}

extern "C" int b2ot_emp_init(b2ot_emp* h, const void* f0, const void* g0) {
  B2OT_REQUIRE(h && f0 && g0, "b2ot_emp_init: null argument");
  const size_t sz = esize(h->dtype);
  B2OT_CUDA(cudaMemsetAsync(h->st, 0, 256, h->stream));
  B2OT_CUDA(cudaMemcpyAsync(h->fring, f0, (size_t)h->n * sz, cudaMemcpyDeviceToDevice, h->stream));
  B2OT_CUDA(cudaMemcpyAsync(h->g, g0, (size_t)h->m * sz, cudaMemcpyDeviceToDevice, h->stream));
  if (h->dtype == B2OT_F32 && !h->use_tc) {
    int rc = launch_sqnorm<float>(h->Xs, h->n, h->d, h->ldxs, h->xx, h->stream);
    if (rc) return rc;
    return launch_sqnorm<float>(h->Xt, h->m, h->d, h->ldxt, h->yy, h->stream);
  }
  if (h->dtype == B2OT_F32) {
    // tensor-core sweep: centre both clouds on the mean of the target cloud, then norms, split (hi, lo) and
    // pre-swizzle of the centred coordinates, once per solve
    const float* Xs = (const float*)h->Xs;
    const float* Xt = (const float*)h->Xt;
    float* centre = h->tc_centre;
    double* cpart = (double*)((char*)centre + 256);
    emp_centre_partial_kernel<<<TC_CENTRE_BLOCKS, 256, 0, h->stream>>>(Xt, h->m, h->d, h->ldxt, cpart);
    B2OT_CHECK_LAUNCH();
    emp_centre_final_kernel<<<1, TC_MAX_D, 0, h->stream>>>(cpart, TC_CENTRE_BLOCKS, h->m, h->d, centre);
    B2OT_CHECK_LAUNCH();
    emp_sqnorm_centred_kernel<<<(unsigned)((h->n + 7) / 8), 256, 0, h->stream>>>(Xs, h->n, h->d, h->ldxs, centre,
                                                                               (float*)h->xx);
    B2OT_CHECK_LAUNCH();
    emp_sqnorm_centred_kernel<<<(unsigned)((h->m + 7) / 8), 256, 0, h->stream>>>(Xt, h->m, h->d, h->ldxt, centre,
                                                                               (float*)h->yy);
    B2OT_CHECK_LAUNCH();
    const long long us = (h->n + TC_BM - 1) / TC_BM * TC_BM * h->kc * 8;
    const long long ut = (h->m + TC_BM - 1) / TC_BM * TC_BM * h->kc * 8;
    if (h->tc_kind == 0) {
      emp_tc_pack_kernel<<<(unsigned)((us + 255) / 256), 256, 0, h->stream>>>(Xs, h->n, h->d, h->ldxs, h->kc, centre,
                                                                             (float*)h->packs);
      B2OT_CHECK_LAUNCH();
      emp_tc_pack_kernel<<<(unsigned)((ut + 255) / 256), 256, 0, h->stream>>>(Xt, h->m, h->d, h->ldxt, h->kc, centre,
                                                                             (float*)h->packt);
      B2OT_CHECK_LAUNCH();
      return 0;
    }
    // 3xFP16: one power-of-two scale per cloud from its largest |centred coordinate|, then the scaled split
    unsigned int* maxbits = (unsigned int*)(h->tc_scales + 4);
    B2OT_CUDA(cudaMemsetAsync(maxbits, 0, 2 * sizeof(unsigned int), h->stream));
    const unsigned gs = (unsigned)std::min<long long>((h->n * h->d + 255) / 256, 1184);
    const unsigned gt = (unsigned)std::min<long long>((h->m * h->d + 255) / 256, 1184);
    emp_absmax_kernel<<<gs, 256, 0, h->stream>>>(Xs, h->n, h->d, h->ldxs, centre, maxbits);
    B2OT_CHECK_LAUNCH();
    emp_absmax_kernel<<<gt, 256, 0, h->stream>>>(Xt, h->m, h->d, h->ldxt, centre, maxbits + 1);
    B2OT_CHECK_LAUNCH();
    emp_tc_scale_kernel<<<1, 32, 0, h->stream>>>(maxbits, h->tc_scales);
    B2OT_CHECK_LAUNCH();
    emp_tc_pack_f16_kernel<<<(unsigned)((us + 255) / 256), 256, 0, h->stream>>>(
        Xs, h->n, h->d, h->ldxs, h->kc, centre, h->tc_scales, (unsigned char*)h->packs);
    B2OT_CHECK_LAUNCH();
    emp_tc_pack_f16_kernel<<<(unsigned)((ut + 255) / 256), 256, 0, h->stream>>>(
        Xt, h->m, h->d, h->ldxt, h->kc, centre, h->tc_scales + 1, (unsigned char*)h->packt);
    B2OT_CHECK_LAUNCH();
    return 0;
  }
  int rc = launch_sqnorm<double>(h->Xs, h->n, h->d, h->ldxs, h->xx, h->stream);
  if (rc) return rc;
  return launch_sqnorm<double>(h->Xt, h->m, h->d, h->ldxt, h->yy, h->stream);
}

extern "C" int b2ot_emp_iterate(b2ot_emp* h, int n_iters) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_iterate: null handle");
  return h->dtype == B2OT_F32 ? emp_iterate_t<float>(h, n_iters) : emp_iterate_t<double>(h, n_iters);
}

// Measurement aid: n_iters iterations with CUDA events (on the solve's stream) around each of the two sweep
// launches of an iteration; reports the average duration of ONE sweep kernel and of a whole iteration.
extern "C" int b2ot_emp_profile(b2ot_emp* h, int n_iters, float* sweep_ms_avg, float* iter_ms_avg) {
  B2OT_REQUIRE(h != nullptr && n_iters >= 1 && n_iters <= 1024, "b2ot_emp_profile: bad arguments");
  std::vector<cudaEvent_t> ev(4 * (size_t)n_iters + 1);
  for (auto& e : ev) B2OT_CUDA(cudaEventCreate(&e));
  const bool f32 = h->dtype == B2OT_F32;
  int rc = 0;
  for (int it = 0; it < n_iters && rc == 0; ++it) {
    B2OT_CUDA(cudaEventRecord(ev[4 * it], h->stream));
    rc = f32 ? launch_lse_pass<float>(h, true) : launch_lse_pass<double>(h, true);
    B2OT_CUDA(cudaEventRecord(ev[4 * it + 1], h->stream));
    if (rc == 0) rc = f32 ? launch_fin_f<float>(h) : launch_fin_f<double>(h);
    B2OT_CUDA(cudaEventRecord(ev[4 * it + 2], h->stream));
    if (rc == 0) rc = f32 ? launch_lse_pass<float>(h, false) : launch_lse_pass<double>(h, false);
    B2OT_CUDA(cudaEventRecord(ev[4 * it + 3], h->stream));
    if (rc == 0) rc = f32 ? launch_fin_g<float>(h) : launch_fin_g<double>(h);
  }
  if (rc == 0) {
    B2OT_CUDA(cudaEventRecord(ev[4 * n_iters], h->stream));
    B2OT_CUDA(cudaStreamSynchronize(h->stream));
    double sweep = 0.0;
    for (int it = 0; it < n_iters; ++it) {
      float a = 0.f, b = 0.f;
      B2OT_CUDA(cudaEventElapsedTime(&a, ev[4 * it], ev[4 * it + 1]));
      B2OT_CUDA(cudaEventElapsedTime(&b, ev[4 * it + 2], ev[4 * it + 3]));
      sweep += a + b;
    }
    float total = 0.f;
    B2OT_CUDA(cudaEventElapsedTime(&total, ev[0], ev[4 * n_iters]));
    if (sweep_ms_avg) *sweep_ms_avg = (float)(sweep / (2.0 * n_iters));
    if (iter_ms_avg) *iter_ms_avg = total / n_iters;
  }
  for (auto& e : ev) cudaEventDestroy(e);
  return rc;
}

extern "C" int b2ot_emp_info(const b2ot_emp* h, int* variant, int* nsplit_rows, int* nsplit_cols) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_info: null handle");
  if (variant) *variant = h->use_tc ? B2OT_EMP_TC : B2OT_EMP_SIMT;
  if (nsplit_rows) *nsplit_rows = h->nsplitA;
  if (nsplit_cols) *nsplit_cols = h->nsplitB;
  return 0;
}

extern "C" int b2ot_emp_poll(b2ot_emp* h, int* status, int* iter) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_poll: null handle");
  B2OT_CUDA(cudaMemcpyAsync(h->h_state, h->st, sizeof(EmpState), cudaMemcpyDeviceToHost, h->stream));
  B2OT_CUDA(cudaStreamSynchronize(h->stream));
  if (status) *status = h->h_state->status;
  if (iter) *iter = h->h_state->iter;
  return 0;
}

extern "C" int b2ot_emp_solve(b2ot_emp* h, const void* f0, const void* g0, int poll_every, int* status, int* niter,
                              int* n_err) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_solve: null handle");
  int rc = b2ot_emp_init(h, f0, g0);
  if (rc) return rc;
  int st = ST_RUNNING, it = 0;
  while (st == ST_RUNNING && it < h->max_iter) {
    int chunk = h->max_iter - it;
    if (poll_every > 0 && chunk > poll_every) chunk = poll_every;
    rc = b2ot_emp_iterate(h, chunk);
    if (rc) return rc;
    if (it + chunk >= h->max_iter && h->max_iter % 10 == 0) {
      rc = h->dtype == B2OT_F32 ? emp_tail_t<float>(h) : emp_tail_t<double>(h);
      if (rc) return rc;
    }
    rc = b2ot_emp_poll(h, &st, &it);
    if (rc) return rc;
  }
  if (status) *status = st;
  if (niter) *niter = h->h_state->niter;
  if (n_err) *n_err = h->h_state->n_err;
  return 0;
}

extern "C" int b2ot_emp_export(b2ot_emp* h, void* f_out, void* g_out, double* err_host, int err_capacity, int* n_err) {
  B2OT_REQUIRE(h != nullptr, "b2ot_emp_export: null handle");
  const long long mx = h->n > h->m ? h->n : h->m;
  const unsigned grid = (unsigned)((mx + 255) / 256);
  if (h->dtype == B2OT_F32)
    emp_export_kernel<float><<<grid, 256, 0, h->stream>>>(h->fring, h->g, h->n, h->m, h->st, f_out, g_out);
  else
    emp_export_kernel<double><<<grid, 256, 0, h->stream>>>(h->fring, h->g, h->n, h->m, h->st, f_out, g_out);
  B2OT_CHECK_LAUNCH();
  B2OT_CUDA(cudaMemcpyAsync(h->h_state, h->st, sizeof(EmpState), cudaMemcpyDeviceToHost, h->stream));
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

// ---- ot.dist and lazy plan blocks ----
template <typename T, int EPI>
static int dist_like(const void* X, const void* Y, const void* f, const void* g, int64_t n, int64_t m, int d,
                     int64_t ldx, int64_t ldy, double reg, int metric, void* out, int64_t ldo, void* scratch,
                     cudaStream_t s) {
  T* xx = (T*)scratch;
  T* yy = xx + n;
  int rc = launch_sqnorm<T>(X, n, d, ldx, xx, s);
  if (rc == 0) rc = launch_sqnorm<T>(Y, m, d, ldy, yy, s);
  if (rc) return rc;
  EmpTileParams p;
  memset(&p, 0, sizeof(p));
  p.X = X;  p.ldx = ldx;  p.nx = n;  p.xx = xx;
  p.Y = Y;  p.ldy = ldy;  p.ny = m;  p.yy = yy;
  p.potx = f;
  p.poty = g;
  p.out = out;
  p.ldo = ldo;
  p.d = d;
  p.metric = metric;
  p.reg = reg;
  return launch_tile<T, EPI>(p, s);
}

extern "C" size_t b2ot_dist_scratch_bytes(int64_t n, int64_t m, int dtype) { return (size_t)(n + m) * esize(dtype); }

extern "C" int b2ot_dist(int dtype, const void* X, const void* Y, int64_t n, int64_t m, int d, int64_t ldX, int64_t ldY,
                         int metric, void* out, int64_t ldOut, void* scratch, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_dist: bad dtype %d", dtype);
  B2OT_REQUIRE(metric == B2OT_METRIC_SQEUCLIDEAN || metric == B2OT_METRIC_EUCLIDEAN, "b2ot_dist: bad metric %d", metric);
  B2OT_REQUIRE(X && Y && out && scratch && n >= 1 && m >= 1 && d >= 1, "b2ot_dist: bad argument");
  return dtype == B2OT_F32 ? dist_like<float, EPI_DIST>(X, Y, nullptr, nullptr, n, m, d, ldX, ldY, 1.0, metric, out,
                                                        ldOut, scratch, (cudaStream_t)stream)
                           : dist_like<double, EPI_DIST>(X, Y, nullptr, nullptr, n, m, d, ldX, ldY, 1.0, metric, out,
                                                         ldOut, scratch, (cudaStream_t)stream);
}

extern "C" int b2ot_emp_plan_block(int dtype, const void* Xs, const void* Xt, const void* f, const void* g, int64_t n,
                                   int64_t m, int d, int64_t ldXs, int64_t ldXt, double reg, int metric, void* out,
                                   int64_t ldOut, void* scratch, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_emp_plan_block: bad dtype %d", dtype);
  B2OT_REQUIRE(Xs && Xt && f && g && out && scratch && n >= 1 && m >= 1 && d >= 1 && reg > 0,
               "b2ot_emp_plan_block: bad argument");
  return dtype == B2OT_F32 ? dist_like<float, EPI_PLAN>(Xs, Xt, f, g, n, m, d, ldXs, ldXt, reg, metric, out, ldOut,
                                                        scratch, (cudaStream_t)stream)
                           : dist_like<double, EPI_PLAN>(Xs, Xt, f, g, n, m, d, ldXs, ldXt, reg, metric, out, ldOut,
                                                         scratch, (cudaStream_t)stream);
}

/* loss = sum_ij C_ij exp(f_i + g_j - C_ij/reg) without materialising cost or plan */
extern "C" size_t b2ot_emp_loss_scratch_bytes(int64_t n, int64_t m, int dtype) {
  const size_t row_tiles = (size_t)((n + EMP_BM - 1) / EMP_BM);
  return align_up((size_t)(n + m) * esize(dtype), 256) + row_tiles * (size_t)pick_nsplit(n, m) * sizeof(double);
}

template <typename T>
static int emp_loss_t(const void* Xs, const void* Xt, const void* f, const void* g, int64_t n, int64_t m, int d,
                      int64_t ldx, int64_t ldy, double reg, int metric, double* loss, void* scratch, cudaStream_t s) {
  T* xx = (T*)scratch;
  T* yy = xx + n;
  double* blk = (double*)((char*)scratch + align_up((size_t)(n + m) * sizeof(T), 256));
  int rc = launch_sqnorm<T>(Xs, n, d, ldx, xx, s);
  if (rc == 0) rc = launch_sqnorm<T>(Xt, m, d, ldy, yy, s);
  if (rc) return rc;
  EmpTileParams p;
  memset(&p, 0, sizeof(p));
  p.X = Xs;  p.ldx = ldx;  p.nx = n;  p.xx = xx;
  p.Y = Xt;  p.ldy = ldy;  p.ny = m;  p.yy = yy;
  p.potx = f;
  p.poty = g;
  p.blk_out = blk;
  p.nsplit = pick_nsplit(n, m);
  p.d = d;
  p.metric = metric;
  p.reg = reg;
  rc = launch_tile<T, EPI_LOSS>(p, s);
  if (rc) return rc;
  const int nblk = (int)(((n + EMP_BM - 1) / EMP_BM) * p.nsplit);
  emp_sum_blocks_kernel<<<1, 256, 0, s>>>(blk, nblk, loss);
  B2OT_CHECK_LAUNCH();
  return 0;
}

extern "C" int b2ot_emp_loss(int dtype, const void* Xs, const void* Xt, const void* f, const void* g, int64_t n,
                             int64_t m, int d, int64_t ldXs, int64_t ldXt, double reg, int metric, double* loss_dev,
                             void* scratch, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_emp_loss: bad dtype %d", dtype);
  B2OT_REQUIRE(Xs && Xt && f && g && loss_dev && scratch && n >= 1 && m >= 1 && d >= 1 && reg > 0,
               "b2ot_emp_loss: bad argument");
  return dtype == B2OT_F32 ? emp_loss_t<float>(Xs, Xt, f, g, n, m, d, ldXs, ldXt, reg, metric, loss_dev, scratch,
                                               (cudaStream_t)stream)
                           : emp_loss_t<double>(Xs, Xt, f, g, n, m, d, ldXs, ldXt, reg, metric, loss_dev, scratch,
                                                (cudaStream_t)stream);
}
