// C ABI of the plan / objective epilogue (see include/b2ot.h, value_kernels.cuh).
#include <algorithm>
#include <cstdint>
#include <mutex>

#include "../../include/b2ot.h"
#include "host_util.h"
#include "value_kernels.cuh"

using namespace b2ot;

static const int kPvBlocks = 148 * 8;
static const int kColSplit = 32;

extern "C" size_t b2ot_dense_plan_value_scratch_bytes(int64_t, int64_t) { return 3 * kPvBlocks * sizeof(double); }

// the callers' combinations get branch-free instantiations (value_kernels.cuh, SPEC); anything else the generic kernel
template <typename T, int MODE>
static int launch_plan_value(const PlanValueParams& p, bool vect, double* values, cudaStream_t s) {
  const int spec = (p.terms & 15) | (p.K ? PV_SPEC_K : 0) | (p.C ? PV_SPEC_C : 0) | (p.out ? PV_SPEC_OUT : 0);
#define B2OT_PV_CASE(SP) \
  case SP:               \
    plan_value_kernel<T, MODE, true, SP><<<p.nblk, 256, 0, s>>>(p); \
    break;
  bool done = vect;
  if (vect) {
    if (MODE == MODE_LOG) {
      switch (spec) {
        B2OT_PV_CASE(PV_LINEAR | PV_KL | PV_SPEC_OUT)       // ot.solve, KL regulariser: M in, plan out
        B2OT_PV_CASE(PV_LINEAR | PV_ENTROPY | PV_SPEC_OUT)  // ot.solve, entropy regulariser
        B2OT_PV_CASE(PV_SPEC_K | PV_SPEC_OUT)               // plan (+ row marginal) from Mr: epsilon scaling
        default:
          done = false;
      }
    } else {
      switch (spec) {
        B2OT_PV_CASE(PV_SPEC_K | PV_LINEAR | PV_KL | PV_SPEC_OUT)  // unbalanced: plan, cost, KL(plan, a b^T)
        B2OT_PV_CASE(PV_KMULQ | PV_SPEC_OUT)                       // K = exp(-M/reg) * a b^T
        default:
          done = false;
      }
    }
  }
#undef B2OT_PV_CASE
  if (done) {
  } else if (vect)
    plan_value_kernel<T, MODE, true><<<p.nblk, 256, 0, s>>>(p);
  else
    plan_value_kernel<T, MODE, false><<<p.nblk, 256, 0, s>>>(p);
  B2OT_CHECK_LAUNCH();
  plan_value_reduce_kernel<<<3, 256, 0, s>>>(p.blk, p.nblk, values);
  B2OT_CHECK_LAUNCH();
  return 0;
}

static inline bool aligned16(const void* q) { return q == nullptr || ((uintptr_t)q % 16) == 0; }

extern "C" int b2ot_dense_plan_value(int mode, int dtype, const void* K, const void* M, const void* Cref, double reg,
                                     const void* u, const void* v, const void* a, const void* b, void* plan_out,
                                     void* rowsum_out, int64_t n, int64_t m, int64_t ldK, int64_t ldM, int64_t ldC,
                                     int64_t ldOut, int terms, double* values_dev, void* scratch_dev, void* stream) {
  B2OT_REQUIRE(mode == B2OT_MODE_KNOPP || mode == B2OT_MODE_LOG, "b2ot_dense_plan_value: bad mode %d", mode);
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_dense_plan_value: bad dtype %d", dtype);
  B2OT_REQUIRE(n >= 1 && m >= 1, "b2ot_dense_plan_value: bad shape n=%lld m=%lld", (long long)n, (long long)m);
  B2OT_REQUIRE(u && v && values_dev && scratch_dev, "b2ot_dense_plan_value: null device pointer");
  B2OT_REQUIRE(K != nullptr || M != nullptr, "b2ot_dense_plan_value: need K or the cost M");
  B2OT_REQUIRE(!(terms & B2OT_VALUE_LINEAR) || M != nullptr, "b2ot_dense_plan_value: the linear term needs M");
  B2OT_REQUIRE(K != nullptr || reg > 0.0, "b2ot_dense_plan_value: rebuilding K from the cost needs reg > 0");
  const bool need_q = (terms & B2OT_VALUE_KL) || ((terms & B2OT_VALUE_KMULQ) && K == nullptr);
  B2OT_REQUIRE(!need_q || Cref != nullptr || (a != nullptr && b != nullptr),
               "b2ot_dense_plan_value: the KL reference needs c or (a, b)");
  PlanValueParams p;
  p.K = K;
  p.M = M;
  p.C = Cref;
  p.u = u;
  p.v = v;
  p.a = a;
  p.b = b;
  p.out = plan_out;
  p.rowsum = rowsum_out;
  p.n = n;
  p.m = m;
  p.ldK = ldK;
  p.ldM = ldM;
  p.ldC = ldC;
  p.ldO = ldOut;
  p.reg = reg;
  p.terms = terms;
  p.nblk = (int)std::min<long long>(n, kPvBlocks);
  p.blk = (double*)scratch_dev;
  const int vec = dtype == B2OT_F64 ? 2 : 4;
  const bool vect = m % vec == 0 && (K == nullptr || ldK % vec == 0) && (M == nullptr || ldM % vec == 0) &&
                    (Cref == nullptr || ldC % vec == 0) && (plan_out == nullptr || ldOut % vec == 0) &&
                    aligned16(K) && aligned16(M) && aligned16(Cref) && aligned16(plan_out) && aligned16(v) &&
                    aligned16(b);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B2OT_F32)
    return mode == B2OT_MODE_KNOPP ? launch_plan_value<float, MODE_KNOPP>(p, vect, values_dev, s)
                                   : launch_plan_value<float, MODE_LOG>(p, vect, values_dev, s);
  return mode == B2OT_MODE_KNOPP ? launch_plan_value<double, MODE_KNOPP>(p, vect, values_dev, s)
                                 : launch_plan_value<double, MODE_LOG>(p, vect, values_dev, s);
}

extern "C" size_t b2ot_colsum_scratch_bytes(int64_t, int64_t m) { return (size_t)kColSplit * (size_t)m * sizeof(double); }

extern "C" int b2ot_colsum(int dtype, const void* P, int64_t n, int64_t m, int64_t ld, void* out, void* scratch_dev,
                           void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_colsum: bad dtype %d", dtype);
  B2OT_REQUIRE(P && out && scratch_dev && n >= 1 && m >= 1 && ld >= m, "b2ot_colsum: bad arguments");
  const int S = (int)std::min<long long>(n, kColSplit);
  dim3 grid((unsigned)((m + 255) / 256), (unsigned)S);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B2OT_F32) {
    colsum_kernel<float><<<grid, 256, 0, s>>>((const float*)P, n, m, ld, (double*)scratch_dev);
    B2OT_CHECK_LAUNCH();
    colsum_reduce_kernel<float><<<grid.x, 256, 0, s>>>((const double*)scratch_dev, m, S, (float*)out);
  } else {
    colsum_kernel<double><<<grid, 256, 0, s>>>((const double*)P, n, m, ld, (double*)scratch_dev);
    B2OT_CHECK_LAUNCH();
    colsum_reduce_kernel<double><<<grid.x, 256, 0, s>>>((const double*)scratch_dev, m, S, (double*)out);
  }
  B2OT_CHECK_LAUNCH();
  return 0;
}

// ---- b2ot_colsum without any allocation: the row-block partials live in a file-scope device array ----
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, serving loops, harnesses that forbid it).  kColSplit x B2OT_COLSUM_STATIC_MAX_COLS doubles = 4 MiB,
// held by every GPU that loads the library until the process exits.  Host code must not pass the array's name to a kernel
// (on the host it is a stand-in symbol, not a device address): the device address is looked up once per device with
// cudaGetSymbolAddress, with that device current, and cached.  All calls share the array: calls that overlap on the GPU
// (different streams, or graph branches without an edge between them) race.  Nothing has to start at zero - a call
// writes every partial [0, S) x [0, m) it later reads (colsum_kernel), so the previous call's contents are dead.
static __device__ __align__(256) double g_colsum_static_scratch[(size_t)kColSplit * B2OT_COLSUM_STATIC_MAX_COLS];

static const int kMaxDevices = 64;
static std::mutex g_colsum_static_mutex;
static double* g_colsum_static_ptr[kMaxDevices] = {};

extern "C" int b2ot_colsum_static(int dtype, const void* P, int64_t n, int64_t m, int64_t ld, void* out, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_colsum_static: bad dtype %d", dtype);
  B2OT_REQUIRE(P && out && n >= 1 && m >= 1 && ld >= m, "b2ot_colsum_static: bad arguments");
  B2OT_REQUIRE(m <= (int64_t)B2OT_COLSUM_STATIC_MAX_COLS,
               "b2ot_colsum_static: m=%lld exceeds the static scratch (at most %d columns); use b2ot_colsum with "
               "caller-provided scratch",
               (long long)m, (int)B2OT_COLSUM_STATIC_MAX_COLS);
  int dev = -1;
  B2OT_CUDA(cudaGetDevice(&dev));
  B2OT_REQUIRE(dev >= 0 && dev < kMaxDevices, "b2ot_colsum_static: device ordinal %d out of range", dev);
  double* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_colsum_static_mutex);
    if (g_colsum_static_ptr[dev] == nullptr) {
      double* ptr = nullptr;
      B2OT_CUDA(cudaGetSymbolAddress((void**)&ptr, g_colsum_static_scratch));
      B2OT_REQUIRE(ptr != nullptr, "b2ot_colsum_static: cudaGetSymbolAddress returned a null address");
      g_colsum_static_ptr[dev] = ptr;
    }
    scratch = g_colsum_static_ptr[dev];
  }
  const int S = (int)std::min<long long>(n, kColSplit);
  dim3 grid((unsigned)((m + 255) / 256), (unsigned)S);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == B2OT_F32) {
    colsum_kernel<float><<<grid, 256, 0, s>>>((const float*)P, n, m, ld, scratch);
    B2OT_CHECK_LAUNCH();
    colsum_reduce_kernel<float><<<grid.x, 256, 0, s>>>(scratch, m, S, (float*)out);
  } else {
    colsum_kernel<double><<<grid, 256, 0, s>>>((const double*)P, n, m, ld, scratch);
    B2OT_CHECK_LAUNCH();
    colsum_reduce_kernel<double><<<grid.x, 256, 0, s>>>(scratch, m, S, (double*)out);
  }
  B2OT_CHECK_LAUNCH();
  return 0;
}
