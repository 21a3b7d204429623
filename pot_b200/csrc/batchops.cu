// C ABI of the batched cost matrices and per-problem reductions of ot.batch (see include/b2ot.h, batchops_kernels.cuh).
#include <algorithm>

#include "../../include/b2ot.h"
#include "batchops_kernels.cuh"
#include "host_util.h"

using namespace b2ot;

namespace {

template <typename T>
static int launch_dist(const BatchDistParams& p, int metric, cudaStream_t s) {
  const unsigned grid = (unsigned)(p.B * p.tiles_n * p.tiles_m);
  switch (metric) {
    case B2OT_BMETRIC_SQEUCLIDEAN: batch_dist_kernel<T, BDM_SQEUCLIDEAN><<<grid, 256, 0, s>>>(p); break;
    case B2OT_BMETRIC_EUCLIDEAN: batch_dist_kernel<T, BDM_EUCLIDEAN><<<grid, 256, 0, s>>>(p); break;
    case B2OT_BMETRIC_MINKOWSKI: batch_dist_kernel<T, BDM_MINKOWSKI><<<grid, 256, 0, s>>>(p); break;
    default: batch_dist_kernel<T, BDM_KL><<<grid, 256, 0, s>>>(p); break;
  }
  B2OT_CHECK_LAUNCH();
  return 0;
}

struct ReducePlan {
  int nsplit;
  long long chunk;
};

// enough blocks to fill the chip (B * nsplit >= ~8 per SM), chunks of whole 1024-element trips
static ReducePlan reduce_plan(long long B, long long per) {
  ReducePlan r;
  const long long want_blocks = 148 * 8;
  long long ns = (want_blocks + B - 1) / B;
  const long long max_ns = std::max<long long>(1, per / 4096);
  ns = std::max<long long>(1, std::min(ns, max_ns));
  long long chunk = (per + ns - 1) / ns;
  chunk = (chunk + 1023) / 1024 * 1024;
  r.chunk = chunk;
  r.nsplit = (int)((per + chunk - 1) / chunk);
  if (r.nsplit < 1) r.nsplit = 1;
  return r;
}

template <typename T>
static int launch_reduce(int op, const void* A, const void* Bm, long long B, long long per, double eps, double sign,
                         void* out, double* partial, cudaStream_t s) {
  const ReducePlan r = reduce_plan(B, per);
  const bool aligned = ((uintptr_t)A % 16 == 0) && (op != B2OT_BREDUCE_DOT || (uintptr_t)Bm % 16 == 0) &&
                       (per * sizeof(T)) % 16 == 0;
  const unsigned grid = (unsigned)(B * r.nsplit);
  if (op == B2OT_BREDUCE_DOT)
    batch_reduce_kernel<T, BR_DOT><<<grid, 256, 0, s>>>((const T*)A, (const T*)Bm, per, r.nsplit, r.chunk, eps,
                                                        aligned ? 1 : 0, partial);
  else
    batch_reduce_kernel<T, BR_ENTROPY><<<grid, 256, 0, s>>>((const T*)A, nullptr, per, r.nsplit, r.chunk, eps,
                                                            aligned ? 1 : 0, partial);
  B2OT_CHECK_LAUNCH();
  batch_reduce_final_kernel<T><<<(unsigned)((B + 255) / 256), 256, 0, s>>>(partial, B, r.nsplit, sign, (T*)out);
  B2OT_CHECK_LAUNCH();
  return 0;
}

}  // namespace

extern "C" {

int b2ot_batch_dist(int dtype, const void* X, const void* Y, int64_t B, int64_t n, int64_t m, int d, int metric,
                    double p, const void* w, double eps, void* out, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_batch_dist: dtype %d", dtype);
  B2OT_REQUIRE(metric >= B2OT_BMETRIC_SQEUCLIDEAN && metric <= B2OT_BMETRIC_KL, "b2ot_batch_dist: metric %d", metric);
  B2OT_REQUIRE(B >= 0 && n >= 0 && m >= 0 && d >= 0, "b2ot_batch_dist: negative size");
  if (B == 0 || n == 0 || m == 0) return 0;
  B2OT_REQUIRE(X && Y && out, "b2ot_batch_dist: null pointer");
  BatchDistParams q;
  q.X = X;
  q.Y = Y;
  q.w = w;
  q.out = out;
  q.B = B;
  q.n = n;
  q.m = m;
  q.d = d;
  q.tiles_n = (n + BD_TILE - 1) / BD_TILE;
  q.tiles_m = (m + BD_TILE - 1) / BD_TILE;
  q.p = p;
  q.eps = eps;
  const size_t es = dtype == B2OT_F64 ? 8 : 4;
  q.vec_ok = ((uintptr_t)out % 16 == 0 && (m * es) % 16 == 0) ? 1 : 0;
  B2OT_REQUIRE((long long)q.B * q.tiles_n * q.tiles_m < (1ll << 31), "b2ot_batch_dist: too many tiles");
  cudaStream_t s = (cudaStream_t)stream;
  return dtype == B2OT_F64 ? launch_dist<double>(q, metric, s) : launch_dist<float>(q, metric, s);
}

size_t b2ot_batch_reduce_scratch_bytes(int64_t B, int64_t per) {
  if (B <= 0 || per <= 0) return 16;
  return (size_t)B * (size_t)reduce_plan(B, per).nsplit * sizeof(double);
}

int b2ot_batch_reduce(int dtype, int op, const void* A, const void* Bm, int64_t B, int64_t per, double eps,
                      double sign, void* out, void* scratch_dev, void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_batch_reduce: dtype %d", dtype);
  B2OT_REQUIRE(op == B2OT_BREDUCE_DOT || op == B2OT_BREDUCE_ENTROPY, "b2ot_batch_reduce: op %d", op);
  B2OT_REQUIRE(B >= 0 && per >= 0, "b2ot_batch_reduce: negative size");
  if (B == 0) return 0;
  B2OT_REQUIRE(out && scratch_dev && (per == 0 || (A && (op != B2OT_BREDUCE_DOT || Bm))), "b2ot_batch_reduce: null pointer");
  B2OT_REQUIRE((long long)B * reduce_plan(B, std::max<int64_t>(per, 1)).nsplit < (1ll << 31), "b2ot_batch_reduce: too many blocks");
  cudaStream_t s = (cudaStream_t)stream;
  const long long pr = per;
  return dtype == B2OT_F64 ? launch_reduce<double>(op, A, Bm, B, pr, eps, sign, out, (double*)scratch_dev, s)
                           : launch_reduce<float>(op, A, Bm, B, pr, eps, sign, out, (double*)scratch_dev, s);
}

}  // extern "C"
