// C ABI of the strided batched matrix product behind the Gromov-Wasserstein contraction (include/b2ot.h,
// gemm_kernels.cuh).
#include "../../include/b2ot.h"
#include "gemm_kernels.cuh"
#include "host_util.h"

using namespace b2ot;

extern "C" int b2ot_strided_gemm(int dtype, int64_t Zb, int64_t Zf, int64_t Sf, int64_t M, int64_t N, int64_t K,
                                 const void* A, int64_t sAb, int64_t sAf, int64_t sAi, int64_t sAk, const void* B,
                                 int64_t sBb, int64_t sBf, int64_t sBk, int64_t sBj, const void* C0, int64_t sC0b,
                                 int64_t sC0i, void* C, int64_t sCb, int64_t sCf, int64_t sCi, double alpha, double beta,
                                 void* stream) {
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_strided_gemm: bad dtype %d", dtype);
  B2OT_REQUIRE(Zb >= 0 && Zf >= 1 && Sf >= 1 && M >= 0 && N >= 0 && K >= 0, "b2ot_strided_gemm: bad shape");
  B2OT_REQUIRE(A && B && C, "b2ot_strided_gemm: null device pointer");
  if (Zb == 0 || M == 0 || N == 0) return 0;
  B2OT_REQUIRE(Zb * Zf <= 65535, "b2ot_strided_gemm: more than 65535 (batch x feature) outputs (%lld)", (long long)(Zb * Zf));
  GemmParams p;
  p.A = A;  p.B = B;  p.C0 = C0;  p.C = C;
  p.M = M;  p.N = N;  p.K = K;  p.Zf = Zf;  p.Sf = Sf;
  p.sAb = sAb;  p.sAf = sAf;  p.sAi = sAi;  p.sAk = sAk;
  p.sBb = sBb;  p.sBf = sBf;  p.sBk = sBk;  p.sBj = sBj;
  p.sC0b = sC0b;  p.sC0i = sC0i;
  p.sCb = sCb;  p.sCf = sCf;  p.sCi = sCi;
  p.alpha = alpha;  p.beta = beta;
  cudaStream_t s = (cudaStream_t)stream;
  // fp32: 128 x 128 tiles (8 x 8 per thread) once they give two waves of blocks, fp64: 128 x 64 (8 x 4) from about one
  // wave on; smaller outputs take 64 x 64 tiles (measured, scripts/bench_gemm.py: n = 1000 fp32 0.21 vs 0.38 ms, fp64
  // 0.46 vs 0.38 ms)
  const long long z = Zb * Zf;
  const bool small = dtype == B2OT_F32 ? ((M + 127) / 128) * ((N + 127) / 128) * z < 296
                                       : ((M + 127) / 128) * ((N + 63) / 64) * z < 120;
  auto launch = [&](auto kern, int bm, int bn) {
    const dim3 grid((unsigned)((N + bn - 1) / bn), (unsigned)((M + bm - 1) / bm), (unsigned)z);
    kern<<<grid, 256, 0, s>>>(p);
  };
  B2OT_REQUIRE((M + 63) / 64 <= 65535, "b2ot_strided_gemm: M too large (%lld)", (long long)M);
  if (dtype == B2OT_F32) {
    if (small) launch(strided_gemm_kernel<float, 4, 4>, 64, 64);
    else launch(strided_gemm_kernel<float, 8, 8>, 128, 128);
  } else {
    if (small) launch(strided_gemm_kernel<double, 4, 4>, 64, 64);
    else launch(strided_gemm_kernel<double, 8, 4>, 128, 64);
  }
  B2OT_CHECK_LAUNCH();
  return 0;
}
