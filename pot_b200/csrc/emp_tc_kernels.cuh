// Online-cost Sinkhorn sweep on the 5th-generation tensor cores (tcgen05 / TMEM), sm_100a, fp32 inputs.
//
// One half-iteration of the empirical solver is  out_i = LSE_j( pot_j - C_ij / reg )  with
// C_ij = |x_i|^2 + |y_j|^2 - 2 x_i.y_j.  The cross term is a GEMM (n x m x d); here it runs as 128 x 128 output tiles
// of tcgen05.mma with the accumulator in tensor memory:
//
//   * three-way operand split: every fp32 coordinate is split once per solve into (hi, lo) halves (scaled FP16 by
//     default, TF32 selectable) and a tile is D = Alo.Bhi + Ahi.Blo + Ahi.Bhi (fp32 accumulate), which keeps the dot
//     product at fp32-level accuracy (dropped term lo.lo ~ 2^-22).  The split operands are stored PRE-SWIZZLED in the
//     canonical K-major SWIZZLE_128B shared-memory layout, tile by tile, so that a plain 1-D TMA bulk copy
//     (cp.async.bulk) lands a tile in shared memory exactly as the UMMA descriptors expect it - no tensor maps needed.
//   * warp roles (480 threads, one CTA per SM): warps 0-11 = epilogue, three warpgroups taking column tiles round-robin
//     (three warps per SM sub-partition); warp 12 = producer (bulk copies: the row tile once, the column operand through
//     a 4-stage ring of 32 KB K chunks); warps 13, 14 = MMA issuers on alternate tiles (12 or 24 MMAs per tile;
//     tcgen05.commit releases the smem stage and publishes the accumulator).  The warp index is made provably
//     warp-uniform (shuffle), so ring stages, descriptors and barrier addresses live in uniform registers.
//   * epilogue: one thread owns one row of the tile: tcgen05.ld 32 lanes x 32 columns (double-buffered), then ONE pass
//     per chunk: t = 2 k2 x.y + B_j - k2 |x|^2 - m_i against a lagging per-row reference exponent m_i, side maximum,
//     2^t, sum - all on packed fp32x2 arithmetic (FADD2 / FFMA2 / FMNMX3; ~4 issue slots per element with the
//     MUFU).  The per-column vectors are prefetched one tile ahead and only touched when the next tile is staged.
//   * 4 accumulator stages (4 x 128 TMEM columns = all 512) decouple the MMA stream from the epilogue.
//   * clamp: the reference clamps the expanded cost at 0 (ot/utils.py:299).  The sweep does not (B2OT_TC_CLAMP=1
//     builds it back in): the clamp only moves entries whose COMPUTED cost is negative, i.e. true cost below the
//     rounding of the expansion itself, |C_ij| <~ 2^-21 (|x_i|^2 + |y_j|^2); leaving it out changes a potential by
//     at most that amount (reg cancels: d(pot) = reg ln2 * d(exponent) = dC), the same size as the rounding every
//     other entry carries, and takes 1.75 of 5 issue slots per element out of the loop (-12 % time, measured).
//     Coincident points (X is Y) are covered by tests/test_empirical_gpu.py::test_tc_coincident_points.
// The n x m cost matrix never exists; HBM sees only the packed point clouds (L2 resident) and the vectors.
// Bound: MUFU (n m ex2, 16 / clk / SM) with the tensor pipe (3 x 2 n m d flop, shared-memory operand reads) close
// behind; measured history and the variants that did not pay: profiles/r02_emp_experiments.md.
#pragma once

#include <cuda_fp16.h>

#include "common.cuh"

namespace b2ot {

constexpr int TC_BM = 128;        // rows of an output tile (TMEM lanes)
constexpr int TC_BN = 128;        // columns of an output tile
constexpr int TC_KCH = 32;        // K elements per swizzle chunk (32 x 4 B = 128 B rows)
constexpr int TC_SUB = 16384;     // bytes of one (chunk, hi|lo) sub-tile: 128 rows x 128 B
constexpr int TC_NB = 4;          // shared-memory stages of the column operand, one K chunk (hi + lo) each
constexpr int TC_STAGE = 2 * TC_SUB;  // bytes of a stage
// Build-time switch (A/B'd as separate builds of the library, scripts/build_variants.sh):
#ifndef B2OT_TC_CLAMP
#define B2OT_TC_CLAMP 0   // 1: clamp the expanded cost at 0 inside the sweep (see the header comment)
#endif
constexpr int TC_NACC = 4;        // accumulator stages in tensor memory
#ifndef B2OT_TC_EPI_WARPS
#define B2OT_TC_EPI_WARPS 12
#endif
constexpr int TC_EPI_WARPS = B2OT_TC_EPI_WARPS;  // epilogue warpgroups (4 warps each) take column tiles round-robin
constexpr int TC_NWG = TC_EPI_WARPS / 4;
// Warp roles.  The SM sub-partition arbiter prefers the highest warp id among eligible warps, so the three service
// warps (producer, two MMA issuers: few instructions, but every one of them on the tile hand-off chain) take the TOP ids
// (measured neutral against ids 0, 1 and 2 + TC_EPI_WARPS, profiles/r02_emp_experiments.md).
constexpr int TC_EPI_WARP0 = 0;                     // first epilogue warp (multiple of 4: warp & 3 = TMEM lane quarter)
constexpr int TC_PROD_WARP = TC_EPI_WARPS;
constexpr int TC_MMA1_WARP = TC_EPI_WARPS + 1;
constexpr int TC_MMA2_WARP = TC_EPI_WARPS + 2;      // second MMA-issuing warp (odd tiles)
constexpr int TC_THREADS = 96 + 32 * TC_EPI_WARPS;  // producer warp, MMA warp, epilogue warps, second MMA warp
constexpr int TC_MAX_D = 64;
constexpr int TC_NSLOT = 2 * TC_NWG;  // column-vector slots: two per warpgroup
constexpr int TC_VEC = TC_BN;         // columns a slot serves
constexpr int TC_SLOT = 2 * TC_VEC;   // floats of a slot: B_j and G_j

__host__ __device__ inline size_t tc_tile_bytes(int kc) { return (size_t)kc * 2 * TC_SUB; }
__host__ __device__ inline size_t tc_pack_bytes(long long rows, int kc) {
  return (size_t)((rows + TC_BM - 1) / TC_BM) * tc_tile_bytes(kc);
}
inline size_t tc_smem_bytes(int kc) {
  return 1024 /*alignment slack*/ + tc_tile_bytes(kc) + (size_t)TC_NB * TC_STAGE + (size_t)TC_NSLOT * TC_SLOT * sizeof(float) /*column vectors: two slots per warpgroup*/ +
         (size_t)(TC_NWG - 1) * 2 * TC_BM * sizeof(float) /*warpgroup merge*/;
}

__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// ---- common centre of the two clouds ----
// C_ij = |x_i - y_j|^2 does not change when both clouds are shifted by the same vector, but its expansion
// |x|^2 + |y|^2 - 2 x.y cancels log2(|x|^2 / C) bits, and the (hi, lo) operand split carries 22 significant bits, two
// fewer than fp32: for a cloud 30 standard deviations off the origin the 3xFP16 sweep missed the 1e-4 parity target
// (1.3e-4, tests/test_empirical_gpu.py::test_tc_wide_dynamic_range[offcentre]).  The tensor-core path therefore works
// on centred coordinates: c = mean of the TARGET cloud (replicated on every rank of a sharded solve, so all ranks
// use the same centre), subtracted in fp32 before the norms, the scale search and the split.  Deterministic: per-block
// partial sums in double, combined in a fixed order.
constexpr int TC_CENTRE_BLOCKS = 128;
__global__ void __launch_bounds__(256) emp_centre_partial_kernel(const float* __restrict__ X, long long n, int d,
                                                                 long long ld, double* __restrict__ partial) {
  __shared__ double red[4][TC_MAX_D];
  const int k = threadIdx.x & 63, sub = threadIdx.x >> 6;
  double s = 0.0;
  if (k < d)
    for (long long i = (long long)blockIdx.x * 4 + sub; i < n; i += (long long)gridDim.x * 4) s += (double)X[i * ld + k];
  red[sub][k] = s;
  __syncthreads();
  if (sub == 0) partial[(size_t)blockIdx.x * TC_MAX_D + k] = (red[0][k] + red[1][k]) + (red[2][k] + red[3][k]);
}
__global__ void emp_centre_final_kernel(const double* __restrict__ partial, int nblocks, long long n, int d,
                                        float* __restrict__ centre) {
  const int k = threadIdx.x;
  if (k >= TC_MAX_D) return;
  double s = 0.0;
  for (int b = 0; b < nblocks; ++b) s += partial[(size_t)b * TC_MAX_D + k];
  const float c = (k < d && n > 0) ? (float)(s / (double)n) : 0.f;
  centre[k] = (fabsf(c) < CUDART_INF_F) ? c : 0.f;  // nan / inf coordinates: no shift (the solve reports them itself)
}
// squared norms of the centred rows (one warp per row)
__global__ void __launch_bounds__(256) emp_sqnorm_centred_kernel(const float* __restrict__ X, long long n, int d,
                                                                 long long ld, const float* __restrict__ centre,
                                                                 float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= n) return;
  float s = 0.f;
  for (int k = lane; k < d; k += 32) {
    const float x = X[i * ld + k] - centre[k];
    s = fmaf(x, x, s);
  }
  s = warp_sum(s);
  if (lane == 0) out[i] = s;
}

// Split + pre-swizzle: packed[tile][chunk c][hi|lo][group g of 8 rows][row r][16-byte unit (j ^ r)][4 floats].
// One thread per (row, 16-byte unit).
__global__ void __launch_bounds__(256) emp_tc_pack_kernel(const float* __restrict__ X, long long n, int d, long long ld,
                                                          int kc, const float* __restrict__ centre,
                                                          float* __restrict__ packed) {
  const long long rows_pad = (n + TC_BM - 1) / TC_BM * TC_BM;
  const int units = kc * 8;  // 16-byte units per row
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= rows_pad * units) return;
  const long long row = idx / units;
  const int u = (int)(idx % units);
  const int c = u >> 3, j = u & 7;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row < n) {
    const int k0 = c * TC_KCH + j * 4;
    const float* src = X + row * ld + k0;
    if (k0 + 0 < d) v.x = src[0] - centre[k0 + 0];
    if (k0 + 1 < d) v.y = src[1] - centre[k0 + 1];
    if (k0 + 2 < d) v.z = src[2] - centre[k0 + 2];
    if (k0 + 3 < d) v.w = src[3] - centre[k0 + 3];
  }
  float4 hi, lo;
  hi.x = to_tf32(v.x);  lo.x = to_tf32(v.x - hi.x);
  hi.y = to_tf32(v.y);  lo.y = to_tf32(v.y - hi.y);
  hi.z = to_tf32(v.z);  lo.z = to_tf32(v.z - hi.z);
  hi.w = to_tf32(v.w);  lo.w = to_tf32(v.w - hi.w);
  const long long tile = row / TC_BM;
  const int rt = (int)(row % TC_BM);
  const int g = rt >> 3, r = rt & 7;
  char* base = reinterpret_cast<char*>(packed) + tile * tc_tile_bytes(kc) + (size_t)(c * 2) * TC_SUB + g * 1024 + r * 128 +
               ((j ^ r) << 4);
  *reinterpret_cast<float4*>(base) = hi;
  *reinterpret_cast<float4*>(base + TC_SUB) = lo;
}

// ---- 3xFP16 operands (kind::f16): same split idea on half the bytes ----
// x * s = hi + lo with hi = fp16(x s), lo = fp16(x s - hi): 11 + 11 significant bits, the same as the TF32 split.
// s is ONE power of two per cloud chosen so that max |x s| lies in [2^13, 2^14): the small halves then stay normal
// fp16 numbers unless |x| < 2^-15 max|x| (absolute error 2^-39 of the largest coordinate).  A 128-byte swizzle row
// holds 64 fp16, so a K chunk covers 64 coordinates and one 128x128x16 MMA does twice the work of a 128x128x8 TF32
// one for the same 8 KB of shared-memory operand traffic -- which is what bounds this kernel (UMMA operand reads +
// TMA fills share the 128 B/clk of shared-memory bandwidth).
__global__ void __launch_bounds__(256) emp_absmax_kernel(const float* __restrict__ X, long long n, int d, long long ld,
                                                         const float* __restrict__ centre,
                                                         unsigned int* __restrict__ out_bits) {
  float mx = 0.f;
  const long long total = n * d;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const float v = fabsf(X[(i / d) * ld + (i % d)] - centre[i % d]);
    if (v < CUDART_INF_F) mx = fmaxf(mx, v);  // nan / inf coordinates poison the result anyway
  }
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(mx));
}
// scales[0] = s_x, scales[1] = s_y, scales[2] = 1 / (s_x s_y)   (all powers of two)
__global__ void emp_tc_scale_kernel(const unsigned int* __restrict__ maxbits, float* __restrict__ scales) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float s[2];
  for (int i = 0; i < 2; ++i) {
    const float mx = __uint_as_float(maxbits[i]);
    int e = 0;
    if (mx > 0.f) frexpf(mx, &e);  // mx = f 2^e, f in [0.5, 1)
    s[i] = (mx > 0.f) ? ldexpf(1.f, 14 - e) : 1.f;
  }
  scales[0] = s[0];
  scales[1] = s[1];
  scales[2] = 1.f / (s[0] * s[1]);
}
// packed[tile][chunk c of 64 coordinates][hi|lo][group g of 8 rows][row r][16-byte unit (j ^ r)][8 halves]
__global__ void __launch_bounds__(256) emp_tc_pack_f16_kernel(const float* __restrict__ X, long long n, int d, long long ld,
                                                              int kc, const float* __restrict__ centre,
                                                              const float* __restrict__ scale,
                                                              unsigned char* __restrict__ packed) {
  const long long rows_pad = (n + TC_BM - 1) / TC_BM * TC_BM;
  const int units = kc * 8;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= rows_pad * units) return;
  const long long row = idx / units;
  const int u = (int)(idx % units);
  const int c = u >> 3, j = u & 7;
  const float sc = *scale;
  __half hi[8], lo[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = c * 64 + j * 8 + e;
    const float v = (row < n && k < d) ? (X[row * ld + k] - centre[k]) * sc : 0.f;
    hi[e] = __float2half_rn(v);
    lo[e] = __float2half_rn(v - __half2float(hi[e]));
  }
  const long long tile = row / TC_BM;
  const int rt = (int)(row % TC_BM);
  const int g = rt >> 3, r = rt & 7;
  unsigned char* base = packed + tile * tc_tile_bytes(kc) + (size_t)(c * 2) * TC_SUB + g * 1024 + r * 128 + ((j ^ r) << 4);
  *reinterpret_cast<uint4*>(base) = *reinterpret_cast<const uint4*>(hi);
  *reinterpret_cast<uint4*>(base + TC_SUB) = *reinterpret_cast<const uint4*>(lo);
}

// ---- tcgen05 PTX wrappers ----
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_holder, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_holder)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// one lane of the (converged) warp; the compiler recognises the elected lane and emits the uniform-datapath
// tcgen05 instructions inside `if (elect_one())` once, instead of a per-active-thread election loop
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xFFFFFFFF;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "elect.sync _|q, 0xFFFFFFFF;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "elect.sync _|q, 0xFFFFFFFF;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// NOTE: the tcgen05.mma / tcgen05.commit wrappers are WARP-level calls: every lane of the converged MMA warp executes
// them with identical (warp-uniform) operands and one lane is elected inside the asm block.  ptxas then keeps the
// descriptors in uniform registers and emits back-to-back UTCHMMA; issuing from `if (lane == 0)` instead costs an
// election loop + R2UR moves per MMA (~85 clk per MMA measured, more than the 64 clk the MMA itself takes).
// arrive on an mbarrier when all MMAs issued so far by this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xFFFFFFFF;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
// same, signalling the barrier at this offset in every CTA of the cluster named by `mask`
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "elect.sync _|q, 0xFFFFFFFF;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// bulk copy global -> the same shared-memory offset of every CTA in `mask`; each destination CTA's mbarrier (same
// offset) receives the byte count
__device__ __forceinline__ void bulk_g2s_mc(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}
// K-major, SWIZZLE_128B operand descriptor: 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
constexpr uint32_t TC_IDESC_F16 = (1u << 4) /*D = f32, A = B = f16*/ | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
constexpr uint32_t TC_IDESC = (1u << 4) /*D = f32*/ | (2u << 7) /*A = tf32*/ | (2u << 10) /*B = tf32*/ |
                              ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

#define B2OT_TMEM_LD32(taddr, v)                                                                                       \
  asm volatile(                                                                                                        \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                        \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                        \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                        \
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),    \
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),         \
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),        \
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                      \
      : "r"(taddr)                                                                                                     \
      : "memory")
// wait for the outstanding tcgen05.ld of this thread; the registers are threaded through as in/out operands so that
// the compiler cannot schedule a use of them above the wait
#define B2OT_TMEM_WAIT32(v)                                                                                            \
  asm volatile("tcgen05.wait::ld.sync.aligned;"                                                                        \
               : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),       \
                 "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), \
                 "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]),            \
                 "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]),            \
                 "+r"(v[30]), "+r"(v[31])                                                                              \
               :                                                                                                       \
               : "memory")
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// shared-memory read the compiler may neither hoist nor merge (keeps its place between the other volatile statements)
__device__ __forceinline__ float4 lds128_volatile(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}

// mbarrier wait with a watchdog: a pipeline bug must end in a trap (launch error), never in a hung GPU
__device__ __forceinline__ void mbar_wait_guard(uint64_t* bar, uint32_t parity) {
  unsigned long long t0 = 0;
  for (uint32_t spins = 0;; ++spins) {
    if (mbar_try_wait(bar, parity)) return;
    if ((spins & 0xFFFu) == 0xFFFu) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000ull) __trap();
    }
  }
}

// Pipeline trace and experiment switches exist only in the diagnosis build (-DB2OT_DIAG, `make DIAG=1`); in the
// product build both macros fold to nothing and the kernel carries no such branches.
#ifdef B2OT_DIAG
#define B2OT_TRACE(slot, t)                                                                     \
  do {                                                                                          \
    if (p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && (t) < 256)                  \
      p.trace[(t) * 8 + (slot)] = clock64();                                                    \
  } while (0)
#define TC_DBG(bit) (p.dbg & (bit))
#else
#define B2OT_TRACE(slot, t) do { } while (0)
#define TC_DBG(bit) false
#endif

struct EmpTcParams {
  const float* Apack;   // row operand, packed tiles
  const float* Bpack;   // column operand, packed tiles
  const float* xx;      // |x_i|^2 (nx)
  const float* yy;      // |y_j|^2 (ny)
  const float* poty;    // potential along the reduced axis (ny)
  const int* pot_sel;   // optional ring slot selector (see EmpTileParams)
  long long pot_stride;
  float* pmax;          // [nsplit][nx]
  float* psum;
  long long nx, ny;
  int kc;               // K chunks of 32 (1 or 2)
  int nsplit;
  float k2;             // log2(e) / reg
  const int* status;
  int expect_status;
  int kind;             // 0: 3xTF32 (kind::tf32, 32 coordinates per K chunk), 1: 3xFP16 (kind::f16, 64 per chunk)
  const float* dotscale; // 3xFP16: accumulator -> x.y factor 1 / (s_x s_y), device scalar; null = 1
  long long* trace;     // experiment: clock64 stamps of CTA (0,0): [tile][8] (null = off)
  int dbg;              // experiment switches (B2OT_EMP_DBG): 1 no MUFU, 2 no epilogue arithmetic, 4 no MMAs
};

__global__ void __launch_bounds__(TC_THREADS, 1) emp_tc_lse_kernel(const EmpTcParams p) {
  extern __shared__ unsigned char tc_smem_raw[];
  __shared__ __align__(8) uint64_t a_full;
  __shared__ __align__(8) uint64_t b_full[TC_NB], b_empty[TC_NB];
  __shared__ __align__(8) uint64_t acc_full[TC_NACC], acc_empty[TC_NACC];
  __shared__ uint32_t tmem_holder;

  if (p.status != nullptr && *p.status != p.expect_status) return;

  // the warp index goes through a shuffle so that the compiler knows it is warp-uniform: everything derived from the
  // role (tile counters, ring stages, UMMA descriptors, barrier addresses) can then live in uniform registers - the
  // MMA issue loop otherwise builds each descriptor in vector registers and moves it over with R2UR (~20 instructions
  // of uniform-datapath latency per MMA on the tile hand-off chain)
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const size_t TB = tc_tile_bytes(p.kc);
  // 1024-byte aligned operand area (SWIZZLE_128B atoms repeat every 1024 B)
  // (offset arithmetic on the shared array itself, so the compiler keeps the shared address space: LDS, not LD)
  unsigned char* smem = tc_smem_raw + ((1024u - (smem_u32(tc_smem_raw) & 1023u)) & 1023u);
  unsigned char* As = smem;
  unsigned char* Bs = smem + TB;
  float* colvec = reinterpret_cast<float*>(smem + TB + (size_t)TC_NB * TC_STAGE);  // [TC_NWG][2 slots][2][TC_BN]: pot * log2e, |y|^2

  // column tiles of this block
  const long long ctiles = (p.ny + TC_BN - 1) / TC_BN;
  const long long per = (ctiles + p.nsplit - 1) / p.nsplit;
  const long long ct0 = (long long)blockIdx.y * per;
  const long long ct1 = (ct0 + per < ctiles) ? ct0 + per : ctiles;
  const long long ntiles = ct1 > ct0 ? ct1 - ct0 : 0;
  const long long row_tile = blockIdx.x;
  // cluster of CS CTAs along the row tiles: same column tiles, so each CTA fetches 1/CS of every column tile and
  // multicasts it to all (L2 -> SM traffic of the column operand / CS)
  const uint32_t cs = cluster_nctarank(), crank = cluster_ctarank();
  const uint16_t cmask = (uint16_t)((1u << cs) - 1u);
  const long long row_tiles = (p.nx + TC_BM - 1) / TC_BM;
  const long long row_tile_ld = row_tile < row_tiles ? row_tile : row_tiles - 1;  // padding CTA of the last cluster

  if (tid == 0) {
    mbar_init(&a_full, 1);
    for (int s = 0; s < TC_NB; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], cs);
    }
    for (int s = 0; s < TC_NACC; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 128);
    }
    fence_mbar_init();
  }
  if (warp == TC_MMA1_WARP) tmem_alloc(&tmem_holder, 512);
  tc_fence_before();
  if (cs > 1)
    cluster_sync_all();  // every CTA's barriers exist before a peer's multicast copy or commit can reach them
  else
    __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_holder;

  if (warp == TC_PROD_WARP) {
    // ===== producer: bulk copies of pre-swizzled operands; the column operand moves through the ring one
    // K chunk (hi + lo sub-tiles, 32 KB) at a time =====
    if (lane == 0 && ntiles > 0) {
      const unsigned char* Ag = reinterpret_cast<const unsigned char*>(p.Apack) + (size_t)row_tile_ld * TB;
      mbar_arrive_expect_tx(&a_full, (uint32_t)TB);
      for (size_t o = 0; o < TB; o += TC_SUB) bulk_g2s(As + o, Ag + o, TC_SUB, &a_full);
      long long qn = 0;  // chunk counter
      for (long long t = 0; t < ntiles; ++t) {
        const unsigned char* Bg = reinterpret_cast<const unsigned char*>(p.Bpack) + (size_t)(ct0 + t) * TB;
        for (int c = 0; c < p.kc; ++c, ++qn) {
          const int s = (int)(qn % TC_NB);
          const uint32_t ph = (uint32_t)((qn / TC_NB) & 1);
          mbar_wait_guard(&b_empty[s], ph ^ 1u);
          B2OT_TRACE(0, t);
          if (TC_DBG(8) && qn >= TC_NB) {  // experiment: no refills (stale operands), handshake only
            mbar_arrive(&b_full[s]);
            continue;
          }
          mbar_arrive_expect_tx(&b_full[s], (uint32_t)TC_STAGE);
          unsigned char* dst = Bs + (size_t)s * TC_STAGE;
          const unsigned char* src = Bg + (size_t)c * TC_STAGE;
          if (cs == 1) {
            bulk_g2s(dst, src, TC_SUB, &b_full[s]);
            bulk_g2s(dst + TC_SUB, src + TC_SUB, TC_SUB, &b_full[s]);
          } else {
            const size_t piece = (size_t)TC_STAGE / cs, base = (size_t)crank * piece;
            for (size_t o = 0; o < piece; o += TC_SUB) {
              const size_t len = (piece - o < (size_t)TC_SUB) ? piece - o : (size_t)TC_SUB;
              bulk_g2s_mc(dst + base + o, src + base + o, (uint32_t)len, &b_full[s], cmask);
            }
          }
        }
      }
    }
  } else if (warp == TC_MMA1_WARP || warp == TC_MMA2_WARP) {
    // ===== MMA issuers: two warps take alternate column tiles.  A tcgen05.mma issue blocks while the tensor pipe's
    // queue is full, so one warp's barrier waits for its next tile overlap the other warp's issue stream =====
    const int mw = (warp == TC_MMA1_WARP) ? 0 : 1;
    if (ntiles > 0) {
      mbar_wait_guard(&a_full, 0);
      const uint32_t a_addr = smem_u32(As);
      // The CTA owns all 512 TMEM columns, so the allocation starts at column 0 / lane 0: using the literal 0 keeps
      // the accumulator address in a uniform register (an address loaded from shared memory would force a
      // per-thread register + election loop around every MMA).
      if (tmem_base != 0u) __trap();
      const uint32_t tmem_u = 0u;
      for (long long t = mw; t < ntiles; t += 2) {
        long long qn = t * p.kc;  // chunk counter of this tile's first K chunk
        const int as = (int)(t % TC_NACC);
        const uint32_t aph = (uint32_t)((t / TC_NACC) & 1);
        mbar_wait_guard(&acc_empty[as], aph ^ 1u);
        if (lane == 0) B2OT_TRACE(1, t);
        const uint32_t d_tmem = tmem_u + (uint32_t)(as * TC_BN);
        uint32_t acc = 0;
        for (int c = 0; c < p.kc; ++c, ++qn) {
          const int s = (int)(qn % TC_NB);
          const uint32_t bph = (uint32_t)((qn / TC_NB) & 1);
          mbar_wait_guard(&b_full[s], bph);
          tc_fence_after();
          if (lane == 0) B2OT_TRACE(2, t);
          __syncwarp();  // converged warp: the uniform-datapath instructions below are emitted without election loops
          {
            const uint32_t b_addr = smem_u32(Bs + (size_t)s * TC_STAGE);
            const uint32_t a_chunk = a_addr + (uint32_t)(c * TC_STAGE);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              const uint32_t off_hi = (uint32_t)(ks * 32);
              const uint32_t off_lo = off_hi + TC_SUB;
              const uint64_t ah = umma_desc_sw128(a_chunk + off_hi), al = umma_desc_sw128(a_chunk + off_lo);
              const uint64_t bh = umma_desc_sw128(b_addr + off_hi), bl = umma_desc_sw128(b_addr + off_lo);
              if (TC_DBG(4)) continue;
              if (p.kind == 0) {
                umma_tf32(d_tmem, al, bh, TC_IDESC, acc);  // small terms first
                umma_tf32(d_tmem, ah, bl, TC_IDESC, 1u);
                umma_tf32(d_tmem, ah, bh, TC_IDESC, 1u);
              } else {
                umma_f16(d_tmem, al, bh, TC_IDESC_F16, acc);
                umma_f16(d_tmem, ah, bl, TC_IDESC_F16, 1u);
                umma_f16(d_tmem, ah, bh, TC_IDESC_F16, 1u);
              }
              acc = 1u;
            }
            if (cs == 1)
              umma_commit(&b_empty[s]);     // smem stage free once these MMAs have read it
            else
              umma_commit_mc(&b_empty[s], cmask);  // ... in every CTA of the cluster (they all write into it)
            if (c == p.kc - 1) umma_commit(&acc_full[as]);   // accumulator complete
            if (lane == 0) B2OT_TRACE(3, t);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ===== epilogue: TC_NWG warpgroups take column tiles round-robin; one thread = one row of a tile =====
    const int ew = warp - TC_EPI_WARP0;    // 0..TC_EPI_WARPS-1
    const int wg = ew >> 2;                // warpgroup
    const int q = warp & 3;                // TMEM lane quarter this warp may access
    const int et = (ew & 3) * 32 + lane;   // column of the tile this thread stages into shared memory
    const int rloc = q * 32 + lane;        // row of the tile this thread owns
    const long long row = row_tile * TC_BM + rloc;
    const float xi = (row < p.nx) ? p.xx[row] : 0.f;
    const float* potg = p.poty;
    if (p.pot_sel != nullptr) potg += (long long)(*p.pot_sel) * p.pot_stride;
    const float LOG2E = 1.4426950408889634f;
    const float k2 = p.k2, k22 = 2.f * p.k2 * (p.dotscale != nullptr ? *p.dotscale : 1.f);
    const float nxk = -k2 * xi;
    float m_run = -CUDART_INF_F, s_run = 0.f;
    // column vectors of a tile are fetched one tile ahead into registers (global-load latency hidden behind a
    // whole tile of work) and staged into shared memory as  B_j = pot_j log2e - k2 |y_j|^2  (and G_j = pot_j log2e for the clamp build)
    // (the loaded values are only touched when the next tile is staged: scaling the potential right here put the
    // global-load latency on the critical path of every tile - 4.4 % of the samples sat on that FMUL and the branch
    // behind it, profiles/r02j_ncu_emp_tc_v2.txt)
    float pg = -CUDART_INF_F, py = 0.f;
    auto fetch = [&](long long t) {
      const long long j = (ct0 + t) * TC_BN + et;
      const bool ok = (t < ntiles) && (j < p.ny);
      pg = ok ? potg[j] : -CUDART_INF_F;
      py = ok ? p.yy[j] : 0.f;
    };
    fetch(wg);
    float* cB = colvec;  // column-vector slot in use (set per tile)
    // One pass per chunk against a LAGGING reference exponent: m_run is a value this row
    // attained in an earlier chunk (or 0 when the first values lie within 2^+-64 of it), so t = r - m_run needs no
    // chunk maximum before the exponentials: cost, 2^t and the sums are independent per element - the
    // instruction-level parallelism the two-phase form lacks (there every pipe sat near 50 % and the rest was
    // dependency latency, profiles/r01d_emp_limiters.log).  The chunk maximum of t is tracked on the side: a chunk
    // that rises more than 64 above the reference (2^t could overflow), or a first chunk far from 0, moves the
    // reference to its own maximum, rescales the running sum and is redone.  Terms more than 126 below the reference
    // flush to zero: they are below 2^-62 of a term that is present in the sum.
    auto chunk = [&](uint32_t (&v)[32], int ch) {
      const uint32_t cb_addr = smem_u32(cB) + (uint32_t)(ch * 32 * sizeof(float));
      for (;;) {
        const bool unset = (m_run == -CUDART_INF_F);
        const float base = unset ? 0.f : m_run;
        const float c1 = nxk - base;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, x0 = -CUDART_INF_F, x1 = -CUDART_INF_F;
        // the column vectors are read two groups ahead with volatile loads: the compiler otherwise hoists all 16
        // LDS.128 of a chunk (and keeps them live across the redo loop), 64 registers that spill
        float4 bq[3];
        bq[0] = lds128_volatile(cb_addr);
        bq[1] = lds128_volatile(cb_addr + 16u);
#if B2OT_TC_CLAMP
        const float ng = -base;
        float4 gq[3];
        gq[0] = lds128_volatile(cb_addr + (uint32_t)(TC_VEC * sizeof(float)));
        gq[1] = lds128_volatile(cb_addr + (uint32_t)(TC_VEC * sizeof(float)) + 16u);
#endif
#pragma unroll
        for (int g4 = 0; g4 < 8; ++g4) {
          if (g4 + 2 < 8) {
            bq[(g4 + 2) % 3] = lds128_volatile(cb_addr + (uint32_t)((g4 + 2) * 16));
#if B2OT_TC_CLAMP
            gq[(g4 + 2) % 3] = lds128_volatile(cb_addr + (uint32_t)(TC_VEC * sizeof(float)) + (uint32_t)((g4 + 2) * 16));
#endif
          }
          float z0 = bq[g4 % 3].x, z1 = bq[g4 % 3].y, z2 = bq[g4 % 3].z, z3 = bq[g4 % 3].w;
          add2(z0, z1, c1, c1);
          add2(z2, z3, c1, c1);
          float t0, t1, t2, t3;
          fma2(t0, t1, __uint_as_float(v[g4 * 4 + 0]), __uint_as_float(v[g4 * 4 + 1]), k22, k22, z0, z1);
          fma2(t2, t3, __uint_as_float(v[g4 * 4 + 2]), __uint_as_float(v[g4 * 4 + 3]), k22, k22, z2, z3);
#if B2OT_TC_CLAMP
          float h0 = gq[g4 % 3].x, h1 = gq[g4 % 3].y, h2 = gq[g4 % 3].z, h3 = gq[g4 % 3].w;
          add2(h0, h1, ng, ng);
          add2(h2, h3, ng, ng);
          t0 = fminf(t0, h0);
          t1 = fminf(t1, h1);
          t2 = fminf(t2, h2);
          t3 = fminf(t3, h3);
#endif
          x0 = max3(x0, t0, t1);
          x1 = max3(x1, t2, t3);
          t0 = ex2_ftz(t0);
          t1 = ex2_ftz(t1);
          t2 = ex2_ftz(t2);
          t3 = ex2_ftz(t3);
          add2(s0, s1, t0, t1);
          add2(s2, s3, t2, t3);
        }
        const float cm = fmaxf(x0, x1);
        if (cm <= 64.f && (!unset || cm >= -64.f || cm == -CUDART_INF_F)) {
          s_run += (s0 + s1) + (s2 + s3);
          if (unset && cm != -CUDART_INF_F) m_run = 0.f;
          break;
        }
        s_run = unset ? 0.f : s_run * ex2_ftz(-cm);
        m_run = base + cm;
      }
    };
    int slot_sel = 0;
    for (long long t = wg; t < ntiles; t += TC_NWG) {
      const int as = (int)(t % TC_NACC);
      const uint32_t aph = (uint32_t)((t / TC_NACC) & 1);
      // column-vector slot: private to the warpgroup, double-buffered over its own tiles (a slot indexed by the
      // accumulator stage would be shared between warpgroups as soon as TC_NWG does not divide TC_NACC: the last
      // chunk of a tile is still read after acc_empty has been signalled)
      cB = colvec + (size_t)(wg * 2 + slot_sel) * TC_SLOT;
      slot_sel ^= 1;
      if (et == 0) B2OT_TRACE(4, t);
      const float Gj = pg * LOG2E;
      cB[et] = fmaf(-k2, py, Gj);
#if B2OT_TC_CLAMP
      cB[TC_VEC + et] = Gj;
#endif
      if (et == 0) B2OT_TRACE(5, t);
      fetch(t + TC_NWG);
      named_bar_sync(1 + wg, 128);
      mbar_wait_guard(&acc_full[as], aph);
      tc_fence_after();
      if (et == 0) B2OT_TRACE(6, t);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * TC_BN);
      if (TC_DBG(16)) {  // experiment: no TMEM reads
        tc_fence_before();
        mbar_arrive(&acc_empty[as]);
        continue;
      }
      uint32_t b0[32], b1[32];
      B2OT_TMEM_LD32(taddr, b0);
      B2OT_TMEM_WAIT32(b0);
      B2OT_TMEM_LD32(taddr + 32u, b1);
      chunk(b0, 0);
      B2OT_TMEM_WAIT32(b1);
      B2OT_TMEM_LD32(taddr + 64u, b0);
      chunk(b1, 1);
      B2OT_TMEM_WAIT32(b0);
      B2OT_TMEM_LD32(taddr + 96u, b1);
      chunk(b0, 2);
      B2OT_TMEM_WAIT32(b1);
      tc_fence_before();
      mbar_arrive(&acc_empty[as]);  // accumulator stage drained: the MMA warp may overwrite it
      chunk(b1, 3);
      if (et == 0) B2OT_TRACE(7, t);
    }
    // merge the warpgroups' running (max, sum) of each row
    float* mrg = colvec + (size_t)TC_NSLOT * TC_SLOT;
    if (wg > 0) {
      mrg[(size_t)(wg - 1) * 2 * TC_BM + rloc] = m_run;
      mrg[(size_t)(wg - 1) * 2 * TC_BM + TC_BM + rloc] = s_run;
    }
    named_bar_sync(1 + TC_NWG, 32 * TC_EPI_WARPS);
    if (wg == 0 && row < p.nx) {
      float M = m_run;
#pragma unroll
      for (int k = 0; k < TC_NWG - 1; ++k) M = fmaxf(M, mrg[(size_t)k * 2 * TC_BM + rloc]);
      const float mref = (M == -CUDART_INF_F) ? 0.f : M;
      float s = s_run * ex2_ftz(m_run - mref);
#pragma unroll
      for (int k = 0; k < TC_NWG - 1; ++k)
        s += mrg[(size_t)k * 2 * TC_BM + TC_BM + rloc] * ex2_ftz(mrg[(size_t)k * 2 * TC_BM + rloc] - mref);
      const float LN2 = 0.6931471805599453f;
      p.pmax[(long long)blockIdx.y * p.nx + row] = M * LN2;   // back to natural-log units
      p.psum[(long long)blockIdx.y * p.nx + row] = s;
    }
  }

  // ---- teardown ----
  tc_fence_before();
  if (cs > 1)
    cluster_sync_all();  // no CTA leaves while a peer can still multicast into its shared memory
  else
    __syncthreads();
  if (warp == TC_MMA1_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace b2ot
