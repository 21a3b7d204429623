// Micro-benchmark: issue rate of tcgen05.mma on B200 for the shapes the online-cost sweep could use.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o gpurun_out/ubench_umma scripts/ubench_umma.cu
// Every SM (or SM pair) issues NMMA back-to-back MMAs on (garbage) shared-memory operands in the canonical
// K-major SWIZZLE_128B layout and waits for the last commit; prints clocks per MMA and the TFLOP/s this implies.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t a) {
  return (uint64_t)((a & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ bool try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

// KIND: 0 = tf32 (K = 8 per MMA), 1 = f16 (K = 16).  CG: cta_group (1 or 2).  N: MMA N.
template <int KIND, int CG, int N>
__global__ void __launch_bounds__(128, 1) umma_rate(int nmma, unsigned long long* out_clk) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t holder;
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t crank = 0;
  if (CG == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  // zero operands (avoid NaN power effects; values do not matter for the rate)
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) ((float*)smem)[i] = 1.0f;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 0) {
    if (CG == 1) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&holder)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&holder)), "r"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (CG == 2) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = holder;
  constexpr uint32_t M = (CG == 2) ? 256 : 128;
  constexpr uint32_t FMT = (KIND == 0) ? 2u : 0u;  // tf32 : f16
  constexpr uint32_t IDESC = (1u << 4) | (FMT << 7) | (FMT << 10) | ((uint32_t)(N >> 3) << 17) | ((M >> 4) << 24);
  unsigned long long t0 = 0, t1 = 0;
  if (warp == 1 && crank == 0) {
    if (lane == 0) {
      const uint32_t a = smem_u32(smem), b = smem_u32(smem + 16384);
      t0 = clock64();
      for (int i = 0; i < nmma; ++i) {
        const uint32_t ks = (uint32_t)(i & 3) * 32;
        const uint32_t d = tmem + (uint32_t)((i >> 2) & 1) * (N <= 256 ? 256 : 0);
        const uint64_t ad = desc_sw128(a + ks), bd = desc_sw128(b + ks);
        const uint32_t acc = 1;
        if (KIND == 0 && CG == 1)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
        if (KIND == 1 && CG == 1)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
        if (KIND == 0 && CG == 2)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
        if (KIND == 1 && CG == 2)
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
      }
      if (CG == 1)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      else
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "h"((uint16_t)1) : "memory");
      while (!try_wait(&bar, 0)) {
      }
      t1 = clock64();
      out_clk[blockIdx.x] = t1 - t0;
    }
    __syncwarp();
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (CG == 2) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (CG == 1)
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
    else
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
  }
}

template <int KIND, int CG, int N>
static void run(const char* name, int nmma) {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int grid = (CG == 2) ? (sms / 2) * 2 : sms;
  unsigned long long* d;
  cudaMalloc(&d, grid * sizeof(unsigned long long));
  cudaMemset(d, 0, grid * sizeof(unsigned long long));
  const size_t smem = 1024 + 16384 + 32768;
  cudaFuncSetAttribute(umma_rate<KIND, CG, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CG;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    cudaError_t err = cudaLaunchKernelEx(&cfg, umma_rate<KIND, CG, N>, nmma, d);
    cudaEventRecord(e1);
    cudaError_t e2 = cudaDeviceSynchronize();
    if (err != cudaSuccess || e2 != cudaSuccess) {
      printf("%s: launch error %s / %s\n", name, cudaGetErrorString(err), cudaGetErrorString(e2));
      return;
    }
  }
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  unsigned long long h[256];
  cudaMemcpy(h, d, grid * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  unsigned long long mx = 0;
  for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
  const double M = (CG == 2) ? 256 : 128, K = (KIND == 0) ? 8 : 16;
  const double issuers = (CG == 2) ? grid / 2 : grid;
  const double flops = 2.0 * M * N * K * nmma * issuers;
  printf("%-28s %8.1f clk/MMA  %8.3f ms  %8.1f TFLOP/s (dense %s)\n", name, (double)mx / nmma, ms, flops / (ms * 1e-3) / 1e12,
         KIND == 0 ? "tf32" : "f16");
  cudaFree(d);
}

int main() {
  const int n = 1 << 15;
  run<0, 1, 128>("tf32 cg1 128x128x8", n);
  run<0, 1, 256>("tf32 cg1 128x256x8", n);
  run<1, 1, 128>("f16  cg1 128x128x16", n);
  run<1, 1, 256>("f16  cg1 128x256x16", n);
  run<0, 2, 128>("tf32 cg2 256x128x8", n);
  run<0, 2, 256>("tf32 cg2 256x256x8", n);
  run<1, 2, 256>("f16  cg2 256x256x16", n);
  return 0;
}
