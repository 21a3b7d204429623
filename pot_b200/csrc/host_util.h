// Host-side helpers shared by the b2ot translation units: thread-local error text,
// launch counter, CUDA error check macro.
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

namespace b2ot {

void set_error(const char* fmt, ...);
std::atomic<unsigned long long>& launch_counter();

inline void count_launch(int n = 1) { launch_counter().fetch_add((unsigned long long)n, std::memory_order_relaxed); }

#define B2OT_CUDA(call)                                                                          \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) {                                                                     \
      ::b2ot::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return -2;                                                                                 \
    }                                                                                            \
  } while (0)

#define B2OT_CHECK_LAUNCH()                                                                      \
  do {                                                                                           \
    cudaError_t e_ = cudaGetLastError();                                                         \
    if (e_ != cudaSuccess) {                                                                     \
      ::b2ot::set_error("%s:%d: kernel launch failed: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return -3;                                                                                 \
    }                                                                                            \
    ::b2ot::count_launch();                                                                      \
  } while (0)

#define B2OT_REQUIRE(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      ::b2ot::set_error(__VA_ARGS__);      \
      return -1;                           \
    }                                      \
  } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Small pinned host blocks (the mirrors of the device status words that polls copy into): cudaMallocHost /
// cudaFreeHost cost ~1 ms + ~0.5 ms per solve, so released blocks are kept on a free list (defined in dense.cu).
constexpr size_t kPinnedBlock = 1024;
void* pinned_acquire(size_t bytes);   // nullptr on failure; bytes <= kPinnedBlock
void pinned_release(void* p);

// Tuning / diagnosis switches (B2OT_STAGES, B2OT_CLUSTERS, B2OT_PDL, B2OT_EMP_DBG, ...) are read from the environment
// ONLY in a diagnosis build (`make DIAG=1`, -DB2OT_DIAG): the product library ignores the environment altogether.
inline const char* diag_env(const char* name) {
#ifdef B2OT_DIAG
  return getenv(name);
#else
  (void)name;
  return nullptr;
#endif
}

// Programmatic dependent launch of the per-iteration kernels (diagnosis build: B2OT_PDL=0 turns it off): the kernel may start while
// the previous kernel of the stream is still finishing; it blocks in griddepcontrol.wait before touching its results.
inline bool pdl_enabled() {
  static const bool on = [] {
    const char* e = diag_env("B2OT_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace b2ot
