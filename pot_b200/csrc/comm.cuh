// One-shot all-reduce over NVLink peer memory for the row-sharded solvers (sm_100a, one process per GPU).
//
// Every rank owns one "exchange buffer" of identical layout, mapped into every peer with CUDA IPC:
//
//   [0, 1 KiB)    flags  : u64 flag[parity][source rank]  -- the sequence number of the last complete put
//   [1 KiB, 4 KiB) local scratch (block ticket)
//   [4 KiB, ...)  data   : slot[parity][source rank][slot_bytes]
//
// An exchange with sequence number s (parity s & 1):
//   put  : each rank reduces its local partial rows and STORES the result straight into slot[s&1][rank] of
//          every peer's buffer (NVLink writes, fire and forget); the last block fences at system scope and
//          then publishes flag[s&1][rank] = s on every peer (st.release.sys).
//   take : the consumer kernel (the solver's finalize kernel itself on the hot path) spins on its LOCAL flags
//          until all ranks have published s (ld.acquire.sys), then sums the nranks slots in rank order -- the
//          same order on every rank, so all ranks see bit-identical totals and take identical stop decisions.
// Two parities suffice: a rank can only start the put of s + 2 after it consumed s + 1, which needs every
// peer's put of s + 1, which those peers issue after they consumed s.  The argument needs EVERY sequence number to be
// put and waited for by every rank: a solver that has stopped still publishes the flag (without data) and its
// consumer still waits for the peers' flags (comm_put_kernel / finalize_kernel), so the chain has no gaps.
// The per-iteration traffic is one m-vector per peer (256 KiB at m = 65536): latency-bound, no NCCL launch,
// no separate reduce kernel on the hot path.
#pragma once

#include "common.cuh"

namespace b2ot {

constexpr int COMM_MAX_RANKS = 8;
constexpr size_t COMM_DATA_OFFSET = 4096;
constexpr int ST_COMM_TIMEOUT = 5;  // solver status: a peer never published its flag

struct CommView {
  unsigned long long* flags[COMM_MAX_RANKS];  // flag area of rank q's buffer
  unsigned char* data[COMM_MAX_RANKS];        // data area of rank q's buffer
  unsigned int* ticket;                       // local
  int nranks, rank;
  unsigned long long slot_bytes;
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

template <typename T>
__device__ __forceinline__ T* comm_slot(const CommView& c, int owner, int parity, int src_rank) {
  return reinterpret_cast<T*>(c.data[owner] + (size_t)(parity * c.nranks + src_rank) * c.slot_bytes);
}

// Block-wide wait until every rank has published sequence number `seq`.  Returns false on timeout (20 s).
__device__ __forceinline__ bool comm_wait(const CommView& c, unsigned long long seq) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if ((int)threadIdx.x < c.nranks) {
    const unsigned long long* f = c.flags[c.rank] + (seq & 1ull) * COMM_MAX_RANKS + threadIdx.x;
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys(f) < seq) {
      if (global_timer_ns() - t0 > 20000000000ull) {
        s_ok = 0;
        break;
      }
    }
  }
  __syncthreads();
  return s_ok != 0;
}

// ---- tagged-word exchange (the dense solver's finalize kernel, dense_kernels.cuh) ----
// A value travels as 8-byte words {32 data bits, 32-bit sequence tag} (fp32: one word, fp64: two), stored straight into
// the peers' slots with single 8-byte stores (single-copy atomic, also over NVLink - the NCCL LL idea) and polled by the
// consumer until the tag equals the sequence number: no flag word, no system-scope fence, no block ticket on the hot path.
// The words live in their own region of a slot (offset ll_off), so the flag protocol above can share the buffer.
__device__ __forceinline__ void llw_store(unsigned long long* p, unsigned int bits, unsigned int tag) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(bits), "r"(tag) : "memory");
}
__device__ __forceinline__ bool llw_poll(const unsigned long long* p, unsigned int tag, unsigned int& bits) {
  unsigned int t;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(bits), "=r"(t) : "l"(p) : "memory");
  return t == tag;
}
template <typename T> struct LLWords;
template <> struct LLWords<float> {
  static constexpr int N = 1;
  static __device__ __forceinline__ void put(unsigned long long* base, long long i, float x, unsigned int tag) {
    llw_store(base + i, __float_as_uint(x), tag);
  }
  static __device__ __forceinline__ bool poll(const unsigned long long* base, long long i, unsigned int tag, float& x) {
    unsigned int b;
    const bool ok = llw_poll(base + i, tag, b);
    x = __uint_as_float(b);
    return ok;
  }
};
template <> struct LLWords<double> {
  static constexpr int N = 2;
  static __device__ __forceinline__ void put(unsigned long long* base, long long i, double x, unsigned int tag) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    llw_store(base + 2 * i, (unsigned int)b, tag);
    llw_store(base + 2 * i + 1, (unsigned int)(b >> 32), tag);
  }
  static __device__ __forceinline__ bool poll(const unsigned long long* base, long long i, unsigned int tag, double& x) {
    unsigned int lo, hi;
    const bool ok0 = llw_poll(base + 2 * i, tag, lo);
    const bool ok1 = llw_poll(base + 2 * i + 1, tag, hi);
    x = __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
    return ok0 && ok1;
  }
};
__device__ __forceinline__ unsigned long long* comm_ll_slot(const CommView& c, int owner, int parity, int src_rank,
                                                            unsigned long long ll_off) {
  return reinterpret_cast<unsigned long long*>(c.data[owner] + (size_t)(parity * c.nranks + src_rank) * c.slot_bytes + ll_off);
}
// value of element i from every rank, summed in rank order (identical on all ranks); false on timeout (20 s)
template <typename T>
__device__ __forceinline__ bool comm_ll_take(const CommView& c, int parity, unsigned long long ll_off, long long i,
                                             unsigned int tag, T& total) {
  T x[COMM_MAX_RANKS];
  unsigned int done = 0;
  const unsigned int all = (1u << c.nranks) - 1u;
  unsigned long long t0 = 0;
  for (unsigned int spins = 0; done != all; ++spins) {
#pragma unroll
    for (int q = 0; q < COMM_MAX_RANKS; ++q)
      if (q < c.nranks && !(done & (1u << q)) && LLWords<T>::poll(comm_ll_slot(c, c.rank, parity, q, ll_off), i, tag, x[q]))
        done |= 1u << q;
    if ((spins & 0x3FFFu) == 0x3FFFu) {
      const unsigned long long now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 20000000000ull) return false;
    }
  }
  total = x[0];
#pragma unroll
  for (int q = 1; q < COMM_MAX_RANKS; ++q)
    if (q < c.nranks) total += x[q];
  return true;
}

// put: dst_q[j] = sum over nsrc local rows of src[row * ld + j] for j < count; optional extra element at index
// `count` carrying (*flag_src != 0).  The DATA is written only while *status == expect_status (status may be null);
// the FLAG of `seq` is published in every case: the two-parity reuse argument in the header needs every sequence
// number to be exchanged by every rank, also after the solve has stopped (all ranks stop at the same iteration, but
// the host keeps enqueueing no-op iterations until its next poll; a gap in the sequence would let the first put
// after the gap land in a slot a slower peer is still reading).
// 16-byte path: one thread sums a 16-byte column vector over the local rows and stores it to every peer with one
// 16-byte NVLink store per peer (the scalar version issued 4-byte peer stores).
template <typename T>
__global__ void __launch_bounds__(256) comm_put_kernel(const CommView c, const T* __restrict__ src, long long ld, int nsrc,
                                                       long long count, const int* flag_src, const int* status,
                                                       int expect_status, unsigned long long seq) {
  using V = typename Vec<T>::type;
  constexpr int VN = Vec<T>::N;
  griddep_wait();
  griddep_launch_dependents();
  __shared__ int s_last;
  const bool live = (status == nullptr || *status == expect_status);
  const int par = (int)(seq & 1ull);
  if (live) {
    const bool vec_ok = (nsrc == 1 || ld % VN == 0) && (reinterpret_cast<uintptr_t>(src) % 16 == 0);
    const long long nvec = vec_ok ? count / VN : 0;
    for (long long jv = (long long)blockIdx.x * 256 + threadIdx.x; jv < nvec; jv += (long long)gridDim.x * 256) {
      V acc = __ldcg(reinterpret_cast<const V*>(src) + jv);
      for (int q = 1; q < nsrc; ++q) {
        const V x = __ldcg(reinterpret_cast<const V*>(src + (long long)q * ld) + jv);
#pragma unroll
        for (int e = 0; e < VN; ++e) elem(acc, e) += elem(x, e);
      }
      for (int q = 0; q < c.nranks; ++q) reinterpret_cast<V*>(comm_slot<T>(c, q, par, c.rank))[jv] = acc;
    }
    const long long total = count + (flag_src != nullptr ? 1 : 0);
    for (long long j = nvec * VN + (long long)blockIdx.x * 256 + threadIdx.x; j < total; j += (long long)gridDim.x * 256) {
      T s = (T)0;
      if (j < count) {
        for (int q = 0; q < nsrc; ++q) s += __ldcg(&src[(long long)q * ld + j]);
      } else {
        s = (T)(atomicOr(const_cast<int*>(flag_src), 0) ? 1 : 0);
      }
      for (int q = 0; q < c.nranks; ++q) comm_slot<T>(c, q, par, c.rank)[j] = s;
    }
    __threadfence_system();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(c.ticket, 1u);
    s_last = (t == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();
  if ((int)threadIdx.x < c.nranks)
    st_release_sys(c.flags[threadIdx.x] + (size_t)par * COMM_MAX_RANKS + c.rank, seq);
  if (threadIdx.x == 0) *c.ticket = 0;
}

// take (generic): dst[j] = reduce over ranks of slot[q][j], in place all-reduce result for the caller.
// op 0 = sum, 1 = max.  On timeout sets *err_status = ST_COMM_TIMEOUT (when given).
template <typename T>
__global__ void __launch_bounds__(256) comm_take_kernel(const CommView c, T* __restrict__ dst, long long count, int op,
                                                        const int* gate_status, int expect_status, int* err_status,
                                                        unsigned long long seq) {
  if (gate_status != nullptr && *gate_status != expect_status) return;
  if (!comm_wait(c, seq)) {
    if (err_status != nullptr && threadIdx.x == 0) *err_status = ST_COMM_TIMEOUT;
    return;
  }
  const int par = (int)(seq & 1ull);
  for (long long j = (long long)blockIdx.x * 256 + threadIdx.x; j < count; j += (long long)gridDim.x * 256) {
    T r = __ldcg(comm_slot<T>(c, c.rank, par, 0) + j);
    for (int q = 1; q < c.nranks; ++q) {
      const T x = __ldcg(comm_slot<T>(c, c.rank, par, q) + j);
      r = (op == 0) ? r + x : max_t(r, x);
    }
    dst[j] = r;
  }
}

}  // namespace b2ot
