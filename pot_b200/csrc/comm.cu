// Host side of the NVLink peer-memory exchange (see comm.cuh): buffer allocation + CUDA IPC plumbing, the
// communicator handle, and the generic two-kernel all-reduce built on it.
#include <algorithm>
#include <cstring>
#include <new>

#include "../../include/b2ot.h"
#include "comm.cuh"
#include "comm_host.h"
#include "host_util.h"

using namespace b2ot;

extern "C" int b2ot_comm_alloc(size_t bytes, void** buf_dev, unsigned char* ipc_handle) {
  B2OT_REQUIRE(buf_dev != nullptr && ipc_handle != nullptr, "b2ot_comm_alloc: null argument");
  B2OT_REQUIRE(bytes >= COMM_DATA_OFFSET + 1024, "b2ot_comm_alloc: buffer too small (%zu bytes)", bytes);
  static_assert(sizeof(cudaIpcMemHandle_t) == B2OT_IPC_HANDLE_BYTES, "IPC handle size");
  void* p = nullptr;
  B2OT_CUDA(cudaMalloc(&p, bytes));
  B2OT_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t hd;
  cudaError_t e = cudaIpcGetMemHandle(&hd, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("b2ot_comm_alloc: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    return -2;
  }
  memcpy(ipc_handle, &hd, sizeof(hd));
  B2OT_CUDA(cudaDeviceSynchronize());
  *buf_dev = p;
  return 0;
}

extern "C" int b2ot_comm_free(void* buf_dev) {
  if (buf_dev) B2OT_CUDA(cudaFree(buf_dev));
  return 0;
}

extern "C" int b2ot_comm_open(const unsigned char* ipc_handle, void** peer_ptr) {
  B2OT_REQUIRE(ipc_handle != nullptr && peer_ptr != nullptr, "b2ot_comm_open: null argument");
  cudaIpcMemHandle_t hd;
  memcpy(&hd, ipc_handle, sizeof(hd));
  B2OT_CUDA(cudaIpcOpenMemHandle(peer_ptr, hd, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int b2ot_comm_close(void* peer_ptr) {
  if (peer_ptr) B2OT_CUDA(cudaIpcCloseMemHandle(peer_ptr));
  return 0;
}

extern "C" int b2ot_comm_create(b2ot_comm** comm, int nranks, int rank, void* const* bufs, size_t bytes) {
  B2OT_REQUIRE(comm != nullptr && bufs != nullptr, "b2ot_comm_create: null argument");
  B2OT_REQUIRE(nranks >= 1 && nranks <= COMM_MAX_RANKS && rank >= 0 && rank < nranks,
               "b2ot_comm_create: bad rank %d / nranks %d (at most %d ranks)", rank, nranks, COMM_MAX_RANKS);
  B2OT_REQUIRE(bytes >= COMM_DATA_OFFSET + 1024, "b2ot_comm_create: buffer too small");
  b2ot_comm* c = new (std::nothrow) b2ot_comm();
  B2OT_REQUIRE(c != nullptr, "b2ot_comm_create: out of host memory");
  memset(c, 0, sizeof(*c));
  for (int q = 0; q < nranks; ++q) {
    if (bufs[q] == nullptr) {
      delete c;
      set_error("b2ot_comm_create: null buffer for rank %d", q);
      return -1;
    }
    c->view.flags[q] = (unsigned long long*)bufs[q];
    c->view.data[q] = (unsigned char*)bufs[q] + COMM_DATA_OFFSET;
  }
  c->view.ticket = (unsigned int*)((unsigned char*)bufs[rank] + 1024);
  c->view.nranks = nranks;
  c->view.rank = rank;
  c->view.slot_bytes = ((bytes - COMM_DATA_OFFSET) / (2 * (size_t)nranks)) / 256 * 256;
  c->seq = 0;
  *comm = c;
  return 0;
}

extern "C" int b2ot_comm_destroy(b2ot_comm* comm) {
  delete comm;
  return 0;
}

extern "C" size_t b2ot_comm_slot_bytes(const b2ot_comm* comm) { return comm ? (size_t)comm->view.slot_bytes : 0; }

namespace b2ot {

template <typename T>
static int allreduce_t(b2ot_comm* c, void* buf, long long count, int op, int* err_status, cudaStream_t s) {
  B2OT_REQUIRE((size_t)count * sizeof(T) <= c->view.slot_bytes, "peer all-reduce: %lld elements exceed the slot (%llu B)",
               count, c->view.slot_bytes);
  const unsigned long long seq = ++c->seq;
  const long long vn = 16 / (long long)sizeof(T);
  const unsigned grid = (unsigned)std::max<long long>(1, std::min<long long>((count / vn + 1 + 255) / 256, 148));
  comm_put_kernel<T><<<grid, 256, 0, s>>>(c->view, (const T*)buf, 0, 1, count, nullptr, nullptr, 0, seq);
  B2OT_CHECK_LAUNCH();
  comm_take_kernel<T><<<grid, 256, 0, s>>>(c->view, (T*)buf, count, op, nullptr, 0, err_status, seq);
  B2OT_CHECK_LAUNCH();
  return 0;
}

int comm_allreduce(b2ot_comm* c, void* buf, long long count, int dtype, int op, int* err_status, cudaStream_t s) {
  return dtype == B2OT_F32 ? allreduce_t<float>(c, buf, count, op, err_status, s)
                           : allreduce_t<double>(c, buf, count, op, err_status, s);
}

}  // namespace b2ot

// stand-alone entry point (tests, and callers that want the exchange without a solver)
extern "C" int b2ot_comm_allreduce(b2ot_comm* comm, void* buf_dev, int64_t count, int dtype, int op, void* stream) {
  B2OT_REQUIRE(comm != nullptr && buf_dev != nullptr && count >= 1, "b2ot_comm_allreduce: bad argument");
  B2OT_REQUIRE(dtype == B2OT_F32 || dtype == B2OT_F64, "b2ot_comm_allreduce: bad dtype %d", dtype);
  B2OT_REQUIRE(op == B2OT_OP_SUM || op == B2OT_OP_MAX, "b2ot_comm_allreduce: bad op %d", op);
  return comm_allreduce(comm, buf_dev, count, dtype, op, nullptr, (cudaStream_t)stream);
}
