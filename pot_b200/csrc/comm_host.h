// Host-side view of the peer-memory communicator shared by the solver translation units.
#pragma once

#include <cuda_runtime.h>

#include "comm.cuh"

struct b2ot_comm {
  b2ot::CommView view;
  unsigned long long seq;  // sequence number of the last exchange issued on this communicator
};

namespace b2ot {
// in-place all-reduce of `count` elements at buf (device) through the peer buffers: put + take, two launches
int comm_allreduce(b2ot_comm* c, void* buf, long long count, int dtype, int op, int* err_status, cudaStream_t s);
}  // namespace b2ot
