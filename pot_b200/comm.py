"""NVLink peer-memory communicator for the row-sharded solvers (b2ot_comm_* in include/b2ot.h).

One process per GPU (torchrun).  Every rank allocates one exchange buffer through the library, the CUDA-IPC
handles travel through `torch.distributed.all_gather_object` (host plumbing, once), each rank maps its peers'
buffers, and from then on the per-iteration exchange of the solvers is plain NVLink stores + flag polling
inside the library's own kernels (csrc/comm.cuh) - no NCCL launch on the hot path.
"""

import ctypes as C

import torch

from . import _lib

__all__ = ["PeerComm", "get_comm"]

_HEADER = 4096
_cache = {}


class PeerComm:
    """Exchange buffers of all ranks of `group`, peer-mapped; `slot_bytes` is the largest vector one rank can
    contribute to an exchange."""

    def __init__(self, group, device, slot_bytes):
        import torch.distributed as dist

        self.lib = _lib.load()
        self.group, self.device = group, torch.device(device)
        self.nranks = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.nranks > 8:
            raise ValueError("pot_b200: the peer-memory exchange spans one NVSwitch node (at most 8 ranks)")
        slot = (int(slot_bytes) + 255) // 256 * 256
        self.bytes = _HEADER + 2 * self.nranks * slot
        own = C.c_void_p()
        handle = C.create_string_buffer(_lib.IPC_HANDLE_BYTES)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_comm_alloc(self.bytes, C.byref(own), handle), "b2ot_comm_alloc")
        self.own = own
        handles = [None] * self.nranks
        dist.all_gather_object(handles, bytes(handle.raw), group=group)
        self.peers = []
        ptrs = (C.c_void_p * self.nranks)()
        with torch.cuda.device(self.device):
            for q in range(self.nranks):
                if q == self.rank:
                    ptrs[q] = own.value
                    continue
                p = C.c_void_p()
                _lib.check(self.lib.b2ot_comm_open(handles[q], C.byref(p)), f"b2ot_comm_open(rank {q})")
                self.peers.append(p)
                ptrs[q] = p.value
            h = C.c_void_p()
            _lib.check(self.lib.b2ot_comm_create(C.byref(h), self.nranks, self.rank, ptrs, self.bytes),
                       "b2ot_comm_create")
        self.h = h
        self.slot_bytes = self.lib.b2ot_comm_slot_bytes(h)
        dist.barrier(group=group)  # every rank has mapped every buffer before anyone stores into one

    def all_reduce(self, t, op="sum"):
        """In-place all-reduce of a contiguous CUDA tensor through the exchange buffers (tests / utilities)."""
        from ._dense import dtype_code, stream_ptr

        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2ot_comm_allreduce(self.h, t.data_ptr(), t.numel(), dtype_code(t.dtype),
                                                    _lib.OP_MAX if op == "max" else _lib.OP_SUM,
                                                    stream_ptr(self.device)), "b2ot_comm_allreduce")
        return t

    def close(self):
        if getattr(self, "h", None):
            torch.cuda.synchronize(self.device)
            self.lib.b2ot_comm_destroy(self.h)
            self.h = None
            with torch.cuda.device(self.device):
                for p in self.peers:
                    self.lib.b2ot_comm_close(p)
                self.lib.b2ot_comm_free(self.own)
            self.peers = []


def get_comm(group, device, slot_bytes):
    """Cached communicator of (group, device) with slots of at least `slot_bytes` (collective call)."""
    key = (id(group), str(device))
    c = _cache.get(key)
    if c is None or c.slot_bytes < slot_bytes:
        if c is not None:
            c.close()
        c = PeerComm(group, device, max(int(slot_bytes), 1 << 20))
        _cache[key] = c
    return c
