"""Host side of the data-parallel gradient exchange over peer memory (include/urlg_b200.h urlg_peer_allreduce).

`PeerExchange(n, device)` allocates one rank's exchange buffer in symmetric memory (torch.distributed._symmetric_memory:
CUDA VMM allocations that every rank of the group maps over NVLink), exchanges the peer-mapped pointers at rendezvous and
fills the C struct; `all_reduce(tensor)` then replaces the NCCL all-reduce of PPO.update's flat gradient by the library's
one-shot publish / reduce kernels.  The reference is single-GPU; the call site this serves is [UPSTREAM rsl_rl v1.0.2]
algorithms/ppo.py update (gradient step per minibatch), sharded as SURVEY.md 8(e) describes.

`LocalPeerGroup` builds the same structs for several logical ranks inside ONE process and ONE device (plain device memory,
every "peer" pointer is a local pointer): how tests/test_peer_exchange_cuda.py drives the kernels on a single GPU.
"""
import ctypes as C

import torch

from .. import _lib


def _round4(n):
    return (int(n) + 3) // 4 * 4


class _Exchange:
    """Common part: the C struct, its device-resident pointer tables and counters."""

    @staticmethod
    def floats_needed(count):
        """[2 halves | result area (two-shot) | publish flags + slice flags]"""
        return 3 * count + 2 * _lib.PEER_MAX_WORLD

    def _fill(self, world, rank, buf_ptrs, count, device, mode=None):
        if world > _lib.PEER_MAX_WORLD:
            raise ValueError("peer exchange supports at most %d ranks, got %d" % (_lib.PEER_MAX_WORLD, world))
        self.world, self.rank, self.count = world, rank, count
        flag_ptrs = [p + 3 * count * 4 for p in buf_ptrs]
        # one-shot reads world x n floats per rank (fine at 2 ranks); two-shot moves 2 (world - 1) / world x n
        self.mode = (1 if world > 2 else 0) if mode is None else int(mode)
        self._tables = torch.tensor([buf_ptrs, flag_ptrs], dtype=torch.int64, device=device)
        self._counters = torch.zeros(4, dtype=torch.int32, device=device)      # epoch, ticket[2], pad
        x = self.x = _lib.PeerExchange()
        x.world, x.rank, x.count, x.mode = world, rank, count, self.mode
        x.bufs = self._tables[0].data_ptr()
        x.flags = self._tables[1].data_ptr()
        x.epoch = self._counters.data_ptr()
        x.ticket = self._counters.data_ptr() + 4

    def all_reduce(self, t, stream=None):
        """t (contiguous fp32 CUDA tensor, <= count elements) <- sum over ranks, in rank order on every rank."""
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() > self.count:
            raise ValueError("peer all_reduce: contiguous fp32 tensor of at most %d elements" % self.count)
        _lib.check(_lib.lib().urlg_peer_allreduce(C.byref(self.x), _lib.ptr(t), t.numel(),
                                                  stream if stream is not None else _lib.cur_stream()))


class PeerExchange(_Exchange):
    """One rank of a torch.distributed group (NCCL backend, one process per GPU)."""

    def __init__(self, n, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        count = _round4(n)
        total = self.floats_needed(count)
        self.buf = symm_mem.empty(total, dtype=torch.float32, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, group.group_name)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        import os
        mode = os.environ.get("URLG_PEER_MODE")
        self._fill(dist.get_world_size(group), dist.get_rank(group), ptrs, count, device, None if mode is None else int(mode))
        torch.cuda.synchronize(device)
        dist.barrier(group)            # every rank's buffer and flag row are zero before anyone publishes


class LocalPeerGroup:
    """`world` logical ranks on one device (tests): rank r's exchange is .ranks[r]."""

    def __init__(self, world, n, device, mode=None):
        count = _round4(n)
        total = _Exchange.floats_needed(count)
        self.bufs = [torch.zeros(total, dtype=torch.float32, device=device) for _ in range(world)]
        ptrs = [b.data_ptr() for b in self.bufs]
        self.ranks = []
        for r in range(world):
            e = _Exchange()
            e._fill(world, r, ptrs, count, device, mode)
            self.ranks.append(e)
