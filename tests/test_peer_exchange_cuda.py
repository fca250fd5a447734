"""urlg_peer_allreduce through the C ABI on ONE GPU: `world` logical ranks (LocalPeerGroup: every "peer" pointer is a
local pointer), each on its own stream, so rank 0's reduce kernel really spins on its flag row until the other ranks'
publish kernels have run.  The sum must be the rank-ordered fp32 sum, bit for bit, on every rank, exchange after exchange
(both buffer halves, growing epochs).  The multi-process path over NVLink is covered by tools/multi_gpu_parity.py."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,n", [(1, 1000), (2, 4099), (3, 10001), (8, 20_002), (2, 380_004)])
def test_one_shot_allreduce_is_the_rank_ordered_sum(world, n):
    from unitree_rl_gym_b200.rsl_rl.peer_exchange import LocalPeerGroup
    dev = torch.device("cuda", 0)
    grp = LocalPeerGroup(world, n, dev)
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    g = torch.Generator(device="cuda").manual_seed(5)
    for it in range(5):                                   # 5 exchanges: halves 0,1,0,1,0 and epochs 0..4
        data = [torch.randn(n, device=dev, generator=g) * (1 + r) for r in range(world)]
        want = torch.zeros(n, device=dev)
        for r in range(world):
            want = want + data[r]                         # rank order, fp32: what every rank must hold afterwards
        torch.cuda.synchronize()
        # (on one GPU the spinning reduce grids of all logical ranks must leave room for the later ranks' publish kernels:
        # world x blocks stays far below the 148 x 8 resident blocks in every case here)
        for r in range(world):
            grp.ranks[r].all_reduce(data[r], C.c_void_p(streams[r].cuda_stream))
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(data[r], want), (it, r)
    assert int(grp.ranks[0]._counters[0]) == 5            # device-resident epoch


def test_rejects_bad_arguments():
    from unitree_rl_gym_b200 import _lib
    from unitree_rl_gym_b200.rsl_rl.peer_exchange import LocalPeerGroup
    grp = LocalPeerGroup(2, 64, torch.device("cuda", 0))
    t = torch.zeros(128, device="cuda")
    L = _lib.lib()
    assert L.urlg_peer_allreduce(C.byref(grp.ranks[0].x), _lib.ptr(t), 128, _lib.cur_stream()) != 0      # larger than the buffer
    assert b"do not fit" in L.urlg_last_error()
    bad = _lib.PeerExchange()
    C.memmove(C.byref(bad), C.byref(grp.ranks[0].x), C.sizeof(bad))
    bad.world = _lib.PEER_MAX_WORLD + 1
    assert L.urlg_peer_allreduce(C.byref(bad), _lib.ptr(t), 64, _lib.cur_stream()) != 0
