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


def _filled_ppo(seed, state=None, N=256, T=8):
    """A product PPO (feed-forward Go2 shape) whose storage holds one synthetic rollout; `state` = parameters to start from."""
    import copy
    from tests.test_ppo_cuda import ALG
    from unitree_rl_gym_b200.rsl_rl import modules as M, algorithms as A
    torch.manual_seed(0)
    ac = M.ActorCritic(48, 48, 12, [512, 256, 128], [512, 256, 128], "elu", init_noise_std=0.8)
    if state is not None:
        ac.load_state_dict(copy.deepcopy(state))
    alg = A.PPO(ac, device="cuda:0", **ALG)
    alg.init_storage(N, T, [48], [None], [12])
    g = torch.Generator(device="cuda").manual_seed(seed)
    rn = lambda *s: torch.randn(*s, device="cuda", generator=g)
    for t in range(T):
        obs = rn(N, 48)
        alg.act(obs, obs)
        dones = torch.rand(N, device="cuda", generator=g) < 0.03
        alg.process_env_step(rn(N), dones, {"time_outs": dones & (torch.rand(N, device="cuda", generator=g) < 0.5)})
    alg.compute_returns(rn(N, 48))
    return alg


def _one_step(alg, idx, h, phase):
    from unitree_rl_gym_b200 import _lib
    eng, s, L = alg._check_engine(), alg.storage, _lib.lib()
    if phase == "grads":
        _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(s.desc()), _lib.ptr(idx), idx.numel(), 0, 0, C.byref(h), _lib.cur_stream()))
    else:
        _lib.check(L.urlg_ppo_apply(eng.handle, C.byref(h), _lib.cur_stream()))


@pytest.mark.parametrize("mode", [0, 1])
def test_fused_exchange_equals_separate_allreduce_and_keeps_ranks_identical(mode):
    """Two logical ranks on one GPU, three minibatch steps, one-shot (mode 0) and two-shot (mode 1: every rank reduces its
    slice, stores it into both result areas and waits for the other's -- the two apply kernels run on separate streams so
    that each can see the other's slice arrive).  Fused: urlg_policy_bind_peer_exchange (finalize publishes, the
    norm kernel reduces).  Reference: the same gradients summed in rank order between urlg_ppo_minibatch_grads and
    urlg_ppo_apply.  The ranks of the fused run stay bit-identical to each other and agree with the reference to rounding
    (the two norm kernels add the squares in a different association)."""
    from unitree_rl_gym_b200 import _lib
    from unitree_rl_gym_b200.rsl_rl.peer_exchange import LocalPeerGroup
    base = _filled_ppo(11)
    state = {k: v.detach().clone() for k, v in base.actor_critic.state_dict().items()}
    fused = [_filled_ppo(11 + r, state) for r in range(2)]
    ref = [_filled_ppo(11 + r, state) for r in range(2)]
    n = fused[0].engine.grads_and_sums.numel()
    grp = LocalPeerGroup(2, n, torch.device("cuda", 0), mode=mode)
    streams = [torch.cuda.Stream() for _ in range(2)]
    for r in range(2):
        _lib.check(_lib.lib().urlg_policy_bind_peer_exchange(fused[r].engine.handle, C.byref(grp.ranks[r].x)))
    N, T = 256, 8
    mini = N * T // 4
    perm = torch.randperm(N * T, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    for step in range(3):
        idx = perm[step * mini:(step + 1) * mini].contiguous()
        hs = []
        for alg in fused + ref:
            h = alg._hparams()
            h.grad_scale, h.kl_count_scale = 0.5, 2.0
            hs.append(h)
        for r in range(2):
            _one_step(fused[r], idx, hs[r], "grads")           # publishes into rank r's half, flags both rows
        torch.cuda.synchronize()
        for r in range(2):
            with torch.cuda.stream(streams[r]):
                _one_step(fused[r], idx, hs[r], "apply")       # sums the halves in rank order inside the norm kernel
        torch.cuda.synchronize()
        for r in range(2):
            _one_step(ref[r], idx, hs[2 + r], "grads")
        torch.cuda.synchronize()
        total = torch.zeros_like(ref[0].engine.grads_and_sums)
        for r in range(2):
            total = total + ref[r].engine.grads_and_sums
        for r in range(2):
            ref[r].engine.grads_and_sums.copy_(total)
            _one_step(ref[r], idx, hs[2 + r], "apply")
        torch.cuda.synchronize()
        assert torch.equal(fused[0].engine.flat, fused[1].engine.flat), step
        assert torch.equal(fused[0].engine.grads_and_sums, total), step          # the reduced gradient itself is bit-equal
        d = (fused[0].engine.flat - ref[0].engine.flat).abs().max().item()
        moved = (ref[0].engine.flat - base.engine.flat).abs().max().item()
        assert moved > 1e-4 and d <= 1e-6 + 1e-4 * moved, (step, d, moved)
    assert int(grp.ranks[0]._counters[0]) == 3
    assert _lib.lib().urlg_policy_bind_peer_exchange(fused[0].engine.handle, None) == 0
