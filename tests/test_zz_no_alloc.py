"""urlg_gae_noalloc (include/urlg_b200.h): the allocation-free variant of urlg_gae whose moments scratch is a static
device array of the library (no counterpart in the reference).  Called
through the C-ABI and checked against urlg_gae on the same inputs, on two calls in a row (the second with fewer envs:
partial sums left in the shared scratch by the first call must not leak into it), plus the size limit and a check that
the call leaves the device's free memory untouched.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GAMMA, LAM = 0.99, 0.95


def _storage(N, T, seed):
    from unitree_rl_gym_b200.rsl_rl.storage import RolloutStorage
    g = torch.Generator().manual_seed(seed)
    s = RolloutStorage(N, T, [4], [None], [2], "cuda:0")
    s.rewards.copy_(torch.randn(T, N, 1, generator=g))
    s.values.copy_(torch.randn(T, N, 1, generator=g))
    s.dones.copy_((torch.rand(T, N, 1, generator=g) < 0.1).byte())
    last = torch.randn(N, 1, generator=g).cuda().contiguous()
    return s, last


def _pair(N, T, seed):
    from unitree_rl_gym_b200 import _lib
    L, st = _lib.lib(), _lib.cur_stream()
    a, last = _storage(N, T, seed)
    b, _ = _storage(N, T, seed)
    moments = torch.zeros(3, dtype=torch.float64, device="cuda:0")
    _lib.check(L.urlg_gae(C.byref(a.desc()), _lib.ptr(last), GAMMA, LAM, _lib.ptr(moments), 3, st))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(L.urlg_gae_noalloc(C.byref(b.desc()), _lib.ptr(last), GAMMA, LAM, st))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return a, b, free0, free1


def test_gae_noalloc_matches_gae_twice_in_a_row():
    # tolerance of test_gae_matches_oracle (tests/test_ppo_cuda.py); the two differ only in the order of double sums
    frees = []
    for N, T, seed in ((4097, 24, 1), (300, 24, 2)):
        a, b, free0, free1 = _pair(N, T, seed)
        assert torch.equal(a.returns, b.returns)                                  # same fp32 scan
        np.testing.assert_allclose(b.advantages.cpu().numpy(), a.advantages.cpu().numpy(), rtol=1e-4, atol=1e-4)
        assert float((a.advantages - b.advantages).abs().max()) <= 1e-5
        assert torch.isfinite(b.advantages).all()
        frees.append((free0, free1))
    # the second call runs with the module loaded and the symbol resolved: it must not change the device's free memory
    assert frees[1][0] == frees[1][1], frees


def test_gae_noalloc_rejects_sizes_beyond_the_static_scratch():
    from unitree_rl_gym_b200 import _lib
    L = _lib.lib()
    s, last = _storage(64, 4, 3)
    d = _lib.Storage()
    C.memmove(C.byref(d), C.byref(s.desc()), C.sizeof(_lib.Storage))
    d.N = 1048576 + 1                       # URLG_GAE_NOALLOC_MAX_ENVS + 1; rejected before anything is launched
    before = s.advantages.clone()
    rc = L.urlg_gae_noalloc(C.byref(d), _lib.ptr(last), GAMMA, LAM, _lib.cur_stream())
    assert rc != 0
    msg = L.urlg_last_error().decode()
    assert "1048576" in msg and "exceeds" in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(before, s.advantages)
    rc = L.urlg_gae_noalloc(C.byref(s.desc()), None, GAMMA, LAM, _lib.cur_stream())
    assert rc != 0 and "last_values" in L.urlg_last_error().decode()
