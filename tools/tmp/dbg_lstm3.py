import sys, os, json, ctypes as C
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from tests import test_ppo_cuda as T
from unitree_rl_gym_b200 import _lib
N, TT = 4096, 24
oppo, pppo = T.make_pair(True, 47, 50, 12, [32], N, TT, init_noise_std=0.8)
# storage content does not matter for hazard checks: fill with plausible random data instead of the slow CPU rollout
s = pppo.storage
g = torch.Generator(device="cuda").manual_seed(0)
for name in ("observations", "privileged_observations", "actions", "mu", "rewards", "values", "returns", "advantages", "actions_log_prob"):
    t = getattr(s, name); t.copy_(torch.randn(t.shape, device="cuda", generator=g))
s.sigma.fill_(0.8)
s.dones.copy_((torch.rand(s.dones.shape, device="cuda", generator=g) < 0.05).to(torch.uint8))
eng, h = pppo._check_engine(), pppo._hparams()
L = _lib.lib()
for mb in range(2):
    _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(s.desc()), None, 0, mb * 1024, 1024, C.byref(h), _lib.cur_stream()))
    _lib.check(L.urlg_ppo_apply(eng.handle, C.byref(h), _lib.cur_stream()))
torch.cuda.synchronize()
print("done", float(eng.grads.abs().sum()))
