import sys, os, ctypes as C
sys.path.insert(0, os.getcwd())
import torch
from tests import test_ppo_cuda as T
from unitree_rl_gym_b200 import _lib
from unitree_rl_gym_b200.rsl_rl import modules as M, algorithms as A
from unitree_rl_gym_b200.rsl_rl.algorithms import ppo as P
N, TT = 4096, 24
g = torch.Generator(device="cuda").manual_seed(0)
fill = None
hist = []
def loop(self, idx):
    eng, s = self.engine, self.storage
    L, st = _lib.lib(), _lib.cur_stream()
    h = self._hparams()
    sd = s.desc()
    mini = (N * TT) // 4
    out = []
    for ep in range(5):
        for i in range(4):
            ind = idx[i * mini:(i + 1) * mini]
            _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(sd), _lib.ptr(ind), mini, 0, 0, C.byref(h), st))
            out.append(("g", eng.grads.clone(), eng.stats.clone()))
            _lib.check(L.urlg_ppo_apply(eng.handle, C.byref(h), st))
            out.append(("p", eng.flat.clone(), eng.stats.clone()))
    torch.cuda.synchronize()
    hist.append(out)
P.PPO._minibatch_loop = loop
for run in range(3):
    torch.manual_seed(0)
    ac = M.ActorCritic(48, 48, 12, [512, 256, 128], [512, 256, 128], "elu", init_noise_std=0.8)
    ppo = A.PPO(ac, device="cuda:0", **T.ALG)
    ppo.init_storage(N, TT, [48], [None], [12])
    s = ppo.storage
    names = ["observations", "actions", "mu", "rewards", "values", "returns", "advantages", "actions_log_prob"]
    if fill is None:
        fill = {k: torch.randn(getattr(s, k).shape, device="cuda", generator=g) for k in names}
    for k, v in fill.items():
        getattr(s, k).copy_(v)
    s.sigma.fill_(0.8)
    s.step = TT
    ppo.perm = torch.randperm(N * TT, generator=torch.Generator().manual_seed(2))
    if run == 0:
        p_init = ppo.engine.flat.clone()
    else:
        print("run", run, "initial params equal:", bool(torch.equal(p_init, ppo.engine.flat)))
    ppo.update()
    if run > 0:
        for k, (a, b) in enumerate(zip(hist[0], hist[run])):
            if not torch.equal(a[1], b[1]) or not torch.equal(a[2], b[2]):
                print("run", run, "first difference at event", k, a[0], "max", float((a[1] - b[1]).abs().max()),
                      "stats diff idx", (a[2] != b[2]).nonzero().flatten().tolist(), a[2].tolist()[:8], b[2].tolist()[:8], flush=True)
                break
        else:
            print("run", run, "identical", flush=True)
