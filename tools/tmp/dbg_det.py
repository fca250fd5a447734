import sys, os, ctypes as C
sys.path.insert(0, os.getcwd())
import torch
from tests import test_ppo_cuda as T
from unitree_rl_gym_b200 import _lib
from unitree_rl_gym_b200.rsl_rl import modules as M, algorithms as A
N, TT = 4096, 24
torch.manual_seed(0)
ac = M.ActorCritic(48, 48, 12, [512, 256, 128], [512, 256, 128], "elu", init_noise_std=0.8)
ppo = A.PPO(ac, device="cuda:0", **T.ALG)
ppo.init_storage(N, TT, [48], [None], [12])
s = ppo.storage
g = torch.Generator(device="cuda").manual_seed(0)
for k in ["observations", "actions", "mu", "rewards", "values", "returns", "advantages", "actions_log_prob"]:
    t = getattr(s, k); t.copy_(torch.randn(t.shape, device="cuda", generator=g))
s.sigma.fill_(0.8)
eng, h = ppo._check_engine(), ppo._hparams()
idx = torch.randperm(N * TT, generator=torch.Generator().manual_seed(2)).cuda()[:N * TT // 4].contiguous()
L = _lib.lib()
ref = None
names = [(n, p.numel()) for n, p in ac.named_parameters()]
for rep in range(6):
    _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(s.desc()), _lib.ptr(idx), idx.numel(), 0, 0, C.byref(h), _lib.cur_stream()))
    torch.cuda.synchronize()
    gr = eng.grads.clone(); st = eng.stats[:4].clone()
    if ref is None:
        ref = (gr, st)
    else:
        off = 0; bad = []
        for n, k in names:
            d = (gr[off:off+k] != ref[0][off:off+k]).sum().item()
            if d: bad.append((n, d, float((gr[off:off+k]-ref[0][off:off+k]).abs().max())))
            off += k
        print("rep", rep, "diff tensors:", bad, "stats equal", bool((st == ref[1]).all()), flush=True)
# apply determinism
p0 = eng.flat.clone(); m0 = eng.exp_avg.clone(); v0 = eng.exp_avg_sq.clone(); st0 = eng.stats.clone()
outs = []
for rep in range(3):
    eng.flat.copy_(p0); eng.exp_avg.copy_(m0); eng.exp_avg_sq.copy_(v0); eng.stats.copy_(st0); eng.grads.copy_(ref[0])
    _lib.check(L.urlg_ppo_apply(eng.handle, C.byref(h), _lib.cur_stream()))
    torch.cuda.synchronize()
    outs.append(eng.flat.clone())
print("apply deterministic:", bool(torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])))
