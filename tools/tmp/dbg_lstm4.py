import sys, os, json, ctypes as C
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from tests import test_ppo_cuda as T
from unitree_rl_gym_b200 import _lib
from unitree_rl_gym_b200.rsl_rl.algorithms import ppo as P

rec = []
orig = P.PPO._minibatch_loop
def loop(self, idx):
    eng, s = self.engine, self.storage
    L, st = _lib.lib(), _lib.cur_stream()
    h = self._hparams()
    N = s.num_envs
    eng.stats[_lib.STAT_ACC_VLOSS:_lib.STAT_ACC_N + 1] = 0.0
    sd = s.desc()
    mb = N // self.num_mini_batches
    out = []
    for ep in range(self.num_learning_epochs):
        for i in range(self.num_mini_batches):
            _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(sd), None, 0, i * mb, mb, C.byref(h), st))
            out.append(eng.grads.clone())
            _lib.check(L.urlg_ppo_apply(eng.handle, C.byref(h), st))
    torch.cuda.synchronize()
    rec.append(torch.stack(out).cpu())
P.PPO._minibatch_loop = loop

names = None
for run in range(10):
    oppo, pppo = T.make_pair(True, 47, 50, 12, [32], 4096, 24, init_noise_std=0.8)
    if run == 0:
        last = T.rollout(oppo, pppo, 4096, 24, 47, 50, 12, done_p=0.05)
        with torch.no_grad():
            oppo.compute_returns(last)
        keep = oppo.storage
    T.copy_storage(keep, pppo.storage)
    names = [(n, p.numel()) for n, p in pppo.actor_critic.named_parameters()]
    pppo.update()
    if run > 0:
        a, b = rec[0], rec[run]
        for k in range(a.shape[0]):
            d = (a[k] - b[k]).abs()
            if float(d.max()) > 1e-4 * float(a[k].abs().max()):
                off, bad = 0, []
                for n, sz in names:
                    m = float(d[off:off + sz].max()); r = float(a[k][off:off + sz].abs().max())
                    if m > 1e-4 * (r + 1e-12):
                        bad.append((n, "%.2e" % m, "%.2e" % r))
                    off += sz
                print("run", run, "first divergent minibatch", k, bad, flush=True)
                break
        else:
            print("run", run, "identical to run 0 (1e-4)", flush=True)
