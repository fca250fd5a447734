import sys, os, json, ctypes as C
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from tests import test_ppo_cuda as T
from unitree_rl_gym_b200 import _lib
N, TT = 4096, 24
oppo, pppo = T.make_pair(True, 47, 50, 12, [32], N, TT, init_noise_std=0.8)
last = T.rollout(oppo, pppo, N, TT, 47, 50, 12, done_p=0.05)
with torch.no_grad():
    oppo.compute_returns(last)
T.copy_storage(oppo.storage, pppo.storage)
eng, s, h = pppo._check_engine(), pppo.storage, pppo._hparams()
names = [(n, p.numel()) for n, p in pppo.actor_critic.named_parameters()]
ref = {}
L = _lib.lib()
for rep in range(8):
    for mb in range(4):
        _lib.check(L.urlg_ppo_minibatch_grads(eng.handle, C.byref(s.desc()), None, 0, mb * 1024, 1024, C.byref(h), _lib.cur_stream()))
        torch.cuda.synchronize()
        g = eng.grads.clone()
        st = eng.stats[:4].clone()
        if rep == 0:
            ref[mb] = (g, st)
        else:
            d = (g - ref[mb][0]).abs()
            if float(d.max()) > 1e-5 * float(ref[mb][0].abs().max()):
                off = 0
                bad = []
                for n, k in names:
                    m = float(d[off:off + k].max())
                    if m > 1e-5 * float(ref[mb][0][off:off+k].abs().max() + 1e-12):
                        bad.append((n, m, float(ref[mb][0][off:off + k].abs().max())))
                    off += k
                print("rep", rep, "mb", mb, "DIFF", bad, "stats", (st - ref[mb][1]).tolist(), flush=True)
print("done")
