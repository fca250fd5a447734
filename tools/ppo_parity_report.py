#!/usr/bin/env python
"""Prints the numerical distance between the CUDA PPO half and the oracle (gradient of one minibatch and the
parameters after one full update) — the numbers behind the tolerances in tests/test_ppo_cuda.py."""
import copy
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from tests import test_ppo_cuda as T
from unitree_rl_gym_b200 import _lib

CASES = [(False, 48, None, 12, [512, 256, 128], 512, 24), (True, 47, 50, 12, [32], 256, 24)]
for rec, no, npv, na, hid, N, TT in CASES:
    oppo, pppo = T.make_pair(rec, no, npv, na, hid, N, TT, init_noise_std=0.8)
    errs = []
    last = T.rollout(oppo, pppo, N, TT, no, npv, na, done_p=0.05, check=lambda t, a, b: errs.append(float((a - b).abs().max())))
    torch.cuda.synchronize()
    out = {"recurrent": rec, "N": N, "T": TT, "precision": pppo.engine.precision, "act_max_abs_err": max(errs)}
    for k in ("mu", "values", "actions_log_prob"):
        out["store_" + k] = float((getattr(pppo.storage, k).cpu() - getattr(oppo.storage, k)).abs().max())
    with torch.no_grad():
        oppo.compute_returns(last)
    T.copy_storage(oppo.storage, pppo.storage)
    torch.manual_seed(5)
    with torch.no_grad():
        for p in oppo.actor_critic.parameters():
            p.add_(0.02 * torch.randn_like(p) * p.abs().mean())
    pppo.actor_critic.load_state_dict(copy.deepcopy(oppo.actor_critic.state_dict()))
    mini = (N * TT) // 4
    perm = torch.randperm(4 * mini, generator=torch.Generator().manual_seed(1))
    want, surr, vl = T.oracle_minibatch_grads(oppo, perm)
    eng, s, h = pppo._check_engine(), pppo.storage, pppo._hparams()
    if rec:
        args = (None, 0, 0, N // 4)
    else:
        ind = perm[:mini].cuda().contiguous()
        args = (_lib.ptr(ind), mini, 0, 0)
    _lib.check(_lib.lib().urlg_ppo_minibatch_grads(eng.handle, C.byref(s.desc()), *args, C.byref(h), _lib.cur_stream()))
    torch.cuda.synchronize()
    got = {n: p.grad.detach().cpu() for n, p in pppo.actor_critic.named_parameters()}
    out["grad_rel_l2"] = {n: round(T.rel_l2(got[n].numpy(), want[n].numpy()), 6) for n in want}
    oppo.perm, pppo.perm = perm, perm
    before = {n: p.detach().clone() for n, p in oppo.actor_critic.named_parameters()}
    ov, osur = oppo.update()
    pv, psur = pppo.update()
    out["losses"] = [ov, pv, osur, psur]
    out["lr"] = [oppo.learning_rate, pppo.learning_rate]
    worst, rl = 0.0, {}
    op = dict(oppo.actor_critic.named_parameters())
    frac = []
    for n, p in pppo.actor_critic.named_parameters():
        a, b = p.detach().cpu().numpy(), op[n].detach().numpy()
        rl[n] = round(T.rel_l2(a, b), 6)
        worst = max(worst, float(np.abs(a - b).max()))
        frac.append(float(np.mean(np.abs(a - b) <= 1e-3)))
    out["update_param_rel_l2"] = rl
    out["update_param_max_abs"] = worst
    out["update_frac_within_1e-3"] = min(frac)
    print(json.dumps(out), flush=True)
