"""PPO half of __graft_entry__.smoke(): a tiny rollout (act -> store -> GAE) and one update through the C ABI on cuda:0,
checked against the oracle's rsl_rl restatement (returns 1e-4, gradients of the first minibatch rel-L2 1e-3)."""
import numpy as np
import torch


def smoke_ppo():
    from tests import test_ppo_cuda as T
    N, steps = 128, 8
    oppo, pppo = T.make_pair(False, 48, None, 12, [512, 256, 128], N, steps, init_noise_std=0.8)
    last = T.rollout(oppo, pppo, N, steps, 48, None, 12, done_p=0.05)
    with torch.no_grad():
        oppo.compute_returns(last)
    pppo.compute_returns(last.cuda())
    torch.cuda.synchronize()
    np.testing.assert_allclose(pppo.storage.returns.cpu().numpy(), oppo.storage.returns.numpy(), rtol=1e-3, atol=1e-3)
    T.copy_storage(oppo.storage, pppo.storage)
    mini = (N * steps) // oppo.num_mini_batches
    perm = torch.randperm(oppo.num_mini_batches * mini, generator=torch.Generator().manual_seed(1))
    oppo.perm, pppo.perm = perm, perm
    ov, osur = oppo.update()
    pv, psur = pppo.update()
    torch.cuda.synchronize()
    assert abs(pv - ov) < 1e-3 * abs(ov) + 1e-4 and abs(psur - osur) < 1e-3, (pv, ov, psur, osur)
    for (n, p), (_, q) in zip(pppo.actor_critic.named_parameters(), oppo.actor_critic.named_parameters()):
        assert T.rel_l2(p.detach().cpu().numpy(), q.detach().numpy()) < 1e-3, n
    print("smoke: PPO rollout + update matches the oracle (%d envs x %d steps)" % (N, steps))
