"""On-disk formats either side of the path: `model_<it>.pt` written by OnPolicyRunner.save
(`legged_gym/utils/helpers.py:73-95`, `scripts/play.py:36-40`; dict layout of rsl_rl v1.0.2 OnPolicyRunner.save) and the
TorchScript policies `policy_1.pt` / `policy_lstm_1.pt` (`utils/helpers.py:150-189`) that deploy_mujoco.py:77 /
deploy_real.py:29 load.

  * a checkpoint written by the B200 runner loads into the ORACLE's ActorCritic[Recurrent] + a stock torch.optim.Adam
    (what the unmodified reference would construct), and the other way round;
  * save -> load is the identity on parameters, Adam moments and step counter;
  * the exported TorchScript modules reproduce the CUDA act_inference path within the forward tolerance (1e-4).
"""
import copy
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def make_runner(task, n, seed=0):
    from tests.test_graph_cuda import make_runner as mk
    return mk(task, n)


def one_iteration(runner, seed):
    from tests.test_graph_cuda import eager_rollout, finish
    T = runner.num_steps_per_env
    n = runner.env.num_envs
    mini = (n * T) // 4
    perm = torch.randperm(4 * mini, generator=torch.Generator().manual_seed(seed))
    return finish(runner, eager_rollout(runner), perm)


def oracle_policy(task, runner):
    """The reference-side objects: oracle port of rsl_rl's module classes + stock Adam."""
    from oracle import rsl_rl_port as O
    import bench
    cfg = bench.train_cfg_dict(task)
    pol = dict(cfg["policy"])
    env = runner.env
    n_critic = env.num_privileged_obs if env.num_privileged_obs is not None else env.num_obs
    cls = O.ActorCriticRecurrent if cfg["runner"]["policy_class_name"] == "ActorCriticRecurrent" else O.ActorCritic
    ac = cls(env.num_obs, n_critic, env.num_actions, **pol)
    return ac, torch.optim.Adam(ac.parameters(), lr=1e-3)


@pytest.mark.parametrize("task", ["go2", "g1"])
def test_checkpoint_interchanges_with_reference_classes(task, tmp_path):
    r = make_runner(task, 128)
    for it in range(2):
        one_iteration(r, 30 + it)
    r.current_learning_iteration = 2
    path = os.path.join(str(tmp_path), "model_2.pt")
    r.save(path, infos={"note": "x"})
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert set(ck) == {"model_state_dict", "optimizer_state_dict", "iter", "infos"}
    assert ck["iter"] == 2 and ck["infos"] == {"note": "x"}

    # -> reference side: strict load into the oracle's module + stock Adam
    oac, oopt = oracle_policy(task, r)
    assert list(ck["model_state_dict"]) == list(oac.state_dict())
    oac.load_state_dict(ck["model_state_dict"], strict=True)
    oopt.load_state_dict(ck["optimizer_state_dict"])
    st = oopt.state_dict()["state"]
    assert len(st) == len(list(oac.parameters()))
    for i, p in enumerate(oac.parameters()):
        assert st[i]["exp_avg"].shape == p.shape and float(st[i]["step"]) == 40.0     # 2 iterations x 20 minibatches

    # <- reference side: a checkpoint written by stock torch objects loads into the B200 runner
    path2 = os.path.join(str(tmp_path), "model_ref.pt")
    torch.save({"model_state_dict": oac.state_dict(), "optimizer_state_dict": oopt.state_dict(), "iter": 7, "infos": None}, path2)
    r2 = make_runner(task, 128)
    one_iteration(r2, 99)                                   # make its state differ first
    assert r2.load(path2) is None and r2.current_learning_iteration == 7
    for (k, p), (_, q) in zip(r.alg.actor_critic.named_parameters(), r2.alg.actor_critic.named_parameters()):
        assert torch.equal(p, q), k
    a, b = r.alg.optimizer.state_dict(), r2.alg.optimizer.state_dict()
    for i in a["state"]:
        for key in ("exp_avg", "exp_avg_sq"):
            assert torch.equal(a["state"][i][key], b["state"][i][key]), (i, key)
        assert float(a["state"][i]["step"]) == float(b["state"][i]["step"])


def test_exported_mlp_policy_matches_cuda_inference(tmp_path):
    from unitree_rl_gym_b200.legged_gym.utils.helpers import export_policy_as_jit
    r = make_runner("go2", 128)
    one_iteration(r, 5)
    export_policy_as_jit(r.alg.actor_critic, str(tmp_path))
    pol = torch.jit.load(os.path.join(str(tmp_path), "policy_1.pt"), map_location="cpu")
    obs = torch.randn(128, r.env.num_obs, generator=torch.Generator().manual_seed(1))
    want = pol(obs)
    got = r.get_inference_policy(device="cuda:0")(obs.cuda())
    np.testing.assert_allclose(got.cpu().numpy(), want.detach().numpy(), rtol=1e-4, atol=1e-4)


def test_exported_lstm_policy_matches_cuda_inference(tmp_path):
    from unitree_rl_gym_b200.legged_gym.utils.helpers import export_policy_as_jit
    r = make_runner("g1", 128)
    one_iteration(r, 5)
    ac = r.alg.actor_critic
    export_policy_as_jit(ac, str(tmp_path))
    pol = torch.jit.load(os.path.join(str(tmp_path), "policy_lstm_1.pt"), map_location="cpu")
    assert {"hidden_state", "cell_state"} <= {k for k, _ in pol.named_buffers()}
    pol.reset_memory()
    act = r.get_inference_policy(device="cuda:0")
    ac.reset(torch.ones(128, dtype=torch.bool, device="cuda:0"))           # zero hidden state on every env
    g = torch.Generator().manual_seed(2)
    for t in range(6):                                                     # the exported module is a batch-1 recurrence
        obs = torch.randn(1, r.env.num_obs, generator=g)
        want = pol(obs)
        got = act(obs.expand(128, -1).contiguous().cuda())
        np.testing.assert_allclose(got[:1].cpu().numpy(), want.detach().numpy(), rtol=1e-4, atol=1e-4, err_msg=f"t={t}")
        assert torch.equal(got[0], got[127])
