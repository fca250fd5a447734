"""Logging of OnPolicyRunner.learn ([UPSTREAM rsl_rl v1.0.2] runners/on_policy_runner.py learn() / log(); SURVEY.md §5
metrics row, §8(f) item 4): the per-step bookkeeping lives on the device (_EpisodeBook) and is read once per iteration,
so the rollout CUDA graph stays on with log_dir set.  Checked here: (1) the TensorBoard tags are the reference's,
(2) graph-mode and eager-mode runs write the same scalars, (3) the values equal the reference's host-side bookkeeping
(restated below with its per-step read-backs) on the same rollout."""
import statistics

import pytest
import torch

pytestmark = pytest.mark.gpu

REFERENCE_TAGS = {"Loss/value_function", "Loss/surrogate", "Loss/learning_rate", "Policy/mean_noise_std",
                  "Perf/total_fps", "Perf/collection time", "Perf/learning_time", "Train/mean_reward",
                  "Train/mean_episode_length", "Train/mean_reward/time", "Train/mean_episode_length/time"}
TIMING_TAGS = {"Perf/total_fps", "Perf/collection time", "Perf/learning_time", "Train/mean_reward/time",
               "Train/mean_episode_length/time"}


def make_runner(task, n, log_dir, cuda_graph):
    import bench
    from unitree_rl_gym_b200.rsl_rl.runners import OnPolicyRunner
    spec, frames = bench.make_frames(task, n, 4, seed=5)
    env, cfg = bench.make_env(task, n, frames, "cuda:0")
    torch.manual_seed(0)
    runner = OnPolicyRunner(env, bench.train_cfg_dict(task, cuda_graph=cuda_graph), log_dir=log_dir, device="cuda:0")
    g = torch.Generator().manual_seed(3)
    env.episode_length_buf = torch.randint(0, 1000, (n,), generator=g).cuda()
    return runner


def read_scalars(log_dir):
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    acc = EventAccumulator(str(log_dir), size_guidance={"scalars": 0})
    acc.Reload()
    return {tag: [(e.step, e.value) for e in acc.Scalars(tag)] for tag in acc.Tags()["scalars"]}


@pytest.mark.parametrize("task", ["go2", "g1"])
def test_tensorboard_scalars_graph_equals_eager(task, tmp_path, capsys):
    n, iters = 256, 3
    out = {}
    for mode in ("eager", "graph"):
        r = make_runner(task, n, str(tmp_path / mode), cuda_graph=(mode == "graph"))
        torch.manual_seed(11)
        torch.cuda.manual_seed(11)                       # the update's randperm
        r.learn(iters)
        r.writer.close()
        if mode == "graph":
            assert getattr(r, "_rollout_graph", None) is not None and r._rollout_graph.graph is not None, \
                "log_dir must not switch the rollout graph off"
        out[mode] = read_scalars(tmp_path / mode)
    printed = capsys.readouterr().out
    assert f"Learning iteration 0/{iters}" in printed and f"Learning iteration {iters - 1}/{iters}" in printed
    sum_names = r.env.spec.sum_names
    expected = REFERENCE_TAGS | {"Episode/rew_" + s for s in sum_names}
    for mode in out:
        assert set(out[mode]) == expected, (mode, set(out[mode]) ^ expected)
    for tag in sorted(expected - TIMING_TAGS):
        a, b = out["eager"][tag], out["graph"][tag]
        assert [s for s, _ in a] == [s for s, _ in b] == list(range(iters)), tag
        for (_, x), (_, y) in zip(a, b):
            assert y == pytest.approx(x, rel=1e-5, abs=1e-7), tag


def test_logged_values_equal_reference_bookkeeping(tmp_path):
    """One iteration through learn() against the reference's loop body restated with its own host reads per step
    (cur_reward_sum[new_ids].cpu() ..., torch.cat of infos['episode'][key] then mean) on a twin runner."""
    n = 256
    ra = make_runner("go2", n, str(tmp_path / "a"), cuda_graph=True)
    rb = make_runner("go2", n, None, cuda_graph=False)
    ra.learn(1)
    ra.writer.close()
    env, alg = rb.env, rb.alg
    obs = env.get_observations()
    cobs = obs
    cur_r = torch.zeros(n, device="cuda:0")
    cur_l = torch.zeros(n, device="cuda:0")
    rewbuffer, lenbuffer, ep_infos = [], [], []
    with torch.inference_mode():
        for _ in range(rb.num_steps_per_env):
            act = alg.act(obs, cobs)
            obs, priv, rew, dones, infos = env.step(act)
            cobs = priv if priv is not None else obs
            alg.process_env_step(rew, dones, infos)
            ep_infos.append({k: (v.clone() if torch.is_tensor(v) else v) for k, v in infos["episode"].items()})
            cur_r += rew
            cur_l += 1
            new_ids = (dones > 0).nonzero(as_tuple=False)
            rewbuffer.extend(cur_r[new_ids][:, 0].cpu().numpy().tolist())
            lenbuffer.extend(cur_l[new_ids][:, 0].cpu().numpy().tolist())
            cur_r[new_ids] = 0
            cur_l[new_ids] = 0
    assert len(rewbuffer) > 0
    rewbuffer, lenbuffer = rewbuffer[-100:], lenbuffer[-100:]       # deque(maxlen=100)
    got = read_scalars(tmp_path / "a")
    assert got["Train/mean_reward"][0][1] == pytest.approx(statistics.mean(rewbuffer), rel=1e-5, abs=1e-6)
    assert got["Train/mean_episode_length"][0][1] == pytest.approx(statistics.mean(lenbuffer), rel=1e-6)
    for key in ep_infos[0]:
        t = torch.cat([e[key].reshape(1) if torch.is_tensor(e[key]) else torch.tensor([e[key]], device="cuda:0")
                       for e in ep_infos])
        assert got["Episode/" + key][0][1] == pytest.approx(float(torch.mean(t)), rel=1e-5, abs=1e-7), key
    assert ra.last_stats["learning_rate"] == ra.alg.learning_rate


def test_graphs_refuse_replay_after_engine_rebuild():
    """A captured graph holds raw pointers into the policy engine; replaying after the engine was rebuilt must raise."""
    from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import RolloutGraph
    r = make_runner("go2", 128, None, cuda_graph=False)
    g = RolloutGraph(r).capture()
    g.replay()
    torch.cuda.synchronize()
    r.alg.storage.clear()
    r.alg.engine = r.alg.actor_critic.engine(max_rows=r.alg.engine.max_rows * 2)
    with pytest.raises(RuntimeError, match="rebuilt after capture"):
        g.replay()
