"""CUDA-graph paths replay exactly what the eager public API does: a whole rollout as one graph (RolloutGraph), one
graph per env step (StepGraphs) and the 20 minibatch steps of PPO.update as a graph.  Two identical runners (same
seeds, same synthetic frames; every random draw is counter-based Philox keyed by device-side step counters) must end
with bit-identical storage / parameters whichever way they are launched."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def make_runner(task, n, frames_seed=5):
    import bench
    from unitree_rl_gym_b200.rsl_rl.runners import OnPolicyRunner
    spec, frames = bench.make_frames(task, n, 4, seed=frames_seed)
    env, cfg = bench.make_env(task, n, frames, "cuda:0")
    torch.manual_seed(0)                                 # identical nn.Module initialisation for every runner
    runner = OnPolicyRunner(env, bench.train_cfg_dict(task), log_dir=None, device="cuda:0")
    g = torch.Generator().manual_seed(3)
    env.episode_length_buf = torch.randint(0, 1000, (n,), generator=g).cuda()
    return runner


def eager_rollout(runner):
    env, alg = runner.env, runner.alg
    obs = env.get_observations()
    priv = env.get_privileged_observations()
    cobs = priv if priv is not None else obs
    with torch.inference_mode():
        for _ in range(runner.num_steps_per_env):
            act = alg.act(obs, cobs)
            obs, priv, rew, dones, infos = env.step(act)
            cobs = priv if priv is not None else obs
            alg.process_env_step(rew, dones, infos)
    return cobs


def finish(runner, cobs, perm=None):
    alg = runner.alg
    with torch.inference_mode():
        alg.compute_returns(cobs)
    alg.perm = perm
    return alg.update()


def storage_equal(a, b):
    for name in ("observations", "privileged_observations", "actions", "mu", "sigma", "rewards", "values",
                 "actions_log_prob", "dones"):
        x, y = getattr(a, name), getattr(b, name)
        if x is not None:
            assert torch.equal(x, y), name
    if a.saved_hidden_states_a is not None:
        for i in range(2):
            assert torch.equal(a.saved_hidden_states_a[i], b.saved_hidden_states_a[i])
            assert torch.equal(a.saved_hidden_states_c[i], b.saved_hidden_states_c[i])


@pytest.mark.parametrize("task", ["go2", "g1"])
@pytest.mark.parametrize("mode", ["rollout", "steps"])
def test_graph_replay_equals_eager(task, mode):
    from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import RolloutGraph, StepGraphs
    n = 192
    ra, rb = make_runner(task, n), make_runner(task, n)
    T = ra.num_steps_per_env
    mini = (n * T) // 4
    perm = torch.randperm(4 * mini, generator=torch.Generator().manual_seed(9))
    def sync_params():
        # gradient sums use atomics (split-K wgrad, bias column sums), so two updates agree only to the last bits;
        # copy the parameters across so that the ROLLOUTS under comparison start from identical bits
        rb.alg.engine.flat.copy_(ra.alg.engine.flat)
        rb.alg.engine.sync_weights()
    for r in (ra, rb):                                   # iteration 0 eagerly on both (also the warm-up before capture)
        finish(r, eager_rollout(r), perm)
    sync_params()
    if mode == "rollout":
        g = RolloutGraph(rb).capture()
        replay = lambda: g.replay()[1]
    else:
        sg = StepGraphs(rb).capture()

        def replay():
            for t in range(T):
                out = sg.replay(t)
            return out[1]
    for it in range(2):                                  # two iterations: the device-side counters must keep advancing
        ca = eager_rollout(ra)
        cb = replay()
        torch.cuda.synchronize()
        storage_equal(ra.alg.storage, rb.alg.storage)
        assert torch.equal(ca, cb)
        for name in ("rew_buf", "reset_buf", "episode_length_buf", "commands", "root_states", "dof_state"):
            assert torch.equal(getattr(ra.env, name), getattr(rb.env, name)), name
        finish(ra, ca, perm)
        finish(rb, cb, perm)
        sync_params()


@pytest.mark.parametrize("task", ["go2", "g1"])
def test_update_graph_equals_eager(task):
    n = 128
    ra, rb = make_runner(task, n), make_runner(task, n)
    rb.alg.cuda_graph_update = True
    T = ra.num_steps_per_env
    mini = (n * T) // 4
    for it in range(4):                                  # rb: eager warm-up, capture, then two replays
        perm = torch.randperm(4 * mini, generator=torch.Generator().manual_seed(20 + it))
        la = finish(ra, eager_rollout(ra), perm)
        lb = finish(rb, eager_rollout(rb), perm)
        torch.cuda.synchronize()
        assert rb.alg.learning_rate == ra.alg.learning_rate
        np.testing.assert_allclose(la, lb, rtol=1e-3, atol=1e-5)
        for (k, p), (_, q) in zip(ra.alg.actor_critic.named_parameters(), rb.alg.actor_critic.named_parameters()):
            # split-K / bias-gradient atomics make the two runs differ in the last bits of a gradient sum; for weights
            # whose gradient is at noise level Adam turns that into steps of opposite sign (|step| <= lr = 1e-3), so
            # the two trajectories drift apart by a few 1e-4 over the 80 steps: every element stays within 1e-3, the
            # mean drift within 1e-4
            a, b = p.detach().cpu().numpy(), q.detach().cpu().numpy()
            np.testing.assert_allclose(a, b, rtol=0, atol=1e-3, err_msg=k)
            assert np.abs(a - b).mean() <= 1e-4, k          # typical drift is 2-4e-5
    assert rb.alg._update_graph is not None


def test_release_graphs_drops_and_recaptures():
    """PPO.release_graphs() (run before dist.destroy_process_group when the graph holds NCCL kernels) destroys the captured
    update graph; the next updates warm up / capture again and keep training."""
    n = 128
    r = make_runner("go2", n)
    r.alg.cuda_graph_update = True
    T = r.num_steps_per_env
    perm = torch.randperm(n * T, generator=torch.Generator().manual_seed(5))
    for _ in range(2):
        finish(r, eager_rollout(r), perm)
    assert r.alg._update_graph is not None
    r.alg.release_graphs()
    assert r.alg._update_graph is None
    before = r.alg.engine.flat.clone()
    finish(r, eager_rollout(r), perm)                    # same key: captures again
    torch.cuda.synchronize()
    assert r.alg._update_graph is not None
    assert not torch.equal(before, r.alg.engine.flat)
    assert torch.isfinite(r.alg.engine.flat).all()


def test_rollout_graph_with_host_traffic_hooks():
    """RolloutGraph.capture(pre_step, post_step, epilogue): frame uploads from pinned host memory and per-step reward /
    done read-backs recorded on forked streams inside the capture (one launch per rollout, bench.py's e2e path) leave
    storage bit-identical to the eager loop, and the host rows equal the storage rows."""
    from unitree_rl_gym_b200 import _lib
    from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import RolloutGraph
    n = 192
    ra, rb = make_runner("go2", n), make_runner("go2", n)
    T = ra.num_steps_per_env
    perm = torch.randperm(n * T, generator=torch.Generator().manual_seed(9))
    for r in (ra, rb):
        finish(r, eager_rollout(r), perm)
    rb.alg.engine.flat.copy_(ra.alg.engine.flat)
    rb.alg.engine.sync_weights()
    pool = rb.env.gym.frames
    nf = rb.env.gym.num_frames
    pin = {k: v.cpu().pin_memory() for k, v in pool.items() if v is not None}
    up = [_lib.CopyBatch([(pool[k][f], pin[k][f]) for k in pin]) for f in range(nf)]
    st = rb.alg.storage
    rew_rows = torch.zeros(T, n, pin_memory=True)
    done_rows = torch.zeros(T, n, dtype=st.dones.dtype, pin_memory=True)
    down = [_lib.CopyBatch([(rew_rows[i], st.rewards[i]), (done_rows[i], st.dones[i])]) for i in range(T)]
    side, back = torch.cuda.Stream(), torch.cuda.Stream()
    done_ev = [torch.cuda.Event() for _ in range(T)]
    up_ev = [torch.cuda.Event() for _ in range(T)]
    k0 = (int(rb.env.gym.calls_base.item()) + rb.env.gym.sim_calls) // rb.env.cfg.control.decimation

    def pre(i):
        c = torch.cuda.current_stream()
        if i == 0:
            side.wait_stream(c)
            back.wait_stream(c)
        else:
            side.wait_event(done_ev[i - 1])
        up[(k0 + i) % nf].enqueue(side)                   # same bytes the pool already holds: only ordering is at stake
        up_ev[i].record(side)
        c.wait_event(up_ev[i])

    def post(i):
        c = torch.cuda.current_stream()
        done_ev[i].record(c)
        back.wait_event(done_ev[i])
        down[i].enqueue(back)

    def join():
        c = torch.cuda.current_stream()
        c.wait_stream(side)
        c.wait_stream(back)

    g = RolloutGraph(rb).capture(pre, post, join)
    ca = eager_rollout(ra)
    cb = g.replay()[1]
    torch.cuda.synchronize()
    storage_equal(ra.alg.storage, rb.alg.storage)
    assert torch.equal(ca, cb)
    assert torch.equal(rew_rows, ra.alg.storage.rewards.view(T, n).cpu())
    assert torch.equal(done_rows, ra.alg.storage.dones.view(T, n).cpu())


def test_graph_captured_before_first_rollout_keeps_recurrent_state():
    """Capture on a FRESH recurrent runner (nothing has run yet): buffers allocated on first use -- the (h, c) rollout state
    -- must be created before the capture, or their zero-fill becomes a graph node and every replay restarts the LSTM
    from zero.  Two iterations: the second replay continues from the first one's hidden state exactly like the eager loop."""
    n = 128
    ra, rb = make_runner("g1", n), make_runner("g1", n)
    from unitree_rl_gym_b200.rsl_rl.runners.rollout_graph import RolloutGraph
    g = RolloutGraph(rb).capture()
    for _ in range(2):
        ca = eager_rollout(ra)
        cb = g.replay()[1]
        torch.cuda.synchronize()
        storage_equal(ra.alg.storage, rb.alg.storage)
        assert torch.equal(ca, cb)
        assert torch.equal(ra.alg.actor_critic._hidden, rb.alg.actor_critic._hidden)
        finish(ra, ca)
        finish(rb, cb)
    assert torch.equal(ra.alg.engine.flat, rb.alg.engine.flat)
