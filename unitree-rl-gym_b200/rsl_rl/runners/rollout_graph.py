"""One rollout (num_steps_per_env x [PPO.act -> env.step -> PPO.process_env_step]) captured once as a CUDA
graph and replayed per PPO iteration.

Everything in the loop is stream-ordered device work with no host read (the reference's three host syncs per
env step, SURVEY.md §7, are gone), and every per-step counter the kernels consume is (device base + host offset)
(include/urlg_b200.h conventions), so the loop of OnPolicyRunner.learn ([UPSTREAM rsl_rl v1.0.2]
runners/on_policy_runner.py) is capturable as-is.  The graph's last nodes advance the device
bases, so successive replays continue the env-step / act-step / physics-frame sequences exactly like the eager
loop would.

Requirements (checked): the env exposes begin_relative / advance_base / end_relative (the drop-in LeggedRobot
does) and num_steps_per_env is even (the obs ping-pong slot must return to its starting slot so the graph's
first act() reads the buffer the previous replay wrote last).
"""
import torch


def _baked_identity(runner):
    """What a captured graph holds raw pointers into: the policy engine (weights, workspaces) and the storage slabs."""
    alg = runner.alg
    return (alg.engine, alg.storage, alg.storage.observations.data_ptr(), getattr(runner.env, "_handle", None))


def _check_baked(runner, baked, who):
    alg = runner.alg
    now = (alg.engine, alg.storage, alg.storage.observations.data_ptr(), getattr(runner.env, "_handle", None))
    if any(a is not b if not isinstance(a, int) else a != b for a, b in zip(now, baked)):
        raise RuntimeError(who + ": the policy engine, rollout storage or env was rebuilt after capture (e.g. "
                           "actor_critic.to(), a new init_storage); capture a new graph")


def _materialize_lazy_state(runner):
    """Everything the loop allocates on first use must exist BEFORE capture: an allocation made inside the capture
    lands in the graph's private pool and its initialising kernel becomes a graph node, i.e. it would be re-run by
    every replay (the recurrent policy's (h, c) buffer zeroed at the start of every rollout)."""
    alg = runner.alg
    alg._check_engine()
    ac = alg.actor_critic
    if ac.is_recurrent:
        ac.hidden_buffer(alg.storage.num_envs)


class RolloutGraph:
    def __init__(self, runner):
        self.runner = runner
        self.graph = None
        self.launches = 0           # library kernels recorded in the graph (urlg_launch_count delta)

    @staticmethod
    def supported(runner):
        env = runner.env
        return (all(hasattr(env, m) for m in ("begin_relative", "advance_base", "end_relative"))
                and runner.num_steps_per_env % 2 == 0 and torch.device(runner.device).type == "cuda")

    def _loop(self, obs, critic_obs, pre_step=None, post_step=None):
        env, alg = self.runner.env, self.runner.alg
        for t in range(self.runner.num_steps_per_env):
            if pre_step is not None:
                pre_step(t)
            actions = alg.act(obs, critic_obs)
            obs, privileged_obs, rewards, dones, infos = env.step(actions)
            critic_obs = privileged_obs if privileged_obs is not None else obs
            alg.process_env_step(rewards, dones, infos)
            self.last = (rewards, dones, infos)      # what post_step hooks of this step may consume (device tensors)
            if post_step is not None:
                post_step(t)
        return obs, critic_obs

    def capture(self, pre_step=None, post_step=None, epilogue=None):
        """pre_step(t) / post_step(t) / epilogue() run INSIDE the capture, around env step t and after the last one: a
        caller whose physics state lives elsewhere records its own stream-ordered work there (e.g. the upload of step
        t + 1's state frame on a forked copy stream, the read-back of step t's rewards), so the whole rollout including
        that traffic stays one graph launch.  Forked streams must be joined back (in `epilogue`) before capture ends."""
        from ... import _lib
        r = self.runner
        env, alg = r.env, r.alg
        if not self.supported(r):
            raise RuntimeError("RolloutGraph: env / runner configuration is not capturable")
        if alg.storage.step != 0:
            raise RuntimeError("RolloutGraph.capture: storage must be empty (call after update())")
        obs = env.get_observations()
        priv = env.get_privileged_observations()
        critic_obs = priv if priv is not None else obs
        slot0 = env._obs_slot
        _materialize_lazy_state(r)
        torch.cuda.synchronize()
        env.begin_relative()
        alg.begin_relative()
        g = torch.cuda.CUDAGraph()
        n0 = _lib.lib().urlg_launch_count()
        with torch.inference_mode(), torch.cuda.graph(g):
            obs_out, critic_out = self._loop(obs, critic_obs, pre_step, post_step)
            env.advance_base()
            alg.advance_base()
            if epilogue is not None:
                epilogue()
        self.launches = _lib.lib().urlg_launch_count() - n0
        env.end_relative()
        alg.end_relative()
        alg.storage.clear()         # capture recorded work but executed nothing
        assert env._obs_slot == slot0
        self.graph = g
        self.obs_out, self.critic_out = obs_out, critic_out
        self._baked = _baked_identity(r)
        return self

    def replay(self):
        """Run one rollout; returns (obs, critic_obs) after the last step (inputs of compute_returns)."""
        _check_baked(self.runner, self._baked, "RolloutGraph")
        self.graph.replay()
        self.runner.alg.storage.step = self.runner.num_steps_per_env
        return self.obs_out, self.critic_out


class StepGraphs:
    """One CUDA graph per rollout position t = 0..T-1 (PPO.act -> env.step -> PPO.process_env_step writing storage slab
    t), for callers whose physics state arrives from outside between env steps (host-resident or a separate simulator):
    they refresh the physics tensors, call `replay(t)`, and read rewards / dones, one launch per env step instead of ~25.
    Positions must be replayed in order 0..T-1 (the obs ping-pong slot and the storage slab are baked per position)."""

    def __init__(self, runner):
        self.runner = runner
        self.graphs = []
        self.launches = 0

    def capture(self):
        from ... import _lib
        r = self.runner
        env, alg = r.env, r.alg
        if not RolloutGraph.supported(r):
            raise RuntimeError("StepGraphs: env / runner configuration is not capturable")
        if alg.storage.step != 0:
            raise RuntimeError("StepGraphs.capture: storage must be empty (call after update())")
        slot0 = env._obs_slot
        _materialize_lazy_state(r)
        pool = torch.cuda.graph_pool_handle()
        self.outs = []
        n0 = _lib.lib().urlg_launch_count()
        torch.cuda.synchronize()
        for t in range(r.num_steps_per_env):
            obs = env.get_observations()
            priv = env.get_privileged_observations()
            critic_obs = priv if priv is not None else obs
            env.begin_relative()
            alg.begin_relative()
            g = torch.cuda.CUDAGraph()
            with torch.inference_mode(), torch.cuda.graph(g, pool=pool):
                actions = alg.act(obs, critic_obs)
                obs2, priv2, rew, dones, infos = env.step(actions)
                alg.process_env_step(rew, dones, infos)
                env.advance_base()
                alg.advance_base()
            env.end_relative()
            alg.end_relative()
            self.graphs.append(g)
            self.outs.append((obs2, priv2 if priv2 is not None else obs2, rew, dones))
        self.launches = (_lib.lib().urlg_launch_count() - n0) // max(len(self.graphs), 1)
        alg.storage.clear()
        assert env._obs_slot == slot0
        self._baked = _baked_identity(r)
        return self

    def replay(self, t):
        """Run env step t of the rollout; returns (obs, critic_obs, rewards, dones) device tensors of that step."""
        _check_baked(self.runner, self._baked, "StepGraphs")
        self.graphs[t].replay()
        self.runner.alg.storage.step = t + 1
        return self.outs[t]
