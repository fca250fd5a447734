"""ORACLE (test infrastructure, never on the product path): the CPU baseline leg of bench.py.

Times the CPU restatement of the hot path — oracle.env_port.EnvPort (LeggedRobot.step minus PhysX,
legged_robot.py:615-811) under oracle.rsl_rl_port's OnPolicyRunner / PPO (rsl_rl v1.0.2 restated,
rl_device=cpu as README.md:58-59 of the reference describes) — on the host cores of the box, for
bench.py's `cpu_baseline` object and its `--impl reference` arm.  kind = "port": the unmodified
reference is Python that imports isaacgym and lives in /root/reference, neither of which exists on
the GPU box, so the port (pinned to the unmodified reference by tests/golden/env_*.npz) stands in.

Only bench.py imports this module.
"""
import time

import torch

from . import env_port
from . import rsl_rl_port as R


class PortVecEnv:
    """VecEnv protocol (R1) over EnvPort with the synthetic physics seam (frame pool cycled like
    unitree_rl_gym_b200.physics.SyntheticGym: frame = (simulate calls - 1) // decimation % P)."""

    def __init__(self, spec, frames, device="cpu"):
        self.port = env_port.EnvPort(spec, device=device, fast_rng=True)   # torch.rand like the reference (timing only)
        self.spec = spec
        self.frames = frames
        self.num_frames = frames["root"].shape[0]
        self.sim_calls = 0
        self.num_envs, self.num_obs = spec.num_envs, spec.num_obs
        self.num_privileged_obs = spec.num_priv if spec.num_priv else None
        self.num_actions = spec.num_dof
        self.max_episode_length = spec.max_episode_length
        self.device = device

    def _physics(self, sub):
        if sub is not None:
            self.sim_calls += 1
        f = ((max(self.sim_calls, 1) - 1) // self.spec.decimation) % self.num_frames
        self.port.load_frame(self.frames, f, sub)

    # attributes rsl_rl touches
    @property
    def episode_length_buf(self):
        return self.port.episode_length_buf

    @episode_length_buf.setter
    def episode_length_buf(self, v):
        self.port.episode_length_buf = v

    @property
    def extras(self):
        return self.port.extras

    def get_observations(self):
        return self.port.obs_buf

    def get_privileged_observations(self):
        return self.port.privileged_obs_buf

    def reset(self):
        return self.port.reset(self._physics)

    def step(self, actions):
        return self.port.step(actions, self._physics)


def run(spec, frames, train_cfg, iterations, warmup, threads=None, device="cpu"):
    """Returns dict(rollout_s=[...], update_s=[...]) per timed iteration (one iteration = 24-step rollout +
    compute_returns + 5x4 minibatch update), after `warmup` untimed iterations.  device="cuda" runs the same eager
    torch code on the GPU (bench.py's `eager_gpu` context number: ~300 launches and 3 host syncs per env step)."""
    if threads:
        torch.set_num_threads(int(threads))
    dev = torch.device(device)
    if dev.type != "cpu":
        frames = {k: (v.to(dev) if v is not None else None) for k, v in frames.items()}
    out = {"rollout_s": [], "update_s": [], "iter_wall_s": [], "threads": torch.get_num_threads()}
    with torch.device(dev):                  # the port's factory calls (torch.zeros(...)) follow the run's device
        env = PortVecEnv(spec, frames, device=str(dev))
        torch.manual_seed(1)
        runner = R.OnPolicyRunner(env, train_cfg, log_dir=None, device=str(dev))
        env.episode_length_buf = torch.randint_like(env.episode_length_buf, high=int(env.max_episode_length))
        for it in range(warmup + iterations):
            if dev.type != "cpu":
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            runner.learn(1, init_at_random_ep_len=False)
            if dev.type != "cpu":
                torch.cuda.synchronize()
            st = runner.last_stats
            if it >= warmup:
                out["rollout_s"].append(st["collection_time"])
                out["update_s"].append(st["learn_time"])
                out["iter_wall_s"].append(time.perf_counter() - t0)
            out["last_iter_wall_s"] = time.perf_counter() - t0
    return out
