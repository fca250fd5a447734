"""Generates tests/golden/env_<case>.npz by running the UNMODIFIED reference env classes
(/root/reference/legged_gym) under oracle/refharness, and tests/golden/motion_pt.npz from the
shipped TorchScript policies.  Container-only (needs /root/reference); fixtures are committed.

    python tests/golden/gen_golden.py

Each env fixture holds: the synthetic physics frames, the per-step actions, the planted
episode lengths, and after reset() and after every step a snapshot of every public buffer.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)

from oracle.refharness.run_reference import RefRun  # noqa: E402
from tests.helpers import TWEAKS  # noqa: E402

CASES = {
    # name: (task, num_envs, only_positive_rewards, steps)
    "go2_n64": ("go2", 64, True, 6),
    "go2_n64_signed": ("go2", 64, False, 6),
    "h1_n40": ("h1", 40, False, 6),
    "g1_n64": ("g1", 64, True, 6),
    "h1_2_n33": ("h1_2", 33, False, 5),
    # reward terms / control types no shipped task enables (config overrides in tests/helpers.py TWEAKS)
    "go2_n48_allterms": ("go2", 48, False, 6),
    "h1_2_n35_allterms": ("h1_2", 35, True, 5),
    "go2_n40_ctrlV": ("go2", 40, True, 4),
    "go2_n40_ctrlT": ("go2", 40, True, 4),
    # SURVEY.md 8(f) row 1 ("next"): the unmodified GO2HandstandEnv on the same synthetic frames.  Reference vectors for
    # the round that ports the task (46-dim obs with contact flags, gravity-based termination, 13 extra reward terms,
    # termination = -5); no product code consumes this fixture yet (tests only check it is well-formed).
    "go2_handstand_n64": ("go2_handstand", 64, False, 6),
}
SEED = 1
NUM_FRAMES = 4


def plant_episode_lengths(n, max_len, resample, push, gen):
    ep = torch.randint(0, int(max_len), (n,), generator=gen)
    special = [resample - 1, resample - 2, int(max_len), int(max_len) - 1, push - 1, push - 2, 0, 2 * push - 1]
    for i, v in enumerate(special):
        if i < n:
            ep[i] = v
    return ep


def run_case(name):
    task, n, only_pos, steps = CASES[name]
    def hook(cfg):
        cfg.rewards.only_positive_rewards = only_pos
        if name in TWEAKS:
            TWEAKS[name](cfg)
    r = RefRun(task, n, seed=SEED, num_frames=NUM_FRAMES, cfg_hook=hook)
    env = r.env
    out = {}
    for k, v in r.frames.items():
        out["frames/" + k] = v.numpy()
    gen = torch.Generator().manual_seed(1234)
    env.reset()
    for k, v in r.snapshot().items():
        out[f"reset/{k}"] = v.numpy()
    ep = plant_episode_lengths(n, env.max_episode_length, int(env.cfg.commands.resampling_time / env.dt),
                               int(env.cfg.domain_rand.push_interval), gen)
    env.episode_length_buf[:] = ep
    out["planted_episode_length"] = ep.numpy()
    for t in range(steps):
        a = torch.randn(n, env.num_actions, generator=gen) * (0.3 if t % 2 else 1.0)
        if t == 2:
            a[0] = 250.0      # exercises clip_actions
        out[f"actions/{t}"] = a.numpy()
        env.step(a.clone())
        for k, v in r.snapshot().items():
            out[f"step{t}/{k}"] = v.numpy()
    # constants the drop-in must derive identically from the config
    out["const/p_gains"] = env.p_gains.numpy()
    out["const/d_gains"] = env.d_gains.numpy()
    out["const/default_dof_pos"] = env.default_dof_pos.numpy()
    out["const/torque_limits"] = env.torque_limits.numpy()
    out["const/dof_vel_limits"] = env.dof_vel_limits.numpy()
    out["const/dof_pos_limits"] = env.dof_pos_limits.numpy()
    out["const/noise_scale_vec"] = env.noise_scale_vec.numpy()
    out["const/feet_indices"] = env.feet_indices.numpy()
    out["const/penalised_contact_indices"] = env.penalised_contact_indices.numpy()
    out["const/termination_contact_indices"] = env.termination_contact_indices.numpy()
    out["const/env_origins"] = env.env_origins.numpy()
    out["const/base_init_state"] = env.base_init_state.numpy()
    if hasattr(env, "rear_dof_idx"):          # go2_handstand: constants of its PD / pose overrides
        out["const/handstand/rear_dof_idx"] = env.rear_dof_idx.numpy()
        out["const/handstand/rear_center"] = env.rear_center.numpy()
        out["const/handstand/rear_limit"] = np.array(env.rear_limit, dtype=np.float64)
        out["const/handstand/pose_targets"] = env.pose_targets[0].numpy()
        out["const/handstand/base_height_target"] = np.array(env.cfg.rewards.base_height_target, dtype=np.float64)
        out["const/handstand/action_scale_clip"] = np.array([env.cfg.control.action_scale,
                                                             env.cfg.normalization.clip_actions], dtype=np.float64)
    out["const/reward_names"] = np.array(env.reward_names)
    out["const/reward_scales"] = np.array([env.reward_scales[k] for k in env.reward_names], dtype=np.float64)
    out["const/scalars"] = np.array([env.dt, env.max_episode_length, env.cfg.domain_rand.push_interval,
                                     int(env.cfg.commands.resampling_time / env.dt), steps, NUM_FRAMES, SEED],
                                    dtype=np.float64)
    r.close()
    np.savez_compressed(os.path.join(HERE, f"env_{name}.npz"), **out)
    resets = [int(out[f"step{t}/reset_buf"].sum()) for t in range(steps)]
    print(name, "resets per step", resets, "rew[0:3]", out[f"step{steps-1}/rew_buf"][:3])


def motion_pt():
    """Known-answer vectors for the LSTM + MLP actor; also stores the weights so the
    fixture travels without /root/reference."""
    out = {}
    for robot in ("g1", "h1", "h1_2"):
        m = torch.jit.load(f"/root/reference/deploy/pre_train/{robot}/motion.pt")
        sd = {k: v.clone() for k, v in m.state_dict().items()}
        for k, v in sd.items():
            out[f"{robot}/sd/{k}"] = v.numpy()
        n_in = sd["memory.weight_ih_l0"].shape[1]
        m.reset_memory()
        z = torch.zeros(1, n_in)
        out[f"{robot}/zeros_call1"] = m(z).detach().numpy()
        out[f"{robot}/zeros_call2"] = m(z).detach().numpy()
        m.reset_memory()
        x = torch.rand(8, 1, n_in, generator=torch.Generator().manual_seed(0)) * 2 - 1
        ys = []
        for t in range(8):
            ys.append(m(x[t]).detach().numpy())
        out[f"{robot}/seq_in"] = x.numpy()
        out[f"{robot}/seq_out"] = np.stack(ys)
        print(robot, out[f"{robot}/zeros_call1"][0, :4])
    np.savez_compressed(os.path.join(HERE, "motion_pt.npz"), **out)


if __name__ == "__main__":
    only = sys.argv[1:]                    # optional: case names (and/or "motion_pt") to (re)generate
    for c in CASES:
        if not only or c in only:
            run_case(c)
    if not only or "motion_pt" in only:
        motion_pt()
