"""ORACLE (test infrastructure).  Runs the UNMODIFIED reference env classes from
/root/reference on CPU under the stub isaacgym, fed with synthetic physics
frames and the shared counter-based uniform table (oracle/philox.py).

Only usable in the build container (needs /root/reference); used by
tests/golden/gen_golden.py to produce committed fixtures and by the container-only
tests that pin oracle/env_port.py against the real reference code.

RNG slot map per (env, step) — shared with oracle/env_port.py and
unitree-rl-gym_b200/csrc/env_kernels.cu (A = num_dof, O = num_obs):
    0..2          callback _resample_commands  (x, y, heading|yaw)  legged_robot.py:511-512,530-535
    3..3+A-1      _reset_dofs                                       :566
    3+A..8+A      _reset_root_states (6 velocities)                 :590
    9+A..11+A     reset_idx -> _resample_commands                   :749
    12+A,13+A     _push_robots                                      :549
    14+A..14+A+O-1  observation noise                               :811, h1_env.py:95
"""
import os
import sys
import types

import numpy as np
import torch

REF_ROOT = "/root/reference"
_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "legged_gym"))


def _install_paths():
    for p in (_REPO, _HERE, REF_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)


def import_reference():
    """Import legged_gym from /root/reference with stub isaacgym + oracle rsl_rl."""
    _install_paths()
    import isaacgym  # noqa: F401  (the stub, found via _HERE on sys.path)
    assert isaacgym.__file__.startswith(_HERE), "a real isaacgym shadows the stub"
    # envs must be imported before utils (circular import otherwise)
    import legged_gym.envs as envs  # noqa: F401
    import legged_gym
    assert legged_gym.__file__.startswith(REF_ROOT)
    return envs


def num_rng_slots(num_dof, num_obs):
    return 14 + num_dof + num_obs


class RefRun:
    """One reference env instance + its fake gym + RNG plumbing."""

    def __init__(self, task, num_envs, seed=1, env_offset=0, num_frames=4, frames=None, frame_seed=1,
                 add_noise=True, push=True, cfg_hook=None):
        from oracle import philox
        envs = import_reference()
        from isaacgym import gymapi, torch_utils
        from legged_gym.utils.task_registry import task_registry
        from legged_gym.utils.helpers import class_to_dict
        sys.path.insert(0, _REPO)
        from unitree_rl_gym_b200.robots import robot_for_asset
        from unitree_rl_gym_b200 import synthetic

        self.task = task
        self.torch_utils = torch_utils
        env_cfg, train_cfg = task_registry.get_cfgs(task)
        import copy
        env_cfg = copy.deepcopy(env_cfg)      # the registry hands out one shared instance per task; overrides stay local
        env_cfg.env.num_envs = num_envs
        env_cfg.noise.add_noise = add_noise
        env_cfg.domain_rand.push_robots = push
        if cfg_hook is not None:          # config overrides applied before construction (same hook as tests/helpers.py)
            cfg_hook(env_cfg)
        self.env_cfg, self.train_cfg = env_cfg, train_cfg
        robot = robot_for_asset(env_cfg.asset.file)
        self.robot = robot

        gym = gymapi._new_gym()
        gym.robot = robot
        gym.decimation = env_cfg.control.decimation
        self.gym = gym

        sim_params = gymapi.SimParams()
        from isaacgym import gymutil
        gymutil.parse_sim_config(class_to_dict(env_cfg.sim), sim_params)
        sim_params.use_gpu_pipeline = False

        torch.manual_seed(seed)
        np.random.seed(seed)
        cls = task_registry.get_task_class(task)
        self.env = cls(cfg=env_cfg, sim_params=sim_params, physics_engine=gymapi.SIM_PHYSX,
                       sim_device="cpu", headless=True)
        env = self.env
        A, O = env.num_dof, env.num_obs
        self.S = num_rng_slots(A, O)
        self.seed, self.env_offset = seed, env_offset

        if frames is None:
            frames = synthetic.make_frames(
                robot, None, num_envs, num_frames, env_cfg.control.decimation, seed=frame_seed,
                feet_idx=env.feet_indices.tolist(), pen_idx=env.penalised_contact_indices.tolist(),
                term_idx=env.termination_contact_indices.tolist(), init_z=env_cfg.init_state.pos[2],
                default_dof_pos=env.default_dof_pos[0].clone(), env_spacing=env_cfg.env.env_spacing)
        self.frames = frames
        gym.frames = frames

        # ---- RNG redirection: reference code unchanged, only the source of uniforms ----
        rng = torch_utils.RNG
        cache = {}

        def table(step):
            if step not in cache:
                cache.clear()
                ids = np.arange(num_envs, dtype=np.int64) + env_offset
                cache[step] = torch.from_numpy(
                    philox.uniform_table(seed, philox.STREAM_ENV, ids, step, self.S).copy())
            return cache[step]

        rng.table_fn = table
        self._rng = rng
        all_ids = torch.arange(num_envs)
        state = {"in_reset": False}

        def pre(name, fn):
            orig = getattr(env, name)

            def wrapped(*a, **k):
                fn(*a, **k)
                return orig(*a, **k)
            setattr(env, name, wrapped)

        def pre_resample(env_ids):
            rng.step = env.common_step_counter
            rng.env_ids = env_ids
            rng.slot = (9 + A) if state["in_reset"] else 0

        def pre_reset_dofs(env_ids):
            rng.step = env.common_step_counter
            rng.env_ids = env_ids
            rng.slot = 3

        def pre_reset_root(env_ids):
            rng.step = env.common_step_counter
            rng.env_ids = env_ids
            rng.slot = 3 + A

        def pre_push():
            rng.step = env.common_step_counter
            rng.env_ids = all_ids
            rng.slot = 12 + A

        pre("_resample_commands", pre_resample)
        pre("_reset_dofs", pre_reset_dofs)
        pre("_reset_root_states", pre_reset_root)
        pre("_push_robots", pre_push)
        orig_reset_idx = env.reset_idx

        def reset_idx(env_ids):
            state["in_reset"] = True
            try:
                return orig_reset_idx(env_ids)
            finally:
                state["in_reset"] = False
        env.reset_idx = reset_idx

        orig_obs = env.compute_observations
        real_rand_like = torch.rand_like

        def compute_observations():
            def fake_rand_like(t, *a, **k):
                tab = table(env.common_step_counter)
                return tab[:, 14 + A:14 + A + t.shape[1]].to(t.dtype)
            torch.rand_like = fake_rand_like
            try:
                return orig_obs()
            finally:
                torch.rand_like = real_rand_like
        env.compute_observations = compute_observations

    def close(self):
        self._rng.table_fn = None
        self._rng.env_ids = None

    # ---- snapshot of every public buffer the drop-in must reproduce ----
    def snapshot(self):
        e = self.env
        out = {
            "obs_buf": e.obs_buf, "rew_buf": e.rew_buf, "reset_buf": e.reset_buf.to(torch.bool),
            "time_out_buf": e.time_out_buf, "episode_length_buf": e.episode_length_buf,
            "torques": e.torques, "commands": e.commands, "actions": e.actions,
            "last_actions": e.last_actions, "last_dof_vel": e.last_dof_vel,
            "last_root_vel": e.last_root_vel, "feet_air_time": e.feet_air_time,
            "last_contacts": e.last_contacts, "rpy": e.rpy, "base_lin_vel": e.base_lin_vel,
            "base_ang_vel": e.base_ang_vel, "projected_gravity": e.projected_gravity,
            "root_states": e.root_states, "dof_state": e.dof_state,
        }
        if e.privileged_obs_buf is not None:
            out["privileged_obs_buf"] = e.privileged_obs_buf
        for k, v in e.episode_sums.items():
            out["episode_sums/" + k] = v
        if "episode" in e.extras:
            for k, v in e.extras["episode"].items():
                out["extras_episode/" + k] = torch.as_tensor(v).reshape(1)
        if "time_outs" in e.extras:
            out["extras_time_outs"] = e.extras["time_outs"]
        return {k: v.detach().clone() for k, v in out.items()}
