"""Drop-in LeggedRobot whose step-time work runs in the fused sm_100a kernels of liburlg_b200.

Public surface (constructor signature, VecEnv protocol, every public tensor name / shape / dtype)
follows legged_gym/envs/base/legged_robot.py:24-49, 615-647 and base_task.py:31-86.  What differs
by design: check_termination / compute_reward / compute_observations / _post_physics_step_callback /
_push_robots / _resample_commands are not separate eager passes any more — they are stages of one
kernel (csrc/env_kernels.cu).  What the reference's own subclasses change by overriding them (H1Robot,
GO2HandstandEnv) is a *variant* of that kernel selected by the class's `_family`; a subclass that overrides
a fused stage (or defines a `_reward_<name>`) with Python code of its own is rejected at construction
instead of being silently ignored.  Host syncs of the reference (`nonzero` x2, boolean-mask `len`) are gone;
the indexed setters of the physics seam (:552-555, 569-572, 591-594) get device-resident id lists + counts.
"""
import ctypes as C

import numpy as np
import torch

from .base_task import BaseTask
from .legged_robot_config import LeggedRobotCfg
from .... import _lib
from ....envspec import build_env_spec, term_id, FAMILY_BASE, FAMILY_H1
from ....physics import SyntheticGym
from ....robots import robot_for_asset

_FUSED_HOOKS = ("check_termination", "compute_reward", "compute_observations", "_post_physics_step_callback",
                "_push_robots", "_resample_commands", "_reset_dofs", "_reset_root_states")


class LeggedRobot(BaseTask):
    _family = FAMILY_BASE

    def __init__(self, cfg: LeggedRobotCfg, sim_params, physics_engine, sim_device, headless):
        self.cfg = cfg
        self.sim_params = sim_params
        self.height_samples = None
        self.debug_viz = False
        self.init_done = False
        for name in _FUSED_HOOKS:
            if getattr(type(self), name) is not getattr(LeggedRobot, name):
                raise NotImplementedError(
                    f"{type(self).__name__}.{name} overrides a stage that is fused into urlg_post_physics; Python code "
                    "cannot run inside the kernel — add the behaviour as a kernel variant / reward term "
                    "(csrc/env_kernels.cu, like GO2HandstandEnv's) instead")
        for name in dir(type(self)):
            if name.startswith("_reward_") and callable(getattr(type(self), name, None)):
                raise NotImplementedError(
                    f"{type(self).__name__}.{name} is a Python reward function; rewards are terms of the fused step "
                    "kernel (urlg_reward_term in include/urlg_b200.h) — add it there")
        self._parse_cfg(self.cfg)
        super().__init__(self.cfg, sim_params, physics_engine, sim_device, headless)
        self._init_buffers()
        self._prepare_reward_function()
        self._bind()
        self.init_done = True

    # ------------------------------------------------------------------ init-time (host)
    def _parse_cfg(self, cfg):
        from ...utils.helpers import class_to_dict
        self.robot = robot_for_asset(cfg.asset.file)
        seed = getattr(cfg, "seed", 1)
        self.spec = build_env_spec(getattr(cfg.asset, "name", "robot"), cfg, self.robot, self.sim_params.dt,
                                   seed=seed if seed is not None and seed >= 0 else 1,
                                   env_offset=getattr(cfg.env, "env_offset", 0), family=self._family)
        s = self.spec
        self.dt = s.dt
        self.obs_scales = cfg.normalization.obs_scales
        self.reward_scales = class_to_dict(cfg.rewards.scales)
        self.command_ranges = class_to_dict(cfg.commands.ranges)
        self.max_episode_length_s = s.max_episode_length_s
        self.max_episode_length = np.float64(s.max_episode_length)
        cfg.domain_rand.push_interval = np.float64(s.push_interval)

    def create_sim(self):
        self.up_axis_idx = 2
        pe = self.physics_engine
        if pe is not None and hasattr(pe, "acquire_dof_state_tensor"):
            self.gym = pe
        else:
            self.gym = SyntheticGym(self.robot, self.num_envs, self.cfg.control.decimation, self.device)
        self.sim = None
        s = self.spec
        self.num_dof = self.num_dofs = s.num_dof
        self.num_bodies = s.num_bodies
        self.dof_names = list(s.dof_names)
        dev = self.device
        t = lambda x, dt=torch.float: torch.as_tensor(np.asarray(x), dtype=dt, device=dev)
        self.dof_pos_limits = t(s.dof_pos_limits)
        self.dof_vel_limits = t(s.dof_vel_limits)
        self.torque_limits = t(s.torque_limits)
        self.feet_indices = t(s.feet_idx, torch.long)
        self.penalised_contact_indices = t(s.pen_idx, torch.long)
        self.termination_contact_indices = t(s.term_idx, torch.long)
        self.base_init_state = t(s.base_init_state)
        self.custom_origins = False
        self.env_origins = t(s.env_origins)

    def _init_buffers(self):
        s, dev, N, A = self.spec, self.device, self.num_envs, self.num_dof
        f = dict(dtype=torch.float, device=dev, requires_grad=False)
        g = self.gym
        self.root_states = g.acquire_actor_root_state_tensor(self.sim)
        self.dof_state = g.acquire_dof_state_tensor(self.sim)
        self.dof_pos = self.dof_state.view(N, A, 2)[..., 0]
        self.dof_vel = self.dof_state.view(N, A, 2)[..., 1]
        self.base_quat = self.root_states[:, 3:7]
        self.base_pos = self.root_states[:N, 0:3]
        self.contact_forces = g.acquire_net_contact_force_tensor(self.sim).view(N, -1, 3)
        self.rigid_body_states = None
        if self._family != FAMILY_BASE:
            self.rigid_body_states = g.acquire_rigid_body_state_tensor(self.sim)
            self.rigid_body_states_view = self.rigid_body_states.view(N, -1, 13)
            self.feet_num = len(s.feet_idx)
        if self._family == FAMILY_H1:
            self.phase = torch.zeros(N, **f)
            self.leg_phase = torch.zeros(N, 2, **f)
        self.common_step_counter = 0
        self._step_base = torch.zeros(1, dtype=torch.int64, device=dev)
        self.extras = {}
        self.noise_scale_vec = self._get_noise_scale_vec(self.cfg)
        self.gravity_vec = torch.tensor([0., 0., -1.], **f).repeat((N, 1))
        self.forward_vec = torch.tensor([1., 0., 0.], **f).repeat((N, 1))
        self.torques = torch.zeros(N, self.num_actions, **f)
        self.p_gains = torch.tensor(s.p_gains, **f)
        self.d_gains = torch.tensor(s.d_gains, **f)
        self.actions = torch.zeros(N, self.num_actions, **f)
        self.last_actions = torch.zeros(N, self.num_actions, **f)
        self.last_dof_vel = torch.zeros(N, A, **f)
        self.last_root_vel = torch.zeros(N, 6, **f)
        self.commands = torch.zeros(N, self.cfg.commands.num_commands, **f)
        self.commands_scale = torch.tensor([self.obs_scales.lin_vel, self.obs_scales.lin_vel, self.obs_scales.ang_vel], **f)
        self.feet_air_time = torch.zeros(N, len(s.feet_idx), **f)
        self.last_contacts = torch.zeros(N, len(s.feet_idx), dtype=torch.bool, device=dev)
        self.rpy = torch.zeros(N, 3, **f)
        self.base_lin_vel = torch.zeros(N, 3, **f)
        self.base_ang_vel = torch.zeros(N, 3, **f)
        self.projected_gravity = torch.zeros(N, 3, **f)
        self.projected_gravity[:, 2] = -1.0
        self.default_dof_pos = torch.tensor(s.default_dof_pos, **f).unsqueeze(0)

    def _get_noise_scale_vec(self, cfg):
        self.add_noise = cfg.noise.add_noise
        return torch.as_tensor(self.spec.noise_scale_vec, dtype=torch.float, device=self.device)

    def _prepare_reward_function(self):
        s = self.spec
        for key in list(self.reward_scales.keys()):        # same bookkeeping as :822-827
            if self.reward_scales[key] == 0:
                self.reward_scales.pop(key)
            else:
                self.reward_scales[key] *= self.dt
        self.reward_names = list(s.reward_names)
        names = s.sum_names
        self._episode_sums_buf = torch.zeros(len(names), self.num_envs, dtype=torch.float, device=self.device)
        self.episode_sums = {n: self._episode_sums_buf[i] for i, n in enumerate(names)}
        self._extras_episode = torch.zeros(len(names), dtype=torch.float, device=self.device)
        self._extras_time_outs = torch.zeros(self.num_envs, dtype=torch.bool, device=self.device)
        # compacted int32 id lists + device counts for the indexed setters of the physics seam (:552-555, 569-572, 591-594)
        self._reset_ids = torch.zeros(self.num_envs, dtype=torch.int32, device=self.device)
        self._reset_count = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._root_ids = torch.zeros(self.num_envs, dtype=torch.int32, device=self.device)
        self._root_count = torch.zeros(1, dtype=torch.int32, device=self.device)

    # ------------------------------------------------------------------ C-ABI binding
    def _desc(self):
        s = self.spec
        d = _lib.EnvDesc()
        d.family, d.num_envs, d.num_dof, d.num_bodies = s.family, s.num_envs, s.num_dof, s.num_bodies
        d.num_obs, d.num_priv = s.num_obs, s.num_priv
        d.num_feet, d.num_pen, d.num_term_bodies, d.num_hip = len(s.feet_idx), len(s.pen_idx), len(s.term_idx), len(s.hip_idx)
        d.decimation, d.control_type = s.decimation, ord(str(s.control_type)[0]) if s.control_type else 0
        if s.control_type not in ("P", "V", "T"):
            raise NameError(f"Unknown controller type: {s.control_type}")
        d.resample_steps, d.push_interval = s.resample_steps, s.push_interval
        d.push_robots, d.heading_command, d.add_noise = int(s.push_robots), int(s.heading_command), int(s.add_noise)
        d.only_positive, d.send_timeouts = int(s.only_positive), int(s.send_timeouts)
        names = s.sum_names
        d.num_terms, d.num_sums = len(s.reward_names), len(names)
        d.has_termination = int(s.termination_scale is not None)
        d.termination_sum_row = names.index("termination") if s.termination_scale is not None else -1
        d.termination_scale = s.termination_scale or 0.0
        for i, (n, sc) in enumerate(zip(s.reward_names, s.reward_scales)):
            d.term_ids[i], d.term_sum_row[i], d.term_scales[i] = term_id(n, s.family), names.index(n), sc
        for k in ("dt", "sim_dt", "max_episode_length", "max_episode_length_s", "action_scale", "clip_actions",
                  "tracking_sigma", "base_height_target", "soft_dof_vel_limit", "soft_torque_limit",
                  "max_contact_force", "obs_scale_lin_vel", "obs_scale_ang_vel", "obs_scale_dof_pos", "obs_scale_dof_vel"):
            setattr(d, k, float(getattr(s, k)))
        d.max_push_vel, d.clip_obs = s.max_push_vel, s.clip_obs
        for k in ("cmd_x", "cmd_y", "cmd_yaw", "cmd_heading"):
            v = getattr(s, k)
            getattr(d, k)[0], getattr(d, k)[1] = v[0], v[1]
        for j in range(s.num_dof):
            d.p_gains[j], d.d_gains[j], d.default_dof_pos[j] = s.p_gains[j], s.d_gains[j], s.default_dof_pos[j]
            d.torque_limits[j], d.dof_vel_limits[j] = s.torque_limits[j], s.dof_vel_limits[j]
            d.dof_pos_lo[j], d.dof_pos_hi[j] = s.dof_pos_limits[j, 0], s.dof_pos_limits[j, 1]
        for dst, src in ((d.feet_idx, s.feet_idx), (d.pen_idx, s.pen_idx), (d.term_idx, s.term_idx), (d.hip_idx, s.hip_idx)):
            if len(src) > _lib.MAX_IDX:
                raise ValueError("index list longer than URLG_MAX_IDX")
            for i, v in enumerate(src):
                dst[i] = v
        for i, v in enumerate(s.noise_scale_vec):
            d.noise_scale_vec[i] = v
        for i, v in enumerate(s.base_init_state):
            d.base_init_state[i] = v
        d.seed_lo, d.seed_hi, d.env_offset = s.seed & 0xFFFFFFFF, (s.seed >> 32) & 0xFFFFFFFF, s.env_offset
        d.termination_mode, d.callback_mode, d.clip_torques = s.termination_mode, s.callback_mode, int(s.clip_torques)
        d.num_front_dof = s.num_front_dof
        for j in range(s.num_dof):
            d.pd_target_lo[j], d.pd_target_hi[j], d.pose_targets[j] = s.pd_target_lo[j], s.pd_target_hi[j], s.pose_targets[j]
        return d

    def _bind(self, rng_override=None):
        L = _lib.lib()
        p = _lib.ptr
        if getattr(self, "_handle", None) is None:
            h = C.c_void_p()
            _lib.check(L.urlg_env_create(C.byref(self._desc()), C.byref(h)))
            self._handle = h
        b = _lib.EnvBuffers()
        b.root_states, b.dof_state = p(self.root_states), p(self.dof_state)
        b.contact_forces = p(self.contact_forces)
        b.rigid_body_states = p(self.rigid_body_states)
        for i in range(2):
            b.obs_buf[i] = self._obs_slots[i].data_ptr()
            b.priv_obs_buf[i] = self._priv_slots[i].data_ptr() if self._priv_slots else None
        b.rew_buf, b.reset_buf, b.time_out_buf = p(self.rew_buf), p(self.reset_buf), p(self.time_out_buf)
        b.episode_length_buf = p(self._episode_length_buf)
        b.torques, b.actions, b.last_actions = p(self.torques), p(self.actions), p(self.last_actions)
        b.last_dof_vel, b.last_root_vel = p(self.last_dof_vel), p(self.last_root_vel)
        b.commands, b.feet_air_time, b.last_contacts = p(self.commands), p(self.feet_air_time), p(self.last_contacts)
        b.rpy, b.base_lin_vel, b.base_ang_vel = p(self.rpy), p(self.base_lin_vel), p(self.base_ang_vel)
        b.projected_gravity = p(self.projected_gravity)
        if self._family == FAMILY_H1:
            b.phase, b.leg_phase = p(self.phase), p(self.leg_phase)
        b.episode_sums, b.env_origins = p(self._episode_sums_buf), p(self.env_origins)
        b.extras_episode, b.extras_time_outs = p(self._extras_episode), p(self._extras_time_outs)
        b.reset_ids, b.reset_count = p(self._reset_ids), p(self._reset_count)
        b.root_ids, b.root_count = p(self._root_ids), p(self._root_count)
        self._rng_override = rng_override
        b.rng_override = p(rng_override)
        _lib.check(L.urlg_env_bind(self._handle, C.byref(b)))
        # pool-indexed physics seam (SyntheticGym(pool_seam=True)): the kernels read the current synthetic frame in place
        self._pool = bool(getattr(self.gym, "pool_seam", False))
        if self._pool:
            self._pool_desc = self.gym.frame_pool_desc()
            _lib.check(L.urlg_env_bind_frame_pool(self._handle, C.byref(self._pool_desc)))

    def __del__(self):
        h = getattr(self, "_handle", None)
        if h is not None:
            try:
                _lib.lib().urlg_env_destroy(h)
            except Exception:
                pass
            self._handle = None

    # ------------------------------------------------------------------ step-time (device)
    def step(self, actions):
        """legged_robot.py:615-647: clip actions, decimation x (PD torques -> physics), fused post-physics."""
        if actions.device != self.actions.device or actions.dtype != torch.float or not actions.is_contiguous():
            actions = actions.to(device=self.device, dtype=torch.float).contiguous()
        L, st, g = _lib.lib(), _lib.cur_stream(), self.gym
        raw = _lib.ptr(actions)
        for k in range(self.cfg.control.decimation):
            if self._pool:
                _lib.check(L.urlg_pd_torques_at(self._handle, raw if k == 0 else None, g.sim_calls, st))
            else:
                _lib.check(L.urlg_pd_torques(self._handle, raw if k == 0 else None, st))
            g.set_dof_actuation_force_tensor(self.sim, self.torques)
            g.simulate(self.sim)
            g.refresh_dof_state_tensor(self.sim)
        self.post_physics_step()
        return self.obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf, self.extras

    def _compute_torques(self, actions):
        """K1 on its own (legged_robot.py:649-671); `actions` are clipped and stored like step() does."""
        actions = actions.to(device=self.device, dtype=torch.float).contiguous()
        if self._pool:
            _lib.check(_lib.lib().urlg_pd_torques_at(self._handle, _lib.ptr(actions), self.gym.sim_calls, _lib.cur_stream()))
        else:
            _lib.check(_lib.lib().urlg_pd_torques(self._handle, _lib.ptr(actions), _lib.cur_stream()))
        return self.torques

    def post_physics_step(self):
        """legged_robot.py:673-709 as one fused kernel + finalize (csrc/env_kernels.cu K2/K2b), then the commits the
        reference makes through the indexed setters: reset envs' DOF state (:569-572), reset + pushed envs' root state
        (:591-594, 552-555) — as device-resident int32 id lists with device counts, so nothing syncs."""
        g = self.gym
        g.refresh_actor_root_state_tensor(self.sim)
        g.refresh_net_contact_force_tensor(self.sim)
        if self._family != FAMILY_BASE:
            g.refresh_rigid_body_state_tensor(self.sim)
        self.common_step_counter += 1
        self._obs_slot ^= 1
        self._launch_post()
        self._commit_indexed()
        self.obs_buf = self._obs_slots[self._obs_slot]
        if self._priv_slots:
            self.privileged_obs_buf = self._priv_slots[self._obs_slot]
        self._publish_extras()

    def _launch_post(self):
        """K2 + K2b for the current step counter / obs slot (pool-indexed or bound physics tensors)."""
        L = _lib.lib()
        if self._pool:
            _lib.check(L.urlg_post_physics_at(self._handle, _lib.ptr(self._step_base), self.common_step_counter,
                                              self._obs_slot, self.gym.sim_calls, _lib.cur_stream()))
        else:
            _lib.check(L.urlg_post_physics(self._handle, _lib.ptr(self._step_base), self.common_step_counter,
                                           self._obs_slot, _lib.cur_stream()))

    def _launch_pd(self, raw_actions=None):
        """K1 for the current substep; raw_actions (device tensor) on the first substep of a step."""
        L = _lib.lib()
        if self._pool:
            _lib.check(L.urlg_pd_torques_at(self._handle, _lib.ptr(raw_actions), self.gym.sim_calls, _lib.cur_stream()))
        else:
            _lib.check(L.urlg_pd_torques(self._handle, _lib.ptr(raw_actions), _lib.cur_stream()))

    def _commit_indexed(self):
        g = self.gym
        g.set_dof_state_tensor_indexed(self.sim, self.dof_state, self._reset_ids, self._reset_count)
        g.set_actor_root_state_tensor_indexed(self.sim, self.root_states, self._root_ids, self._root_count)

    def _publish_extras(self):
        if "episode" not in self.extras:
            self.extras["episode"] = {"rew_" + n: self._extras_episode[i] for i, n in enumerate(self.spec.sum_names)}
            if self.cfg.commands.curriculum:
                self.extras["episode"]["max_command_x"] = self.command_ranges["lin_vel_x"][1]
        if self.cfg.env.send_timeouts and "time_outs" not in self.extras:
            self.extras["time_outs"] = self._extras_time_outs

    def reset_idx(self, env_ids):
        """legged_robot.py:723-768 for an explicit id list (the in-step resets are inside K2)."""
        if len(env_ids) == 0:
            return
        ids = torch.as_tensor(env_ids, device=self.device).to(torch.int64).contiguous()
        _lib.check(_lib.lib().urlg_env_reset_idx(self._handle, _lib.ptr(ids), ids.numel(), _lib.ptr(self._step_base),
                                                 self.common_step_counter, _lib.cur_stream()))
        self._commit_indexed()
        self._publish_extras()

    # ------------------------------------------------------------------ graph support
    def begin_relative(self):
        """Fold host counters into device bases so a CUDA graph can be captured and replayed."""
        self._step_base += self.common_step_counter
        self._host_steps_folded = getattr(self, "_host_steps_folded", 0) + self.common_step_counter
        self.common_step_counter = 0
        if hasattr(self.gym, "begin_relative"):
            self.gym.begin_relative()

    def advance_base(self):
        self._step_base += self.common_step_counter
        if hasattr(self.gym, "advance_base"):
            self.gym.advance_base()

    def end_relative(self):
        self.common_step_counter = 0
        if hasattr(self.gym, "sim_calls"):
            self.gym.sim_calls = 0

    # ------------------------------------------------------------------ fused stages (kept as names)
    def _fused(self, *a, **k):
        raise NotImplementedError("this stage is fused into urlg_post_physics (csrc/env_kernels.cu); "
                                  "call post_physics_step()")
    check_termination = compute_reward = compute_observations = _post_physics_step_callback = _fused
    _push_robots = _resample_commands = _reset_dofs = _reset_root_states = _fused
