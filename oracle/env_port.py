"""ORACLE (test infrastructure, never on the product path; only tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs may import this).

CPU restatement (torch fp32, vectorised over envs) of the env half of the hot path:
LeggedRobot.step minus the PhysX call, for the base (Go2) and H1-family (H1, H1_2, G1)
env classes.  Every function cites the reference lines it follows
(/root/reference/legged_gym/envs/base/legged_robot.py unless noted).

PINNED: against the unmodified reference classes run under oracle/refharness
(tests/test_oracle_vs_reference.py, container only) and against the fixtures that run
produced (tests/golden/env_*.npz, committed; tests/test_oracle_golden.py).

Random numbers come from the shared counter-based table (oracle/philox.py); slot map in
oracle/refharness/run_reference.py.
"""
import math

import numpy as np
import torch

from . import philox

TWO_PI = 2 * math.pi


def quat_rotate_inverse(q, v):
    """[UPSTREAM isaacgym.torch_utils] v*(2w^2-1) - 2w(u x v) + 2u(u.v), q = xyzw."""
    w = q[:, 3]
    u = q[:, :3]
    a = v * (2.0 * w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(u, v, dim=-1) * w.unsqueeze(-1) * 2.0
    c = u * (u * v).sum(-1, keepdim=True) * 2.0
    return a - b + c


def quat_apply(q, v):
    """[UPSTREAM isaacgym.torch_utils] v + w t + u x t, t = 2 (u x v)."""
    u = q[:, :3]
    t = torch.cross(u, v, dim=-1) * 2
    return v + q[:, 3:] * t + torch.cross(u, t, dim=-1)


def euler_xyz(q):
    """legged_gym/utils/isaacgym_utils.py:11-30 — angles in (-pi, pi], pitch saturates."""
    x, y, z, w = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    roll = torch.atan2(2.0 * (w * x + y * z), w * w - x * x - y * y + z * z)
    sinp = 2.0 * (w * y - z * x)
    sat = torch.full_like(sinp, math.pi / 2.0) * torch.sign(sinp)
    pitch = torch.where(sinp.abs() >= 1, sat, torch.asin(sinp))
    yaw = torch.atan2(2.0 * (w * z + x * y), w * w + x * x - y * y - z * z)
    return torch.stack((roll, pitch, yaw), dim=-1)


def wrap_to_pi(a):
    """legged_gym/utils/math.py:15-18."""
    a = torch.remainder(a, TWO_PI)
    return a - TWO_PI * (a > math.pi)


class EnvPort:
    """State + step functions.  `spec` is duck-typed (unitree_rl_gym_b200.envspec.EnvSpec or
    the attributes scraped from a reference env by tests/golden/gen_golden.py)."""

    def __init__(self, spec, threads=None, device="cpu", fast_rng=False):
        """device / fast_rng serve the timed baselines only (bench.py): `device="cuda"` runs the same eager torch ops
        on the GPU (the "eager GPU" context number), `fast_rng` draws the uniforms with torch.rand on that device --
        what the reference itself does -- instead of building the shared Philox table in numpy (parity runs need the
        table; timing runs must not pay for it)."""
        self._device, self._fast_rng = torch.device(device), bool(fast_rng)
        s = self.spec = spec
        N, A, F = s.num_envs, s.num_dof, len(s.feet_idx)
        f = lambda *shape: torch.zeros(*shape, dtype=torch.float32)
        # physics seam tensors (:83-119, h1_env.py:37-42)
        self.root_states = f(N, 13)
        self.root_states[:, 6] = 1.0
        self.dof_state = f(N * A, 2)
        self.contact_forces = f(N, s.num_bodies, 3)
        self.rigid_body_states = f(N, s.num_bodies, 13)
        # env buffers (:125-186, base_task.py:41-49)
        self.obs_buf = f(N, s.num_obs)
        self.privileged_obs_buf = f(N, s.num_priv) if s.num_priv else None
        self.rew_buf = f(N)
        self.reset_buf = torch.ones(N, dtype=torch.bool)
        self.time_out_buf = torch.zeros(N, dtype=torch.bool)
        self.episode_length_buf = torch.zeros(N, dtype=torch.long)
        self.torques, self.actions, self.last_actions, self.last_dof_vel = f(N, A), f(N, A), f(N, A), f(N, A)
        self.last_root_vel = f(N, 6)
        self.commands = f(N, 4)
        self.feet_air_time = f(N, F)
        self.last_contacts = torch.zeros(N, F, dtype=torch.bool)
        self.rpy, self.base_lin_vel, self.base_ang_vel, self.projected_gravity = f(N, 3), f(N, 3), f(N, 3), f(N, 3)
        self.phase = f(N)
        self.leg_phase = f(N, 2)
        self.episode_sums = {k: f(N) for k in s.sum_names}
        self.extras = {}
        self.common_step_counter = 0
        # constants
        t = lambda x: torch.as_tensor(np.asarray(x), dtype=torch.float32)
        self.p_gains, self.d_gains = t(s.p_gains), t(s.d_gains)
        self.default_dof_pos = t(s.default_dof_pos).view(1, A)
        self.torque_limits, self.dof_vel_limits = t(s.torque_limits), t(s.dof_vel_limits)
        self.dof_pos_limits = t(s.dof_pos_limits)
        self.noise_scale_vec = t(s.noise_scale_vec)
        self.base_init_state = t(s.base_init_state)
        self.env_origins = t(s.env_origins)
        self.gravity_vec = torch.tensor([0., 0., -1.]).repeat(N, 1)
        self.forward_vec = torch.tensor([1., 0., 0.]).repeat(N, 1)
        self.commands_scale = torch.tensor([s.obs_scale_lin_vel, s.obs_scale_lin_vel, s.obs_scale_ang_vel])
        self.env_ids_global = np.arange(N, dtype=np.int64) + s.env_offset
        self._ucache = (None, None)
        if self._device.type != "cpu":
            for k, v in list(vars(self).items()):
                if torch.is_tensor(v):
                    setattr(self, k, v.to(self._device))
            self.episode_sums = {k: v.to(self._device) for k, v in self.episode_sums.items()}
        self._gen = torch.Generator(device=self._device).manual_seed(int(s.seed)) if self._fast_rng else None

    # views, as :108-109
    @property
    def dof_pos(self):
        return self.dof_state.view(self.spec.num_envs, self.spec.num_dof, 2)[..., 0]

    @property
    def dof_vel(self):
        return self.dof_state.view(self.spec.num_envs, self.spec.num_dof, 2)[..., 1]

    def uniforms(self, step):
        if self._fast_rng:
            if self._ucache[0] != step:
                self._ucache = (step, torch.rand(self.spec.num_envs, self.spec.num_rng_slots, device=self._device,
                                                 generator=self._gen))
            return self._ucache[1]
        if self._ucache[0] != step:
            u = philox.uniform_table(self.spec.seed, philox.STREAM_ENV, self.env_ids_global, step,
                                     self.spec.num_rng_slots)
            self._ucache = (step, torch.from_numpy(u.copy()))
        return self._ucache[1]

    # ------------------------------------------------------------------ E2: PD law :649-671
    def compute_torques(self):
        s = self.spec
        a = self.actions * s.action_scale
        if s.control_type == "P":
            tq = self.p_gains * (a + self.default_dof_pos - self.dof_pos) - self.d_gains * self.dof_vel
        elif s.control_type == "V":
            tq = self.p_gains * (a - self.dof_vel) - self.d_gains * (self.dof_vel - self.last_dof_vel) / s.sim_dt
        elif s.control_type == "T":
            tq = a
        else:
            raise NameError(f"Unknown controller type: {s.control_type}")
        self.torques = torch.clip(tq, -self.torque_limits, self.torque_limits)
        return self.torques

    def set_actions(self, actions):
        """:623-624."""
        c = self.spec.clip_actions
        self.actions = torch.clip(actions, -c, c).clone()

    # ------------------------------------------------------------------ E6 :519-538
    def _resample(self, mask, U, slot):
        s = self.spec
        lo, hi = s.cmd_x
        cx = (hi - lo) * U[:, slot] + lo
        lo, hi = s.cmd_y
        cy = (hi - lo) * U[:, slot + 1] + lo
        self.commands[:, 0] = torch.where(mask, cx, self.commands[:, 0])
        self.commands[:, 1] = torch.where(mask, cy, self.commands[:, 1])
        if s.heading_command:
            lo, hi = s.cmd_heading
            self.commands[:, 3] = torch.where(mask, (hi - lo) * U[:, slot + 2] + lo, self.commands[:, 3])
        else:
            lo, hi = s.cmd_yaw
            self.commands[:, 2] = torch.where(mask, (hi - lo) * U[:, slot + 2] + lo, self.commands[:, 2])
        keep = (torch.norm(self.commands[:, :2], dim=1) > 0.2).unsqueeze(1)
        self.commands[:, :2] = torch.where(mask.unsqueeze(1), self.commands[:, :2] * keep, self.commands[:, :2])

    # ------------------------------------------------------------------ E10 :723-768
    def reset_idx(self, mask, U):
        s = self.spec
        A = s.num_dof
        if not bool(mask.any()):
            return
        m1 = mask.unsqueeze(1)
        # _reset_dofs :566-567
        new_pos = self.default_dof_pos * ((1.5 - 0.5) * U[:, 3:3 + A] + 0.5)
        st = self.dof_state.view(s.num_envs, A, 2)
        st[..., 0] = torch.where(m1, new_pos, st[..., 0])
        st[..., 1] = torch.where(m1, torch.zeros_like(new_pos), st[..., 1])
        # _reset_root_states :587-590 (custom_origins is always False, :263)
        fresh = self.base_init_state.unsqueeze(0).repeat(s.num_envs, 1)
        fresh[:, :3] += self.env_origins
        fresh[:, 7:13] = (0.5 - (-0.5)) * U[:, 3 + A:9 + A] + (-0.5)
        self.root_states[:] = torch.where(m1, fresh, self.root_states)
        self._resample(mask, U, 9 + A)                                    # :749
        z = torch.zeros(())
        self.actions = torch.where(m1, z, self.actions)                   # :752-756
        self.last_actions = torch.where(m1, z, self.last_actions)
        self.last_dof_vel = torch.where(m1, z, self.last_dof_vel)
        self.feet_air_time = torch.where(m1, z, self.feet_air_time)
        self.episode_length_buf = torch.where(mask, torch.zeros_like(self.episode_length_buf), self.episode_length_buf)
        self.reset_buf = self.reset_buf | mask                            # :757
        ep = {}
        for k in self.episode_sums:                                       # :759-763
            ep["rew_" + k] = torch.mean(self.episode_sums[k][mask]) / s.max_episode_length_s
            self.episode_sums[k] = torch.where(mask, z, self.episode_sums[k])
        self.extras["episode"] = ep
        if s.send_timeouts:                                               # :767-768
            self.extras["time_outs"] = self.time_out_buf

    # ------------------------------------------------------------------ E9 reward terms :843-939, h1_env.py:98-123
    def _term(self, name):
        s = self.spec
        cf = self.contact_forces
        feet = s.feet_idx
        if name == "action_rate":
            return torch.sum(torch.square(self.last_actions - self.actions), dim=1)
        if name == "alive":
            return torch.ones(s.num_envs)
        if name == "ang_vel_xy":
            return torch.sum(torch.square(self.base_ang_vel[:, :2]), dim=1)
        if name == "base_height":
            return torch.square(self.root_states[:, 2] - s.base_height_target)
        if name == "collision":
            return torch.sum(1. * (torch.norm(cf[:, s.pen_idx, :], dim=-1) > 0.1), dim=1)
        if name == "contact":
            res = torch.zeros(s.num_envs)
            for i, b in enumerate(feet):
                stance = self.leg_phase[:, i] < 0.55
                res += ~((cf[:, b, 2] > 1) ^ stance)
            return res
        if name == "contact_no_vel":
            c = torch.norm(cf[:, feet, :3], dim=2) > 1.
            v = self.rigid_body_states[:, feet, 7:10] * c.unsqueeze(-1)
            return torch.sum(torch.square(v), dim=(1, 2))
        if name == "dof_acc":
            return torch.sum(torch.square((self.last_dof_vel - self.dof_vel) / s.dt), dim=1)
        if name == "dof_pos_limits":
            o = -(self.dof_pos - self.dof_pos_limits[:, 0]).clip(max=0.)
            o = o + (self.dof_pos - self.dof_pos_limits[:, 1]).clip(min=0.)
            return torch.sum(o, dim=1)
        if name == "dof_vel":
            return torch.sum(torch.square(self.dof_vel), dim=1)
        if name == "dof_vel_limits":
            return torch.sum((torch.abs(self.dof_vel) - self.dof_vel_limits * s.soft_dof_vel_limit).clip(min=0., max=1.), dim=1)
        if name == "feet_air_time":                      # stateful, :912-923
            contact = cf[:, feet, 2] > 1.
            filt = torch.logical_or(contact, self.last_contacts)
            self.last_contacts = contact
            first = (self.feet_air_time > 0.) * filt
            self.feet_air_time = self.feet_air_time + s.dt
            r = torch.sum((self.feet_air_time - 0.5) * first, dim=1)
            r = r * (torch.norm(self.commands[:, :2], dim=1) > 0.1)
            self.feet_air_time = self.feet_air_time * ~filt
            return r
        if name == "feet_contact_forces":
            return torch.sum((torch.norm(cf[:, feet, :], dim=-1) - s.max_contact_force).clip(min=0.), dim=1)
        if name == "feet_swing_height":
            c = torch.norm(cf[:, feet, :3], dim=2) > 1.
            return torch.sum(torch.square(self.rigid_body_states[:, feet, 2] - 0.08) * ~c, dim=1)
        if name == "hip_pos":
            return torch.sum(torch.square(self.dof_pos[:, s.hip_idx]), dim=1)
        if name == "lin_vel_z":
            return torch.square(self.base_lin_vel[:, 2])
        if name == "orientation":
            return torch.sum(torch.square(self.projected_gravity[:, :2]), dim=1)
        if name == "stand_still":
            return torch.sum(torch.abs(self.dof_pos - self.default_dof_pos), dim=1) * (torch.norm(self.commands[:, :2], dim=1) < 0.1)
        if name == "stumble":
            return torch.any(torch.norm(cf[:, feet, :2], dim=2) > 5 * torch.abs(cf[:, feet, 2]), dim=1).float()
        if name == "torque_limits":
            return torch.sum((torch.abs(self.torques) - self.torque_limits * s.soft_torque_limit).clip(min=0.), dim=1)
        if name == "torques":
            return torch.sum(torch.square(self.torques), dim=1)
        if name == "tracking_ang_vel":
            return torch.exp(-torch.square(self.commands[:, 2] - self.base_ang_vel[:, 2]) / s.tracking_sigma)
        if name == "tracking_lin_vel":
            e = torch.sum(torch.square(self.commands[:, :2] - self.base_lin_vel[:, :2]), dim=1)
            return torch.exp(-e / s.tracking_sigma)
        raise AttributeError("_reward_" + name)

    # ------------------------------------------------------------------ E3: post_physics_step :673-709
    def post_physics_step(self):
        s = self.spec
        A = s.num_dof
        self.episode_length_buf = self.episode_length_buf + 1
        self.common_step_counter += 1
        U = self.uniforms(self.common_step_counter)
        q = self.root_states[:, 3:7]
        # E4 :685-690
        self.rpy = euler_xyz(q)
        self.base_lin_vel = quat_rotate_inverse(q, self.root_states[:, 7:10])
        self.base_ang_vel = quat_rotate_inverse(q, self.root_states[:, 10:13])
        self.projected_gravity = quat_rotate_inverse(q, self.gravity_vec)
        self._post_physics_step_callback(U)
        self.check_termination()
        # E8 :775-787
        self.rew_buf = torch.zeros(s.num_envs)
        for name, scale in zip(s.reward_names, s.reward_scales):
            r = self._term(name) * scale
            self.rew_buf = self.rew_buf + r
            self.episode_sums[name] = self.episode_sums[name] + r
        if s.only_positive:
            self.rew_buf = torch.clip(self.rew_buf, min=0.)
        if s.termination_scale is not None:
            r = (self.reset_buf * ~self.time_out_buf) * s.termination_scale
            self.rew_buf = self.rew_buf + r
            self.episode_sums["termination"] = self.episode_sums["termination"] + r
        # :697-698
        self.reset_idx(self.reset_buf.clone(), U)
        self._push_robots(U)
        # E12 :789-811 / h1_env.py:68-95
        self.compute_observations(U)
        self.last_actions = self.actions.clone()
        self.last_dof_vel = self.dof_vel.clone()
        self.last_root_vel = self.root_states[:, 7:13].clone()

    # E5 callback: h1_env.py:55-65 then :511-516
    def _post_physics_step_callback(self, U):
        s = self.spec
        q = self.root_states[:, 3:7]
        if s.family == 1:
            self.phase = (self.episode_length_buf * s.dt) % 0.8 / 0.8
            self.leg_phase = torch.stack([self.phase, (self.phase + 0.5) % 1], dim=-1)
        due = self.episode_length_buf % s.resample_steps == 0
        self._resample(due, U, 0)
        if s.heading_command:
            fwd = quat_apply(q, self.forward_vec)
            heading = torch.atan2(fwd[:, 1], fwd[:, 0])
            self.commands[:, 2] = torch.clip(0.5 * wrap_to_pi(self.commands[:, 3] - heading), -1., 1.)

    # E7 :715-721
    def check_termination(self):
        s = self.spec
        self.reset_buf = torch.any(torch.norm(self.contact_forces[:, s.term_idx, :], dim=-1) > 1., dim=1)
        self.reset_buf = self.reset_buf | torch.logical_or(torch.abs(self.rpy[:, 1]) > 1.0, torch.abs(self.rpy[:, 0]) > 0.8)
        self.time_out_buf = self.episode_length_buf > s.max_episode_length
        self.reset_buf = self.reset_buf | self.time_out_buf

    # E11 :540-555 (quirk: any due env => all envs rewritten).  Called from post_physics_step itself (:700-701), after
    # reset_idx — not from the callback — so a subclass that overrides the callback still pushes.
    def _push_robots(self, U):
        s = self.spec
        A = s.num_dof
        if s.push_robots and bool((self.episode_length_buf % s.push_interval == 0).any()):
            m = s.max_push_vel
            self.root_states[:, 7:9] = (m - (-m)) * U[:, 12 + A:14 + A] + (-m)

    def compute_observations(self, U):
        s = self.spec
        A = s.num_dof
        common = [self.base_ang_vel * s.obs_scale_ang_vel, self.projected_gravity,
                  self.commands[:, :3] * self.commands_scale,
                  (self.dof_pos - self.default_dof_pos) * s.obs_scale_dof_pos,
                  self.dof_vel * s.obs_scale_dof_vel, self.actions]
        lin = self.base_lin_vel * s.obs_scale_lin_vel
        if s.family == 0:
            self.obs_buf = torch.cat([lin] + common, dim=-1)
        else:
            ph = [torch.sin(TWO_PI * self.phase).unsqueeze(1), torch.cos(TWO_PI * self.phase).unsqueeze(1)]
            self.obs_buf = torch.cat(common + ph, dim=-1)
            self.privileged_obs_buf = torch.cat([lin] + common + ph, dim=-1)
        if s.add_noise:
            self.obs_buf = self.obs_buf + (2 * U[:, 14 + A:14 + A + s.num_obs] - 1) * self.noise_scale_vec

    # ------------------------------------------------------------------ E1 :615-647 with the physics seam as a callback
    def step(self, actions, physics):
        """physics(substep) must refresh dof_state; physics(None) refreshes root/contact/rigid-body."""
        s = self.spec
        self.set_actions(actions)
        for k in range(s.decimation):
            self.compute_torques()
            physics(k)
        physics(None)
        self.post_physics_step()
        c = s.clip_obs
        self.obs_buf = torch.clip(self.obs_buf, -c, c)
        if self.privileged_obs_buf is not None:
            self.privileged_obs_buf = torch.clip(self.privileged_obs_buf, -c, c)
        return self.obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf, self.extras

    def reset(self, physics):
        """base_task.py:82-86."""
        s = self.spec
        U = self.uniforms(self.common_step_counter)
        self.reset_idx(torch.ones(s.num_envs, dtype=torch.bool), U)
        obs, priv, _, _, _ = self.step(torch.zeros(s.num_envs, s.num_dof), physics)
        return obs, priv

    def load_frame(self, frames, f, sub):
        """Synthetic physics: copy frame f (sub = substep index or None for the post-step tensors)."""
        if sub is None:
            self.root_states.copy_(frames["root"][f])
            self.contact_forces.view(-1, 3).copy_(frames["contact"][f])
            if frames.get("rigid_body") is not None:
                self.rigid_body_states.view(-1, 13).copy_(frames["rigid_body"][f])
        else:
            self.dof_state.copy_(frames["dof"][f, sub])
