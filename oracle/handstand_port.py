"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU restatement of the reference's go2_handstand task, pinned to
tests/golden/env_go2_handstand_n64.npz (recorded from the unmodified GO2HandstandEnv under oracle/refharness):
the termination rule, the PD variant, the 46-dim observation, the eleven task-specific reward terms, and
`HandstandPort`, the EnvPort subclass that wires them into a steppable env (tests/test_oracle_golden.py).
"""
import torch

from .env_port import EnvPort


def handstand_termination(contact_forces, termination_contact_indices, projected_gravity, episode_length_buf,
                          max_episode_length):
    """reference: legged_gym/envs/go2/go2_handstand_env.py:178-219 (check_termination override).

    contact_forces [N, B, 3]; returns (reset_buf, time_out_buf), both bool [N]:
      illegal = any_b(||F_b|| > 1.0) over the termination bodies      (:196-205)
      fallen  = -0.2 < projected_gravity.z < 0.2                      (:211-212)
      timeout = episode_length_buf > max_episode_length               (:216)
      reset   = illegal | fallen | timeout                            (:219)
    The base class's roll/pitch rule (legged_robot.py:715-716) is NOT applied by this override.
    """
    n = contact_forces.shape[0]
    if termination_contact_indices.numel() > 0:
        illegal = torch.any(torch.norm(contact_forces[:, termination_contact_indices, :], dim=-1) > 1.0, dim=1)
    else:
        illegal = torch.zeros(n, dtype=torch.bool)
    cos_up = projected_gravity[:, 2]
    fallen = (cos_up > -0.2) & (cos_up < 0.2)
    time_out = episode_length_buf > max_episode_length
    return illegal | fallen | time_out, time_out


def handstand_pd_torques(actions, dof_pos, dof_vel, p_gains, d_gains, default_dof_pos, action_scale, rear_dof_idx,
                         rear_center, rear_limit):
    """reference: legged_gym/envs/go2/go2_handstand_env.py:326-350 (_compute_torques override; `actions` already
    clipped to +-clip_actions by LeggedRobot.step, legged_robot.py:623-624).

      target = default_dof_pos + action_scale * actions                                   (:333)
      target[:, rear] = clamp(target[:, rear], rear_center - rear_limit, rear_center + rear_limit)   (:336-346)
      torque = kp * (target - q) - kd * qd        -- no torque-limit clip, unlike the base class   (:349-350)
    rear = the six RL_/RR_ hip, thigh, calf joints (:27-39), rear_center = handstand pose of those joints (:41),
    rear_limit = 0.15 rad (:42).
    """
    targets = default_dof_pos + action_scale * actions
    if rear_dof_idx.numel() > 0:
        targets[:, rear_dof_idx] = torch.clamp(targets[:, rear_dof_idx], rear_center - rear_limit, rear_center + rear_limit)
    return p_gains * (targets - dof_pos) - d_gains * dof_vel


def handstand_observations(base_ang_vel, projected_gravity, dof_pos, dof_vel, default_dof_pos, actions, contact_forces,
                           front_feet_indices, hind_feet_indices, ang_vel_scale, dof_pos_scale, dof_vel_scale,
                           noise_uniforms=None, noise_scale_vec=None, flips=None):
    """reference: legged_gym/envs/go2/go2_handstand_env.py:140-176 (the second, effective compute_observations).

    46 = [base_ang_vel*s (3), projected_gravity (3), (q - q_default)*s (12), qd*s (12), actions (12),
          1[F_z(front feet) > 5] (2), 1[F_z(hind feet) > 5] (2)]                                   (:141-155)
    noise: obs += (2u - 1) * noise_scale_vec, u uniform per element                                (:157-158)
    then the four contact flags are re-binarised and flipped with probability 0.01:
      flag = (flag > 0.5) xor flip                                                                 (:160-169)
    `flips` (bool [N, 4]) are the Bernoulli draws; None = no flip.  No command and no base linear velocity in this
    observation; privileged_obs_buf is a clone (:172).
    """
    front = (contact_forces[:, front_feet_indices, 2] > 5.0).float()
    hind = (contact_forces[:, hind_feet_indices, 2] > 5.0).float()
    n = base_ang_vel.shape[0]
    obs = torch.cat((base_ang_vel * ang_vel_scale, projected_gravity, (dof_pos - default_dof_pos) * dof_pos_scale,
                     dof_vel * dof_vel_scale, actions, front.view(n, -1), hind.view(n, -1)), dim=-1)
    if noise_uniforms is not None:
        obs = obs + (2.0 * noise_uniforms - 1.0) * noise_scale_vec
        f = flips if flips is not None else torch.zeros(n, 4, dtype=torch.bool)
        obs[:, -4:] = torch.logical_xor(obs[:, -4:] > 0.5, f).float()
    return obs


# ---- reward terms of the handstand task that read only base-frame velocities, gravity, feet contact and hip angles.
#      Each returns [N] f32 BEFORE the scale*dt factor (legged_robot.py:770-787 applies it).  The task's remaining active terms (action_rate,
#      ang_vel_xy, collision, dof_pos_limits, dof_vel, lin_vel_z, torques) are the base class's, already in env_port.py.
def reward_orientation(projected_gravity):
    """go2_handstand_env.py:224-235: exp(-3 (1 - clamp(g_z, -1, 1))^2)."""
    err = 1.0 - projected_gravity[:, 2].clamp(-1.0, 1.0)
    return torch.exp(-3.0 * err * err)


def reward_front_feet_contact(contact_forces, front_feet_indices):
    """go2_handstand_env.py:248-256: mean over the two front feet of 1[F_z > 5]."""
    return (contact_forces[:, front_feet_indices, 2] > 5.0).float().mean(dim=1)


def reward_hind_feet_no_contact(contact_forces, hind_feet_indices):
    """go2_handstand_env.py:258-265: 1 - mean over the two hind feet of 1[F_z > 5]."""
    return 1.0 - (contact_forces[:, hind_feet_indices, 2] > 5.0).float().mean(dim=1)


def reward_stability(base_lin_vel, base_ang_vel):
    """go2_handstand_env.py:309-317: exp(-2 (|v_xy| + 0.5 |w_xy|))."""
    return torch.exp(-2.0 * (torch.norm(base_lin_vel[:, :2], dim=1) + 0.5 * torch.norm(base_ang_vel[:, :2], dim=1)))


def reward_front_hip_neutral(dof_pos, fl_hip, fr_hip):
    """go2_handstand_env.py:352-358: exp(-4 (q_FLhip^2 + q_FRhip^2))."""
    hip = dof_pos[:, [fl_hip, fr_hip]]
    return torch.exp(-4.0 * (hip ** 2).sum(dim=1))


def reward_stay_still(base_lin_vel, base_ang_vel):
    """go2_handstand_env.py:375-380: exp(-(|v_xy| / 0.05)^2 - (|w_z| / 0.2)^2)."""
    lin_xy = torch.norm(base_lin_vel[:, :2], dim=1)
    return torch.exp(-(lin_xy / 0.05) ** 2 - (torch.abs(base_ang_vel[:, 2]) / 0.2) ** 2)


def reward_lin_vel_xy(base_lin_vel):
    """go2_handstand_env.py:382-383: |v_xy|."""
    return torch.norm(base_lin_vel[:, :2], dim=1)


def reward_base_height(root_z, base_height_target):
    """go2_handstand_env.py:237-246: exp(-8 max(target - z, 0)^2) — only heights BELOW the target are penalised."""
    err = (base_height_target - root_z).clamp(min=0.0)
    return torch.exp(-8.0 * err * err)


def reward_pose(dof_pos, pose_targets, front_ids=(0, 1, 2, 3, 4, 5), hind_ids=(6, 7, 8, 9, 10, 11)):
    """go2_handstand_env.py:277-307 (the second, effective _reward_pose): exp(-4 (sum_front e^2 + 0.3 sum_hind e^2)),
    e = q - handstand pose; front = FL/FR hip, thigh, calf, hind = RL/RR hip, thigh, calf."""
    err = dof_pos - pose_targets
    sq = (err[:, list(front_ids)] ** 2).sum(dim=1) + 0.3 * (err[:, list(hind_ids)] ** 2).sum(dim=1)
    return torch.exp(-4.0 * sq)


def reward_front_feet_together(feet_pos):
    """go2_handstand_env.py:360-373: exp(-20 max(|p_FL.xy - p_FR.xy| - 0.12, 0)^2); feet_pos [N, F, 3] are the feet rows
    of the rigid-body state tensor, refreshed in the task's _post_physics_step_callback (:84-96).  That override does
    NOT call the base class's callback, so this task never resamples commands in the callback nor recomputes the heading
    command (legged_robot.py:488-517 is skipped); pushes remain (post_physics_step calls _push_robots itself, :700-701)."""
    dist = torch.norm(feet_pos[:, 0, :2] - feet_pos[:, 1, :2], dim=1)
    err = (dist - 0.12).clamp(min=0.0)
    return torch.exp(-20.0 * err * err)


def reward_energy(torques, dof_vel):
    """go2_handstand_env.py:319-324: sum |tau| |qd| (scale 0 in the shipped config)."""
    return torch.sum(torch.abs(torques) * torch.abs(dof_vel), dim=1)


class HandstandPort(EnvPort):
    """GO2HandstandEnv as overrides of the base port, mirroring which methods the reference subclass overrides."""

    def __init__(self, spec, threads=None):
        super().__init__(spec, threads)
        s = spec
        self.pose_targets = torch.as_tensor(s.pose_targets, dtype=torch.float32).view(1, -1).repeat(s.num_envs, 1)
        idx = {n: i for i, n in enumerate(s.dof_names)}
        self.rear_dof_idx = torch.tensor([idx[n] for n in ("RL_hip_joint", "RL_thigh_joint", "RL_calf_joint",
                                                            "RR_hip_joint", "RR_thigh_joint", "RR_calf_joint")])
        self.rear_center = self.pose_targets[0, self.rear_dof_idx].clone()
        self.rear_limit = 0.15
        self.front_feet = torch.tensor(s.feet_idx[:2])
        self.hind_feet = torch.tensor(s.feet_idx[2:4])
        self.fl_hip, self.fr_hip = idx["FL_hip_joint"], idx["FR_hip_joint"]
        self.privileged_obs_buf = torch.zeros(s.num_envs, s.num_obs)

    def compute_torques(self):                                   # :326-350
        s = self.spec
        self.torques = handstand_pd_torques(self.actions, self.dof_pos, self.dof_vel, self.p_gains, self.d_gains,
                                            self.default_dof_pos, s.action_scale, self.rear_dof_idx, self.rear_center,
                                            self.rear_limit)
        return self.torques

    def _post_physics_step_callback(self, U):                    # :92-96: feet states only
        pass

    def check_termination(self):                                 # :178-219
        s = self.spec
        self.reset_buf, self.time_out_buf = handstand_termination(
            self.contact_forces, torch.tensor(s.term_idx), self.projected_gravity, self.episode_length_buf,
            s.max_episode_length)

    def _term(self, name):
        s = self.spec
        if name == "orientation":
            return reward_orientation(self.projected_gravity)
        if name == "base_height":
            return reward_base_height(self.root_states[:, 2], s.base_height_target)
        if name == "front_feet_contact":
            return reward_front_feet_contact(self.contact_forces, self.front_feet)
        if name == "hind_feet_no_contact":
            return reward_hind_feet_no_contact(self.contact_forces, self.hind_feet)
        if name == "pose":
            return reward_pose(self.dof_pos, self.pose_targets, range(s.num_front_dof), range(s.num_front_dof, s.num_dof))
        if name == "stability":
            return reward_stability(self.base_lin_vel, self.base_ang_vel)
        if name == "energy":
            return reward_energy(self.torques, self.dof_vel)
        if name == "front_hip_neutral":
            return reward_front_hip_neutral(self.dof_pos, self.fl_hip, self.fr_hip)
        if name == "front_feet_together":
            return reward_front_feet_together(self.rigid_body_states[:, s.feet_idx, :3])
        if name == "stay_still":
            return reward_stay_still(self.base_lin_vel, self.base_ang_vel)
        if name == "lin_vel_xy":
            return reward_lin_vel_xy(self.base_lin_vel)
        return super()._term(name)

    def compute_observations(self, U):                           # :140-176; flips from slots 14+A+O .. +3
        s = self.spec
        A, O = s.num_dof, s.num_obs
        noise = U[:, 14 + A:14 + A + O] if s.add_noise else None
        flips = (U[:, 14 + A + O:14 + A + O + 4] < 0.01) if s.add_noise else None
        self.obs_buf = handstand_observations(
            self.base_ang_vel, self.projected_gravity, self.dof_pos, self.dof_vel, self.default_dof_pos, self.actions,
            self.contact_forces, self.front_feet, self.hind_feet, s.obs_scale_ang_vel, s.obs_scale_dof_pos,
            s.obs_scale_dof_vel, noise, self.noise_scale_vec, flips)
        self.privileged_obs_buf = self.obs_buf.clone()
