"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

First pieces of a CPU restatement of the reference's go2_handstand task ("next"), pinned to
tests/golden/env_go2_handstand_n64.npz (recorded from the unmodified GO2HandstandEnv under oracle/refharness).
Restated so far: the termination rule and the PD variant; the observation layout and the 13 extra reward terms
follow when the task is ported.
"""
import torch


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
