"""TEST INFRASTRUCTURE ONLY (never imported by the product path).

First pieces of a CPU restatement of the reference's go2_handstand task ("next"), pinned to
tests/golden/env_go2_handstand_n64.npz (recorded from the unmodified GO2HandstandEnv under oracle/refharness).
Only the termination rule is restated so far; the observation layout, the PD variant and the 13 extra reward terms
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
