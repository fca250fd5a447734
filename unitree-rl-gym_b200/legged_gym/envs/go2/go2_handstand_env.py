"""GO2HandstandEnv: the reference fork's own task (legged_gym/envs/go2/go2_handstand_env.py), which overrides
_compute_torques / _post_physics_step_callback / check_termination / compute_observations / _get_noise_scale_vec and
adds eleven reward functions.  Here each override is a variant of the fused kernels selected by the family flag:

    _compute_torques (:326-350)             K1 with per-DOF target clamp (rear legs +-0.15 rad around the pose), no torque clip
    _post_physics_step_callback (:84-96)    URLG_CALLBACK_NONE: no command resampling / heading command (pushes stay: legged_robot.py:700-701)
    check_termination (:178-219)            URLG_TERMINATION_GRAVITY: contact | -0.2 < g_z < 0.2 | time-out
    compute_observations (:140-176)         46 = ang_vel, gravity, dof_pos - default, dof_vel, actions, 4 contact flags
                                            (+ noise, + 1 % flag flips from Philox slots 14+A+O..+3); privileged = clone
    _reward_* (:224-383)                    term ids URLG_T_*_HS / FRONT_* / HIND_* / POSE / STABILITY / STAY_STILL / ...
"""
import torch

from ..base.legged_robot import LeggedRobot
from ....envspec import FAMILY_HANDSTAND


class GO2HandstandEnv(LeggedRobot):
    _family = FAMILY_HANDSTAND

    def _init_buffers(self):
        super()._init_buffers()
        s = self.spec
        # go2_handstand_env.py:17-42, 66-82: what the reference's constructor derives
        self.front_feet_indices = self.feet_indices[[0, 1]]
        self.hind_feet_indices = self.feet_indices[[2, 3]]
        self.dof_name_to_idx = {name: i for i, name in enumerate(self.dof_names)}
        self.pose_targets = torch.as_tensor(s.pose_targets, device=self.device).unsqueeze(0).repeat(self.num_envs, 1)
        self.rear_dof_idx = torch.tensor([self.dof_name_to_idx[n] for n in (
            "RL_hip_joint", "RL_thigh_joint", "RL_calf_joint", "RR_hip_joint", "RR_thigh_joint", "RR_calf_joint")],
            dtype=torch.long, device=self.device)
        self.rear_center = self.pose_targets[0, self.rear_dof_idx].clone()
        self.rear_limit = 0.15

    def _update_feet_states(self):
        """:84-90 as views (the kernel reads the feet rows directly)."""
        self.gym.refresh_rigid_body_state_tensor(self.sim)
        self.feet_state = self.rigid_body_states_view[:, self.feet_indices, :]
        self.feet_pos = self.feet_state[:, :, :3]
        self.feet_vel = self.feet_state[:, :, 7:10]
