"""G1Robot: H1-family env (gait phase, 9+3A+2 obs, privileged obs, feet state from the rigid-body tensor;
reference legged_gym/envs/g1/g1_env.py).  All of it runs inside the fused kernels; the family
flag selects the observation layout and enables the five extra reward terms."""
from ..base.legged_robot import LeggedRobot
from ....envspec import FAMILY_H1


class G1Robot(LeggedRobot):
    _family = FAMILY_H1

    def update_feet_state(self):
        """h1_env.py:48-53 as views (the kernel reads the feet rows directly)."""
        self.gym.refresh_rigid_body_state_tensor(self.sim)
        self.feet_state = self.rigid_body_states_view[:, self.feet_indices, :]
        self.feet_pos = self.feet_state[:, :, :3]
        self.feet_vel = self.feet_state[:, :, 7:10]
