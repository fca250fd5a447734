"""B200-native hot path of unitree-rl-gym: LeggedRobot.step (minus PhysX) and
the rsl_rl PPO loop, as hand-written sm_100a CUDA behind a C ABI
(include/urlg_b200.h), with the reference's Python APIs kept drop-in:

    from unitree_rl_gym_b200.legged_gym.envs import LeggedRobot, task_registry
    from unitree_rl_gym_b200.rsl_rl.runners import OnPolicyRunner

There is no CPU fallback: compute calls raise if the CUDA library is missing.
"""
__version__ = "0.1.0"
