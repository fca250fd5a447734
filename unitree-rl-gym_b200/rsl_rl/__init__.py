"""Drop-in rsl_rl (v1.0.2 API) whose hot path runs in liburlg_b200's sm_100a kernels.

Same import paths as the package the reference imports (legged_gym/utils/task_registry.py:8-9):
    from rsl_rl.env import VecEnv
    from rsl_rl.runners import OnPolicyRunner
"""
from . import algorithms, env, modules, runners, storage, utils  # noqa: F401
