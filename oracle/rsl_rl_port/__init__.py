"""ORACLE (test infrastructure, never on the product path).

CPU restatement of rsl_rl v1.0.2 — the PPO half of the hot path.  The submodule
is ABSENT from /root/reference (`.SUBMODULES.json`: path rsl_rl, commit
2ad79cf0caa85b91721abfe358105f869a784121, status missing; pinned to upstream tag
v1.0.2 by doc/setup_en.md:96-106), so this restates the published algorithm and
anchors on the reference's call sites: legged_gym/utils/task_registry.py:8-9,
118-126; scripts/train.py:14; scripts/play.py:34,39; utils/helpers.py:150-189;
config keys envs/base/legged_robot_config.py:244-286.

PARITY STATUS: "parity unpinned" for PPO/GAE/storage (no reference tests or
vectors exist); the recurrent actor branch IS pinned by the three shipped
TorchScript policies deploy/pre_train/{g1,h1,h1_2}/motion.pt (tests/golden).
"""
from .modules import ActorCritic, ActorCriticRecurrent, Memory  # noqa: F401
from .storage import RolloutStorage, split_and_pad_trajectories, unpad_trajectories  # noqa: F401
from .ppo import PPO  # noqa: F401
from .runner import OnPolicyRunner, VecEnv  # noqa: F401
