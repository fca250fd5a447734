"""Task registration (mirrors legged_gym/envs/__init__.py:21-27)."""
from .robot_configs import (GO2RoughCfg, GO2RoughCfgPPO, H1RoughCfg, H1RoughCfgPPO,  # noqa: F401
                            H1_2RoughCfg, H1_2RoughCfgPPO, G1RoughCfg, G1RoughCfgPPO,
                            GO2HandstandCfg, GO2HandstandCfgPPO)
from .go2.go2_handstand_env import GO2HandstandEnv  # noqa: F401
from .base.legged_robot import LeggedRobot  # noqa: F401
from .h1.h1_env import H1Robot  # noqa: F401
from .h1_2.h1_2_env import H1_2Robot  # noqa: F401
from .g1.g1_env import G1Robot  # noqa: F401
from ..utils.task_registry import task_registry

task_registry.register("go2", LeggedRobot, GO2RoughCfg(), GO2RoughCfgPPO())
task_registry.register("h1", H1Robot, H1RoughCfg(), H1RoughCfgPPO())
task_registry.register("h1_2", H1_2Robot, H1_2RoughCfg(), H1_2RoughCfgPPO())
task_registry.register("g1", G1Robot, G1RoughCfg(), G1RoughCfgPPO())
task_registry.register("go2_handstand", GO2HandstandEnv, GO2HandstandCfg(), GO2HandstandCfgPPO())
