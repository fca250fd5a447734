from ..robot_configs import H1_2RoughCfg, H1_2RoughCfgPPO  # noqa: F401
