from ..robot_configs import GO2RoughCfg, GO2RoughCfgPPO  # noqa: F401
