from ..robot_configs import G1RoughCfg, G1RoughCfgPPO  # noqa: F401
