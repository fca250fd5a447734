from ..robot_configs import H1RoughCfg, H1RoughCfgPPO  # noqa: F401
