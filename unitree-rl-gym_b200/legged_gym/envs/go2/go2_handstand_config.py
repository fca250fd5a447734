from ..robot_configs import GO2HandstandCfg, GO2HandstandCfgPPO  # noqa: F401
