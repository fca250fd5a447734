from .robot_configs import (GO2RoughCfg, GO2RoughCfgPPO, H1RoughCfg, H1RoughCfgPPO,  # noqa: F401
                            H1_2RoughCfg, H1_2RoughCfgPPO, G1RoughCfg, G1RoughCfgPPO)
