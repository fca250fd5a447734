from oracle.rsl_rl_port import VecEnv  # noqa: F401
