from oracle.rsl_rl_port import OnPolicyRunner  # noqa: F401
