"""ORACLE shim: the reference imports `rsl_rl.env.VecEnv` and
`rsl_rl.runners.OnPolicyRunner` (legged_gym/utils/task_registry.py:8-9); the
submodule is absent from /root/reference (.SUBMODULES.json), so these names
resolve to the oracle's restatement of rsl_rl v1.0.2 (oracle/rsl_rl_port)."""
