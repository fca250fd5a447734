"""Stub gymutil: device-string parsing and sim-config copy only."""
def parse_device_str(s):
    s = str(s)
    if ':' in s:
        kind, idx = s.split(':')
        return kind, int(idx)
    return s, 0

def parse_sim_config(cfg, sim_params):
    for k, v in cfg.items():
        if isinstance(v, dict):
            sub = getattr(sim_params, k, None)
            if sub is not None:
                parse_sim_config(v, sub)
        else:
            setattr(sim_params, k, v)

def parse_arguments(description="", custom_parameters=()):
    raise RuntimeError("stub isaacgym: build args in the harness instead")
