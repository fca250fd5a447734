"""Parity of the fused CUDA env path (through the C ABI, via the drop-in LeggedRobot classes)
against fixtures produced by the unmodified reference and against the oracle port."""
import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu


def make_env(case, fx=None, frames=None, seed=1):
    from unitree_rl_gym_b200.legged_gym.envs import LeggedRobot, H1Robot, H1_2Robot, G1Robot
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    from unitree_rl_gym_b200.physics import SyntheticGym
    from unitree_rl_gym_b200.robots import ROBOTS
    task, n, pos = H.ENV_CASES[case]
    cfg = H.make_cfg(task, n, pos)
    cfg.seed = seed
    sp = SimParams()
    sp.dt = cfg.sim.dt
    gym = SyntheticGym(ROBOTS[task], n, cfg.control.decimation, "cuda:0",
                       frames=frames if frames is not None else H.frames_of(fx))
    cls = {"go2": LeggedRobot, "h1": H1Robot, "h1_2": H1_2Robot, "g1": G1Robot}[task]
    return cls(cfg=cfg, sim_params=sp, physics_engine=gym, sim_device="cuda:0", headless=True)


def env_snapshot(e):
    names = ["obs_buf", "rew_buf", "reset_buf", "time_out_buf", "episode_length_buf", "torques", "commands",
             "actions", "last_actions", "last_dof_vel", "last_root_vel", "feet_air_time", "last_contacts", "rpy",
             "base_lin_vel", "base_ang_vel", "projected_gravity", "root_states", "dof_state"]
    out = {k: getattr(e, k) for k in names}
    if e.privileged_obs_buf is not None:
        out["privileged_obs_buf"] = e.privileged_obs_buf
    for k, v in e.episode_sums.items():
        out["episode_sums/" + k] = v
    for k, v in e.extras.get("episode", {}).items():
        out["extras_episode/" + k] = v.reshape(1)
    if "time_outs" in e.extras:
        out["extras_time_outs"] = e.extras["time_outs"]
    torch.cuda.synchronize()
    return {k: v.detach().cpu().numpy() for k, v in out.items()}


@pytest.mark.parametrize("case", list(H.ENV_CASES))
def test_env_matches_reference_fixture(case):
    fx = H.load_case(case)
    env = make_env(case, fx)
    env.reset()
    snap = env_snapshot(env)
    for k in H.snapshot_keys(fx, "reset"):
        H.compare(k, snap[k], fx["reset/" + k], "reset/")
    env.episode_length_buf = torch.from_numpy(fx["planted_episode_length"]).cuda()
    for t in range(int(fx["const/scalars"][4])):
        env.step(torch.from_numpy(fx[f"actions/{t}"]).cuda())
        snap = env_snapshot(env)
        for k in H.snapshot_keys(fx, f"step{t}"):
            H.compare(k, snap[k], fx[f"step{t}/{k}"], f"step{t}/")


@pytest.mark.parametrize("task,n", [("go2", 4096), ("g1", 1000), ("h1_2", 2049)])
def test_env_matches_oracle_port_larger(task, n):
    """Same seeded inputs at sizes the oracle finishes in seconds; ragged tails (n % 32 != 0)."""
    from oracle.env_port import EnvPort
    from unitree_rl_gym_b200 import synthetic
    from unitree_rl_gym_b200.envspec import build_env_spec
    from unitree_rl_gym_b200.robots import ROBOTS
    case = "tmp"
    H.ENV_CASES[case] = (task, n, False)
    try:
        spec = H.make_spec(case)
        frames = synthetic.make_frames(ROBOTS[task], None, n, 3, spec.decimation, seed=5, feet_idx=spec.feet_idx,
                                       pen_idx=spec.pen_idx, term_idx=spec.term_idx, init_z=float(spec.base_init_state[2]),
                                       default_dof_pos=torch.from_numpy(spec.default_dof_pos))
        env = make_env(case, frames=frames)
    finally:
        del H.ENV_CASES[case]
    port = EnvPort(spec)
    calls = {"n": 0}

    def physics(sub):
        if sub is not None:
            calls["n"] += 1
        port.load_frame(frames, ((calls["n"] - 1) // spec.decimation) % 3, sub)
    env.reset()
    port.reset(physics)
    g = torch.Generator().manual_seed(3)
    ep = torch.randint(0, 1001, (n,), generator=g)
    env.episode_length_buf = ep.cuda()
    port.episode_length_buf = ep.clone()
    for t in range(4):
        a = torch.randn(n, spec.num_dof, generator=g)
        env.step(a.cuda())
        port.step(a, physics)
        snap = env_snapshot(env)
        H.compare("reset_buf", snap["reset_buf"], port.reset_buf.numpy(), f"step{t}/")
        H.compare("time_out_buf", snap["time_out_buf"], port.time_out_buf.numpy(), f"step{t}/")
        H.compare("episode_length_buf", snap["episode_length_buf"], port.episode_length_buf.numpy(), f"step{t}/")
        for k in ("obs_buf", "rew_buf", "torques", "commands", "last_root_vel", "feet_air_time"):
            H.compare(k, snap[k], getattr(port, k).numpy(), f"step{t}/")
        if port.privileged_obs_buf is not None:
            H.compare("privileged_obs_buf", snap["privileged_obs_buf"], port.privileged_obs_buf.numpy(), f"step{t}/")
        for k, v in port.episode_sums.items():
            H.compare("episode_sums/" + k, snap["episode_sums/" + k], v.numpy(), f"step{t}/")
        for k, v in port.extras["episode"].items():
            H.compare("extras_episode/" + k, snap["extras_episode/" + k], v.reshape(1).numpy(), f"step{t}/")
    assert int(port.reset_buf.sum()) >= 0


def test_unknown_controller_raises_nameerror():
    from unitree_rl_gym_b200.legged_gym.envs import LeggedRobot
    from unitree_rl_gym_b200.legged_gym.utils.helpers import SimParams
    cfg = H.make_cfg("go2", 32, True)
    cfg.control.control_type = "X"
    sp = SimParams()
    sp.dt = cfg.sim.dt
    with pytest.raises(NameError):
        LeggedRobot(cfg=cfg, sim_params=sp, physics_engine=None, sim_device="cuda:0", headless=True)
