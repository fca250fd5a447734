"""Shared test plumbing: fixture loading, spec construction, comparison with stated tolerances."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ENV_CASES = {"go2_n64": ("go2", 64, True), "go2_n64_signed": ("go2", 64, False), "h1_n40": ("h1", 40, False),
             "g1_n64": ("g1", 64, True), "h1_2_n33": ("h1_2", 33, False),
             "go2_n48_allterms": ("go2", 48, False), "h1_2_n35_allterms": ("h1_2", 35, True),
             "go2_n40_ctrlV": ("go2", 40, True), "go2_n40_ctrlT": ("go2", 40, True),
             "go2_handstand_n64": ("go2_handstand", 64, False)}



# Config overrides of a case, applied identically to the reference's config object (tests/golden/gen_golden.py, before the
# unmodified reference env is constructed) and to this repo's config (make_cfg): the reward terms / control types that
# no shipped task switches on (legged_robot.py:882-884, 892-899, 925-939; :656-668).
def _tweak_allterms(cfg):
    sc = cfg.rewards.scales
    sc.dof_vel_limits = -1.0
    sc.torque_limits = -0.01
    sc.stumble = -0.5              # the config key `feet_stumble` has no function (quirk 7); `_reward_stumble` does
    sc.stand_still = -0.2
    sc.feet_contact_forces = -0.01
    sc.termination = -5.0
    cfg.rewards.soft_dof_vel_limit = 0.1       # 3.0 rad/s: a third of the synthetic dof_vel ~ N(0, 3^2) exceeds it
    cfg.rewards.soft_torque_limit = 0.5
    cfg.rewards.max_contact_force = 100.0
    cfg.commands.ranges.lin_vel_x = [-0.3, 0.3]   # a third of the resampled commands fall under the 0.2 norm cut
    cfg.commands.ranges.lin_vel_y = [-0.3, 0.3]   # -> zero command -> stand_still is active


def _tweak_ctrl_v(cfg):
    cfg.control.control_type = "V"


def _tweak_ctrl_t(cfg):
    cfg.control.control_type = "T"
    cfg.control.action_scale = 9.0             # scaled actions on both sides of the +-23.7 / 35.55 torque limits


TWEAKS = {"go2_n48_allterms": _tweak_allterms, "h1_2_n35_allterms": _tweak_allterms, "go2_n40_ctrlV": _tweak_ctrl_v,
          "go2_n40_ctrlT": _tweak_ctrl_t}

# north_star tolerances: masks / indices bit-exact; obs, rewards, torques 1e-5 relative (fp32)
RTOL = 1e-5
ATOL = 2e-6     # absolute floor for values near zero (sums of O(1) terms in fp32)
EXACT = {"reset_buf", "time_out_buf", "episode_length_buf", "last_contacts", "extras_time_outs"}


def load_case(name):
    z = np.load(os.path.join(GOLDEN, f"env_{name}.npz"))
    return {k: z[k] for k in z.files}


def make_cfg(task, num_envs, only_positive, tweak=None):
    from unitree_rl_gym_b200.legged_gym.envs import robot_configs as rc
    cls = {"go2": rc.GO2RoughCfg, "h1": rc.H1RoughCfg, "g1": rc.G1RoughCfg, "h1_2": rc.H1_2RoughCfg,
           "go2_handstand": rc.GO2HandstandCfg}[task]
    cfg = cls()
    cfg.env.num_envs = num_envs
    cfg.rewards.only_positive_rewards = only_positive
    if tweak is not None:
        tweak(cfg)
    return cfg


def make_spec(case):
    from unitree_rl_gym_b200.envspec import build_env_spec
    from unitree_rl_gym_b200.robots import ROBOTS
    task, n, pos = ENV_CASES[case]
    cfg = make_cfg(task, n, pos, TWEAKS.get(case))
    return build_env_spec(task, cfg, ROBOTS[task], cfg.sim.dt, seed=1)


def make_port(spec):
    """The oracle env for a spec: EnvPort, or HandstandPort for the handstand family."""
    from oracle.env_port import EnvPort
    if spec.family == 2:
        from oracle.handstand_port import HandstandPort
        return HandstandPort(spec)
    return EnvPort(spec)


def env_class(task):
    from unitree_rl_gym_b200.legged_gym import envs as E
    return {"go2": E.LeggedRobot, "h1": E.H1Robot, "h1_2": E.H1_2Robot, "g1": E.G1Robot,
            "go2_handstand": E.GO2HandstandEnv}[task]


# The handstand reference draws its 1 % contact-flag flips from torch's global generator (go2_handstand_env.py:165),
# which the harness cannot replay: against the REFERENCE fixture the four flag columns (and their privileged clones)
# are compared up to a bounded number of flips; against the oracle port (same Philox slots) they are exact.
def compare_handstand_obs(name, got, want, where=""):
    compare(name, got[:, :-4], want[:, :-4], where)
    assert set(np.unique(got[:, -4:])) <= {0.0, 1.0}, f"{where}{name}: contact flags must be 0/1"
    assert (got[:, -4:] != want[:, -4:]).mean() <= 0.06, f"{where}{name}: contact flags differ beyond the 1 % flips"


def frames_of(fx):
    return {k.split("/", 1)[1]: torch.from_numpy(v) for k, v in fx.items() if k.startswith("frames/")}


def compare(name, got, want, where=""):
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, f"{where}{name}: shape {got.shape} != {want.shape}"
    if name in EXACT or want.dtype == np.bool_ or np.issubdtype(want.dtype, np.integer):
        assert np.array_equal(got.astype(want.dtype), want), f"{where}{name}: mask/index mismatch at " \
            f"{np.argwhere(got.astype(want.dtype) != want)[:5].tolist()}"
    else:
        err = np.abs(got.astype(np.float64) - want.astype(np.float64))
        tol = ATOL + RTOL * np.abs(want.astype(np.float64))
        bad = err > tol
        assert not bad.any(), f"{where}{name}: {bad.sum()} of {bad.size} beyond rtol={RTOL} " \
            f"(max err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}, want {want.flat[err.argmax()]})"


def snapshot_keys(fx, prefix):
    p = prefix + "/"
    return [k[len(p):] for k in fx if k.startswith(p)]
