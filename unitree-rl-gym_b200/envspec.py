"""EnvSpec: every init-time constant of the env hot path, derived from a config exactly the
way the reference derives it, then baked into the CUDA library's constant block.

Derivations follow (file:line in /root/reference/legged_gym/envs/base/legged_robot.py):
  _parse_cfg :52-67 (dt, max_episode_length, push_interval), _process_dof_props :442-470
  (soft limits), _init_buffers :169-186 (gains by substring match), _prepare_reward_function
  :817-840 (drop zero scales, scale*=dt, alphabetical order from class_to_dict),
  _get_noise_scale_vec :188-219 and h1_env.py:10-32, _get_env_origins :258-272.
"""
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch

from .legged_gym.utils.helpers import class_to_dict

# Reward-term ids shared with include/urlg_b200.h (enum urlg_reward_term): alphabetical by reward NAME, like the
# execution order.  'termination' is applied after the positive clamp and is not in this list.  A name whose formula
# the handstand subclass overrides (go2_handstand_env.py:224-246) has a second id with the suffix '@hs'.
TERM_KEYS = [
    "action_rate", "alive", "ang_vel_xy", "base_height", "base_height@hs", "collision", "contact", "contact_no_vel",
    "dof_acc", "dof_pos_limits", "dof_vel", "dof_vel_limits", "energy", "feet_air_time", "feet_contact_forces",
    "feet_swing_height", "front_feet_contact", "front_feet_together", "front_hip_neutral", "hind_feet_no_contact",
    "hip_pos", "lin_vel_xy", "lin_vel_z", "orientation", "orientation@hs", "pose", "stability", "stand_still",
    "stay_still", "stumble", "torque_limits", "torques", "tracking_ang_vel", "tracking_lin_vel",
]
_TERM_KEY_ID = {n: i for i, n in enumerate(TERM_KEYS)}
H1_FAMILY_TERMS = {"alive", "contact", "contact_no_vel", "feet_swing_height", "hip_pos"}
HANDSTAND_TERMS = {"base_height", "orientation", "energy", "front_feet_contact", "front_feet_together",
                   "front_hip_neutral", "hind_feet_no_contact", "lin_vel_xy", "pose", "stability", "stay_still"}
TERM_NAMES = sorted({k.split("@")[0] for k in TERM_KEYS})      # every `_reward_<name>` the fused kernel implements

FAMILY_BASE = 0        # LeggedRobot observation layout (legged_robot.py:800-807)
FAMILY_H1 = 1          # H1/H1_2/G1 layout with gait phase + privileged obs (h1_env.py:68-95)
FAMILY_HANDSTAND = 2   # GO2HandstandEnv: 6+3A+4 obs with contact flags (go2_handstand_env.py:140-176)


def term_id(name, family):
    """Kernel id of reward `name` for an env family; raises AttributeError where the reference's
    getattr(self, '_reward_' + name) would (legged_robot.py:836)."""
    if family == FAMILY_HANDSTAND and name + "@hs" in _TERM_KEY_ID:
        return _TERM_KEY_ID[name + "@hs"]
    if name not in _TERM_KEY_ID:
        raise AttributeError(f"no reward term '_reward_{name}' in the fused step kernel "
                             f"(the reference fails at legged_robot.py:836 for unknown names too)")
    if name in H1_FAMILY_TERMS and family != FAMILY_H1:
        raise AttributeError(f"'_reward_{name}' exists only on the H1-family env classes (h1_env.py:98-123)")
    if name in HANDSTAND_TERMS and name not in ("base_height", "orientation") and family != FAMILY_HANDSTAND:
        raise AttributeError(f"'_reward_{name}' exists only on GO2HandstandEnv (go2_handstand_env.py:224-383)")
    return _TERM_KEY_ID[name]


# kept for callers of the round-1 name: id of a base-class / H1-family term
TERM_ID = {n: i for n, i in _TERM_KEY_ID.items() if "@" not in n}

HIP_POS_IDX = {"h1": [0, 1, 5, 6], "h1_2": [0, 2, 6, 8], "g1": [1, 2, 7, 8]}   # h1_env.py:123 etc.


@dataclass
class EnvSpec:
    task: str
    family: int
    num_envs: int
    num_dof: int
    num_bodies: int
    num_obs: int
    num_priv: int                 # 0 when privileged obs are absent
    sim_dt: float
    dt: float
    decimation: int
    max_episode_length: float     # np.ceil result (float, like the reference)
    max_episode_length_s: float
    resample_steps: int
    push_interval: int
    push_robots: bool
    max_push_vel: float
    action_scale: float
    clip_actions: float
    clip_obs: float
    control_type: str
    p_gains: np.ndarray
    d_gains: np.ndarray
    default_dof_pos: np.ndarray
    torque_limits: np.ndarray
    dof_vel_limits: np.ndarray
    dof_pos_limits: np.ndarray    # [A,2] soft limits
    feet_idx: List[int]
    pen_idx: List[int]
    term_idx: List[int]
    hip_idx: List[int]
    reward_names: List[str]       # active, alphabetical, without 'termination'
    reward_scales: List[float]    # scale * dt (python float, cast to f32 at use)
    termination_scale: Optional[float]   # scale*dt if 'termination' scale != 0 else None
    only_positive: bool
    tracking_sigma: float
    base_height_target: float
    soft_dof_vel_limit: float
    soft_torque_limit: float
    max_contact_force: float
    obs_scale_lin_vel: float
    obs_scale_ang_vel: float
    obs_scale_dof_pos: float
    obs_scale_dof_vel: float
    noise_scale_vec: np.ndarray
    add_noise: bool
    heading_command: bool
    cmd_x: List[float] = field(default_factory=list)
    cmd_y: List[float] = field(default_factory=list)
    cmd_yaw: List[float] = field(default_factory=list)
    cmd_heading: List[float] = field(default_factory=list)
    base_init_state: np.ndarray = None
    env_origins: np.ndarray = None
    seed: int = 1
    env_offset: int = 0           # global id of local env 0 (multi-GPU sharding)
    send_timeouts: bool = True
    dof_names: List[str] = field(default_factory=list)
    body_names: List[str] = field(default_factory=list)
    # ---- variant tables: what a subclass of LeggedRobot changes by overriding methods (go2_handstand_env.py) ----
    termination_mode: int = 0     # 0: roll/pitch rule (legged_robot.py:715-716); 1: projected-gravity band (:178-219)
    callback_mode: int = 0        # 0: base callback (resample / heading, :488-517); 1: none (:84-96)
    clip_torques: bool = True     # legged_robot.py:671; the handstand override has no clip (:349-350)
    pd_target_lo: np.ndarray = None   # [A] P-control target clamp (+-inf = none); rear-leg clamp of :336-346
    pd_target_hi: np.ndarray = None
    pose_targets: np.ndarray = None   # [A] handstand pose (:22-24, 52-57)
    num_front_dof: int = 0

    @property
    def num_feet(self):
        return len(self.feet_idx)

    @property
    def num_rng_slots(self):
        return 14 + self.num_dof + self.num_obs + (4 if self.family == FAMILY_HANDSTAND else 0)

    @property
    def sum_names(self):
        """episode_sums keys in the reference's dict order (legged_robot.py:839-840)."""
        names = list(self.reward_names)
        if self.termination_scale is not None:
            # 'termination' sits in alphabetical position inside reward_scales
            names = sorted(names + ["termination"])
        return names


def _f32(x):
    return float(np.float32(x))


def env_origins_grid(num_envs, spacing):
    cols = np.floor(np.sqrt(num_envs))
    rows = np.ceil(num_envs / cols)
    xx, yy = torch.meshgrid(torch.arange(rows), torch.arange(cols), indexing="ij")
    o = torch.zeros(num_envs, 3)
    o[:, 0] = spacing * xx.flatten()[:num_envs]
    o[:, 1] = spacing * yy.flatten()[:num_envs]
    return o.numpy()


def build_env_spec(task, cfg, robot, sim_dt, seed=1, env_offset=0, family=None) -> EnvSpec:
    A = robot.num_dof
    N = cfg.env.num_envs
    if family is None:
        family = (FAMILY_HANDSTAND if hasattr(cfg, "handstand_pose") else
                  FAMILY_BASE if cfg.env.num_privileged_obs is None else FAMILY_H1)
    if cfg.commands.num_commands != 4:
        raise ValueError("the fused step kernel reads commands as [N, 4] (legged_robot_config.py:92 num_commands = 4); "
                         f"got num_commands = {cfg.commands.num_commands}")
    dt = cfg.control.decimation * sim_dt
    max_len_s = cfg.env.episode_length_s
    max_len = float(np.ceil(max_len_s / dt))
    push_interval = int(np.ceil(cfg.domain_rand.push_interval_s / dt))
    resample_steps = int(cfg.commands.resampling_time / dt)

    # gains: substring match of cfg key in dof name; later keys overwrite earlier ones
    kp, kd, q0 = np.zeros(A, np.float32), np.zeros(A, np.float32), np.zeros(A, np.float32)
    for i, name in enumerate(robot.dof_names):
        q0[i] = cfg.init_state.default_joint_angles[name]
        for key in cfg.control.stiffness.keys():
            if key in name:
                kp[i] = cfg.control.stiffness[key]
                kd[i] = cfg.control.damping[key]

    lim = torch.zeros(A, 2)
    for i in range(A):
        lim[i, 0] = float(np.float32(robot.dof_lower[i]))
        lim[i, 1] = float(np.float32(robot.dof_upper[i]))
        m = (lim[i, 0] + lim[i, 1]) / 2
        r = lim[i, 1] - lim[i, 0]
        lim[i, 0] = m - 0.5 * r * cfg.rewards.soft_dof_pos_limit
        lim[i, 1] = m + 0.5 * r * cfg.rewards.soft_dof_pos_limit

    bodies = robot.body_names
    feet = [i for i, s in enumerate(bodies) if cfg.asset.foot_name in s]
    pen = []
    for key in cfg.asset.penalize_contacts_on:
        pen.extend(i for i, s in enumerate(bodies) if key in s)
    term = []
    for key in cfg.asset.terminate_after_contacts_on:
        term.extend(i for i, s in enumerate(bodies) if key in s)

    scales = class_to_dict(cfg.rewards.scales)
    names, vals, term_scale = [], [], None
    for key, s in scales.items():          # alphabetical (dir() order)
        if s == 0:
            continue
        if key == "termination":
            term_scale = s * dt
            continue
        term_id(key, family)             # AttributeError for a name the env class has no _reward_ function for
        names.append(key)
        vals.append(s * dt)

    os_, ns, lvl = cfg.normalization.obs_scales, cfg.noise.noise_scales, cfg.noise.noise_level
    O = cfg.env.num_observations
    nv = np.zeros(O, np.float32)
    if family == FAMILY_BASE:
        nv[0:3] = ns.lin_vel * lvl * os_.lin_vel
        nv[3:6] = ns.ang_vel * lvl * os_.ang_vel
        nv[6:9] = ns.gravity * lvl
        nv[12:12 + A] = ns.dof_pos * lvl * os_.dof_pos
        nv[12 + A:12 + 2 * A] = ns.dof_vel * lvl * os_.dof_vel
    elif family == FAMILY_H1:
        nv[0:3] = ns.ang_vel * lvl * os_.ang_vel
        nv[3:6] = ns.gravity * lvl
        nv[9:9 + A] = ns.dof_pos * lvl * os_.dof_pos
        nv[9 + A:9 + 2 * A] = ns.dof_vel * lvl * os_.dof_vel
    else:                                    # go2_handstand_env.py:386-422 (gyro, gravity, dof_pos, dof_vel; flags 0)
        nv[0:3] = lvl * ns.gyro * os_.ang_vel
        nv[3:6] = lvl * ns.gravity
        nv[6:6 + A] = lvl * ns.dof_pos * os_.dof_pos
        nv[6 + A:6 + 2 * A] = lvl * ns.dof_vel * os_.dof_vel

    variant = {}
    hip_idx = HIP_POS_IDX.get(robot.name, [])
    if family == FAMILY_HANDSTAND:
        # go2_handstand_env.py:17-42: pose targets from cfg.handstand_pose over the default angles; the six rear-leg
        # joints' PD targets are clamped to pose +- 0.15 rad; front hips for _reward_front_hip_neutral (:352-358)
        idx = {n: i for i, n in enumerate(robot.dof_names)}
        pose = q0.copy()
        for name, ang in cfg.handstand_pose.joint_angles.items():
            if name in idx:
                pose[idx[name]] = ang
        lo = np.full(A, -np.inf, np.float32)
        hi = np.full(A, np.inf, np.float32)
        rear_limit = np.float32(0.15)
        for n_ in ("RL_hip_joint", "RL_thigh_joint", "RL_calf_joint", "RR_hip_joint", "RR_thigh_joint", "RR_calf_joint"):
            if n_ in idx:
                lo[idx[n_]] = np.float32(pose[idx[n_]]) - rear_limit
                hi[idx[n_]] = np.float32(pose[idx[n_]]) + rear_limit
        front = [idx[l + j] for l in ("FL_", "FR_") for j in ("hip_joint", "thigh_joint", "calf_joint")]
        if sorted(front) != list(range(len(front))):
            raise ValueError("handstand pose term expects the front-leg DOFs first (go2.urdf order FL, FR, RL, RR)")
        if len(feet) < 4:
            raise AssertionError("Expected at least 4 feet for Go2.")          # go2_handstand_env.py:19
        hip_idx = [idx["FL_hip_joint"], idx["FR_hip_joint"]]
        variant = dict(termination_mode=1, callback_mode=1, clip_torques=False, pd_target_lo=lo, pd_target_hi=hi,
                       pose_targets=pose.astype(np.float32), num_front_dof=len(front))
    else:
        variant = dict(pd_target_lo=np.full(A, -np.inf, np.float32), pd_target_hi=np.full(A, np.inf, np.float32),
                       pose_targets=np.zeros(A, np.float32))

    ist = cfg.init_state
    rng = cfg.commands.ranges
    return EnvSpec(
        task=task, family=family, num_envs=N, num_dof=A, num_bodies=len(bodies), num_obs=O,
        num_priv=cfg.env.num_privileged_obs or 0, sim_dt=sim_dt, dt=dt, decimation=cfg.control.decimation,
        max_episode_length=max_len, max_episode_length_s=max_len_s, resample_steps=resample_steps,
        push_interval=push_interval, push_robots=bool(cfg.domain_rand.push_robots),
        max_push_vel=cfg.domain_rand.max_push_vel_xy, action_scale=cfg.control.action_scale,
        clip_actions=cfg.normalization.clip_actions, clip_obs=cfg.normalization.clip_observations,
        control_type=cfg.control.control_type, p_gains=kp, d_gains=kd, default_dof_pos=q0,
        torque_limits=np.asarray(robot.dof_effort, np.float32),
        dof_vel_limits=np.asarray(robot.dof_velocity, np.float32), dof_pos_limits=lim.numpy().copy(),
        feet_idx=feet, pen_idx=pen, term_idx=term, hip_idx=hip_idx,
        reward_names=names, reward_scales=vals, termination_scale=term_scale,
        only_positive=bool(cfg.rewards.only_positive_rewards), tracking_sigma=cfg.rewards.tracking_sigma,
        base_height_target=cfg.rewards.base_height_target, soft_dof_vel_limit=cfg.rewards.soft_dof_vel_limit,
        soft_torque_limit=cfg.rewards.soft_torque_limit, max_contact_force=cfg.rewards.max_contact_force,
        obs_scale_lin_vel=os_.lin_vel, obs_scale_ang_vel=os_.ang_vel, obs_scale_dof_pos=os_.dof_pos,
        obs_scale_dof_vel=os_.dof_vel, noise_scale_vec=nv, add_noise=bool(cfg.noise.add_noise),
        heading_command=bool(cfg.commands.heading_command), cmd_x=list(rng.lin_vel_x), cmd_y=list(rng.lin_vel_y),
        cmd_yaw=list(rng.ang_vel_yaw), cmd_heading=list(rng.heading),
        base_init_state=np.asarray(ist.pos + ist.rot + ist.lin_vel + ist.ang_vel, np.float32),
        env_origins=env_origins_grid(N, cfg.env.env_spacing), seed=seed, env_offset=env_offset,
        send_timeouts=bool(cfg.env.send_timeouts), dof_names=list(robot.dof_names), body_names=list(bodies), **variant)
