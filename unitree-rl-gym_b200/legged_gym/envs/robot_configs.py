"""Per-robot env / PPO configs for the five registered tasks (data only).  Values follow
go2_config.py:3-53, h1_config.py:3-118, h1_2_config.py:4-130, g1_config.py:3-108, go2_handstand_config.py:4-219."""
from .base.legged_robot_config import LeggedRobotCfg, LeggedRobotCfgPPO

_ASSET = '{LEGGED_GYM_ROOT_DIR}/resources/robots/'


def _angles(sides, per_side, extra=()):
    d = {f"{s}_{j}": v for s in sides for j, v in per_side.items()}
    d.update({k: 0. for k in extra})
    return d


class _HumanoidRand(LeggedRobotCfg.domain_rand):
    randomize_friction = True; friction_range = [0.1, 1.25]
    randomize_base_mass = True; added_mass_range = [-1., 3.]
    push_robots = True; push_interval_s = 5; max_push_vel_xy = 1.5


class _HumanoidScales(LeggedRobotCfg.rewards.scales):
    tracking_lin_vel = 1.0; tracking_ang_vel = 0.5; lin_vel_z = -2.0; ang_vel_xy = -0.05
    orientation = -1.0; base_height = -10.0; dof_acc = -2.5e-7; feet_air_time = 0.0
    action_rate = -0.01; dof_pos_limits = -5.0; alive = 0.15; hip_pos = -1.0
    contact_no_vel = -0.2; feet_swing_height = -20.0; contact = 0.18


class _RecurrentPolicy:
    init_noise_std = 0.8; actor_hidden_dims = [32]; critic_hidden_dims = [32]; activation = 'elu'
    rnn_type = 'lstm'; rnn_hidden_size = 64; rnn_num_layers = 1


# ------------------------------------------------------------------ Go2
class GO2RoughCfg(LeggedRobotCfg):
    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 0.42]
        default_joint_angles = {f"{leg}_{j}_joint": v
                                for leg in ("FL", "RL", "FR", "RR")
                                for j, v in (("hip", 0.1 if leg[1] == "L" else -0.1),
                                             ("thigh", 0.8 if leg[0] == "F" else 1.),
                                             ("calf", -1.5))}

    class control(LeggedRobotCfg.control):
        control_type = 'P'; stiffness = {'joint': 20.}; damping = {'joint': 0.5}
        action_scale = 0.25; decimation = 4

    class asset(LeggedRobotCfg.asset):
        file = _ASSET + 'go2/urdf/go2.urdf'; name = "go2"; foot_name = "foot"
        penalize_contacts_on = ["thigh", "calf"]; terminate_after_contacts_on = ["base"]
        self_collisions = 1

    class rewards(LeggedRobotCfg.rewards):
        soft_dof_pos_limit = 0.9; base_height_target = 0.25

        class scales(LeggedRobotCfg.rewards.scales):
            torques = -0.0002; dof_pos_limits = -10.0


class GO2RoughCfgPPO(LeggedRobotCfgPPO):
    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01

    class runner(LeggedRobotCfgPPO.runner):
        run_name = ''; experiment_name = 'rough_go2'


# ------------------------------------------------------------------ Go2 handstand (go2_handstand_config.py)
def _leg_angles(hip, thigh, calf):
    """{leg}_{joint}_joint table from per-leg (FL, FR, RL, RR) values."""
    legs = ("FL", "FR", "RL", "RR")
    return {f"{l}_{j}_joint": v[i] for j, v in (("hip", hip), ("thigh", thigh), ("calf", calf)) for i, l in enumerate(legs)}


class GO2HandstandCfg(LeggedRobotCfg):
    class env(LeggedRobotCfg.env):
        num_observations = 46; num_privileged_obs = 46; num_actions = 12

    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 0.28]
        default_joint_angles = _leg_angles((0.3, -0.3, 0.1, -0.1), (1.2, 1.2, 0.3, 0.3), (-2.0, -2.0, -1.4, -1.4))

    class control(LeggedRobotCfg.control):
        control_type = 'P'; stiffness = {'joint': 35.0}; damping = {'joint': 0.5}
        action_scale = 0.3; decimation = 5

    class asset(LeggedRobotCfg.asset):
        file = _ASSET + 'go2/urdf/go2.urdf'; name = "go2"; foot_name = "foot"
        penalize_contacts_on = ["thigh", "calf", "hip"]
        terminate_after_contacts_on = ["base", "thigh", "calf", "hip", "Head_lower", "Head_upper"]
        self_collisions = 1

    class noise(LeggedRobotCfg.noise):
        add_noise = True; noise_level = 1.0

        class noise_scales(LeggedRobotCfg.noise.noise_scales):
            gyro = 0.2; gravity = 0.05; dof_pos = 0.01; dof_vel = 1.5; lin_vel = 0.1; actions = 0.0

    class rewards(LeggedRobotCfg.rewards):
        only_positive_rewards = False; base_height_target = 0.65; soft_dof_pos_limit = 0.9

        class scales(LeggedRobotCfg.rewards.scales):
            orientation = 5.0; base_height = 2.0; front_feet_contact = 3.0; hind_feet_no_contact = 3.0
            pose = 10.0; stability = 1.0; stay_still = 2.0; lin_vel_xy = -3
            lin_vel_z = -0.001; ang_vel_xy = -0.005; torques = -2e-4; dof_vel = -1e-5; dof_acc = 0.0
            action_rate = -0.006; dof_pos_limits = -1.0; dof_vel_limits = 0.0; torque_limits = 0.0
            collision = -2.0; front_hip_neutral = 2.0; front_feet_together = 1.5; feet_contact_forces = 0.0
            termination = -5.0; tracking_lin_vel = 0.0; tracking_ang_vel = 0.0; feet_air_time = 0.0
            stumble = 0.0; energy = 0.0

    class handstand_pose:
        joint_angles = _leg_angles((0.0, 0.0, 0.0, 0.0), (-0.89, -0.89, 1.7, 1.7), (-1.5, -1.5, -1.853, -1.853))

    class footstand_pose:
        joint_angles = _leg_angles((0.0, 0.0, 0.0, 0.0), (0.82, 0.82, 1.82, 1.82), (-1.6, -1.68, -1.16, -1.16))


class GO2HandstandCfgPPO(LeggedRobotCfgPPO):
    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01

    class runner(LeggedRobotCfgPPO.runner):
        run_name = ''; experiment_name = 'go2_handstand'


# ------------------------------------------------------------------ H1
class H1RoughCfg(LeggedRobotCfg):
    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 1.0]
        default_joint_angles = _angles(
            ("left", "right"),
            {"hip_yaw_joint": 0., "hip_roll_joint": 0, "hip_pitch_joint": -0.1, "knee_joint": 0.3,
             "ankle_joint": -0.2, "shoulder_pitch_joint": 0., "shoulder_roll_joint": 0,
             "shoulder_yaw_joint": 0., "elbow_joint": 0.}, extra=("torso_joint",))

    class env(LeggedRobotCfg.env):
        num_observations = 41; num_privileged_obs = 44; num_actions = 10

    class domain_rand(_HumanoidRand):
        pass

    class control(LeggedRobotCfg.control):
        control_type = 'P'
        stiffness = {'hip_yaw': 150, 'hip_roll': 150, 'hip_pitch': 150, 'knee': 200, 'ankle': 40,
                     'torso': 300, 'shoulder': 150, "elbow": 100}
        damping = {'hip_yaw': 2, 'hip_roll': 2, 'hip_pitch': 2, 'knee': 4, 'ankle': 2,
                   'torso': 6, 'shoulder': 2, "elbow": 2}
        action_scale = 0.25; decimation = 4

    class asset(LeggedRobotCfg.asset):
        file = _ASSET + 'h1/urdf/h1.urdf'; name = "h1"; foot_name = "ankle"
        penalize_contacts_on = ["hip", "knee"]; terminate_after_contacts_on = ["pelvis"]
        self_collisions = 0; flip_visual_attachments = False

    class rewards(LeggedRobotCfg.rewards):
        soft_dof_pos_limit = 0.9; base_height_target = 1.05

        class scales(_HumanoidScales):
            collision = -1.0; torques = 0.0


class H1RoughCfgPPO(LeggedRobotCfgPPO):
    class policy(_RecurrentPolicy):
        pass

    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01

    class runner(LeggedRobotCfgPPO.runner):
        policy_class_name = "ActorCriticRecurrent"; max_iterations = 10000
        run_name = ''; experiment_name = 'h1'


# ------------------------------------------------------------------ G1
class G1RoughCfg(LeggedRobotCfg):
    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 0.8]
        default_joint_angles = _angles(
            ("left", "right"),
            {"hip_yaw_joint": 0., "hip_roll_joint": 0, "hip_pitch_joint": -0.1, "knee_joint": 0.3,
             "ankle_pitch_joint": -0.2, "ankle_roll_joint": 0}, extra=("torso_joint",))

    class env(LeggedRobotCfg.env):
        num_observations = 47; num_privileged_obs = 50; num_actions = 12

    class domain_rand(_HumanoidRand):
        pass

    class control(LeggedRobotCfg.control):
        control_type = 'P'
        stiffness = {'hip_yaw': 100, 'hip_roll': 100, 'hip_pitch': 100, 'knee': 150, 'ankle': 40}
        damping = {'hip_yaw': 2, 'hip_roll': 2, 'hip_pitch': 2, 'knee': 4, 'ankle': 2}
        action_scale = 0.25; decimation = 4

    class asset(LeggedRobotCfg.asset):
        file = _ASSET + 'g1_description/g1_12dof.urdf'; name = "g1"; foot_name = "ankle_roll"
        penalize_contacts_on = ["hip", "knee"]; terminate_after_contacts_on = ["pelvis"]
        self_collisions = 0; flip_visual_attachments = False

    class rewards(LeggedRobotCfg.rewards):
        soft_dof_pos_limit = 0.9; base_height_target = 0.78

        class scales(_HumanoidScales):
            dof_vel = -1e-3; collision = 0.0


class G1RoughCfgPPO(LeggedRobotCfgPPO):
    class policy(_RecurrentPolicy):
        pass

    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01

    class runner(LeggedRobotCfgPPO.runner):
        policy_class_name = "ActorCriticRecurrent"; max_iterations = 10000
        run_name = ''; experiment_name = 'g1'


# ------------------------------------------------------------------ H1_2
class H1_2RoughCfg(LeggedRobotCfg):
    class init_state(LeggedRobotCfg.init_state):
        pos = [0.0, 0.0, 1.05]
        default_joint_angles = _angles(
            ("left", "right"),
            {"hip_yaw_joint": 0, "hip_roll_joint": 0, "hip_pitch_joint": -0.16, "knee_joint": 0.36,
             "ankle_pitch_joint": -0.2, "ankle_roll_joint": 0.0, "shoulder_pitch_joint": 0.4,
             "shoulder_roll_joint": 0, "shoulder_yaw_joint": 0, "elbow_pitch_joint": 0.3},
            extra=("torso_joint",))

    class env(LeggedRobotCfg.env):
        num_observations = 47; num_privileged_obs = 50; num_actions = 12

    class control(LeggedRobotCfg.control):
        control_type = 'P'
        stiffness = {'hip_yaw_joint': 200., 'hip_roll_joint': 200., 'hip_pitch_joint': 200.,
                     'knee_joint': 300., 'ankle_pitch_joint': 40., 'ankle_roll_joint': 40.}
        damping = {'hip_yaw_joint': 2.5, 'hip_roll_joint': 2.5, 'hip_pitch_joint': 2.5,
                   'knee_joint': 4, 'ankle_pitch_joint': 2.0, 'ankle_roll_joint': 2.0}
        action_scale = 0.25; decimation = 8

    class sim(LeggedRobotCfg.sim):
        dt = 0.0025

    class domain_rand(_HumanoidRand):
        pass

    class asset(LeggedRobotCfg.asset):
        file = _ASSET + 'h1_2/h1_2_12dof.urdf'; name = "h1_2"; foot_name = "ankle_roll"
        penalize_contacts_on = ["hip", "knee"]; terminate_after_contacts_on = ["pelvis"]
        self_collisions = 0; flip_visual_attachments = False; armature = 1e-3

    class rewards(LeggedRobotCfg.rewards):
        soft_dof_pos_limit = 0.9; base_height_target = 1.0

        class scales(_HumanoidScales):
            dof_vel = -1e-3; collision = 0.0


class H1_2RoughCfgPPO(LeggedRobotCfgPPO):
    class policy(_RecurrentPolicy):
        pass

    class algorithm(LeggedRobotCfgPPO.algorithm):
        entropy_coef = 0.01

    class runner(LeggedRobotCfgPPO.runner):
        policy_class_name = "ActorCriticRecurrent"; max_iterations = 10000
        run_name = ''; experiment_name = 'h1_2'
