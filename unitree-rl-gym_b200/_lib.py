"""ctypes binding of liburlg_b200.so (C ABI in include/urlg_b200.h).

No CPU fallback: `lib()` raises if the shared library is missing, and every compute call needs
CUDA tensors.  `build()` compiles the library in-tree with nvcc for sm_100a.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("URLG_LIB_PATH") or os.path.join(CSRC, "liburlg_b200.so")     # override: A/B builds

MAX_DOF, MAX_IDX, MAX_TERMS, MAX_OBS, MAX_FEET = 16, 16, 32, 128, 4
ABI_VERSION = 3


class EnvDesc(C.Structure):
    _fields_ = [
        ("family", C.c_int32), ("num_envs", C.c_int32), ("num_dof", C.c_int32), ("num_bodies", C.c_int32),
        ("num_obs", C.c_int32), ("num_priv", C.c_int32),
        ("num_feet", C.c_int32), ("num_pen", C.c_int32), ("num_term_bodies", C.c_int32), ("num_hip", C.c_int32),
        ("decimation", C.c_int32), ("control_type", C.c_int32),
        ("resample_steps", C.c_int32), ("push_interval", C.c_int32),
        ("push_robots", C.c_int32), ("heading_command", C.c_int32), ("add_noise", C.c_int32),
        ("only_positive", C.c_int32), ("has_termination", C.c_int32), ("send_timeouts", C.c_int32),
        ("num_terms", C.c_int32), ("num_sums", C.c_int32), ("termination_sum_row", C.c_int32),
        ("term_ids", C.c_int32 * MAX_TERMS), ("term_sum_row", C.c_int32 * MAX_TERMS),
        ("term_scales", C.c_float * MAX_TERMS), ("termination_scale", C.c_float),
        ("dt", C.c_float), ("sim_dt", C.c_float), ("max_episode_length", C.c_float),
        ("max_episode_length_s", C.c_float), ("max_push_vel", C.c_float),
        ("action_scale", C.c_float), ("clip_actions", C.c_float), ("clip_obs", C.c_float),
        ("tracking_sigma", C.c_float), ("base_height_target", C.c_float), ("soft_dof_vel_limit", C.c_float),
        ("soft_torque_limit", C.c_float), ("max_contact_force", C.c_float),
        ("obs_scale_lin_vel", C.c_float), ("obs_scale_ang_vel", C.c_float), ("obs_scale_dof_pos", C.c_float),
        ("obs_scale_dof_vel", C.c_float),
        ("cmd_x", C.c_float * 2), ("cmd_y", C.c_float * 2), ("cmd_yaw", C.c_float * 2), ("cmd_heading", C.c_float * 2),
        ("p_gains", C.c_float * MAX_DOF), ("d_gains", C.c_float * MAX_DOF), ("default_dof_pos", C.c_float * MAX_DOF),
        ("torque_limits", C.c_float * MAX_DOF), ("dof_vel_limits", C.c_float * MAX_DOF),
        ("dof_pos_lo", C.c_float * MAX_DOF), ("dof_pos_hi", C.c_float * MAX_DOF),
        ("feet_idx", C.c_int32 * MAX_IDX), ("pen_idx", C.c_int32 * MAX_IDX), ("term_idx", C.c_int32 * MAX_IDX),
        ("hip_idx", C.c_int32 * MAX_IDX),
        ("noise_scale_vec", C.c_float * MAX_OBS), ("base_init_state", C.c_float * 13),
        ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32), ("env_offset", C.c_int64),
        ("termination_mode", C.c_int32), ("callback_mode", C.c_int32), ("clip_torques", C.c_int32),
        ("num_front_dof", C.c_int32),
        ("pd_target_lo", C.c_float * MAX_DOF), ("pd_target_hi", C.c_float * MAX_DOF),
        ("pose_targets", C.c_float * MAX_DOF),
    ]


_P = C.c_void_p


class EnvBuffers(C.Structure):
    _fields_ = [
        ("root_states", _P), ("dof_state", _P), ("contact_forces", _P), ("rigid_body_states", _P),
        ("obs_buf", _P * 2), ("priv_obs_buf", _P * 2),
        ("rew_buf", _P), ("reset_buf", _P), ("time_out_buf", _P), ("episode_length_buf", _P),
        ("torques", _P), ("actions", _P), ("last_actions", _P), ("last_dof_vel", _P), ("last_root_vel", _P),
        ("commands", _P), ("feet_air_time", _P), ("last_contacts", _P),
        ("rpy", _P), ("base_lin_vel", _P), ("base_ang_vel", _P), ("projected_gravity", _P),
        ("phase", _P), ("leg_phase", _P), ("episode_sums", _P), ("env_origins", _P),
        ("extras_episode", _P), ("extras_time_outs", _P), ("reset_ids", _P), ("reset_count", _P),
        ("root_ids", _P), ("root_count", _P), ("rng_override", _P),
    ]


class FramePool(C.Structure):
    _fields_ = [("root", _P), ("dof", _P), ("contact", _P), ("rigid_body", _P),
                ("num_frames", C.c_int64), ("substeps", C.c_int64), ("calls_base", _P), ("frame_stride", C.c_int64)]


class GemmDesc(C.Structure):
    _fields_ = [
        ("A", _P), ("lda", C.c_int64), ("B", _P), ("ldb", C.c_int64),
        ("a_mn", C.c_int32), ("b_mn", C.c_int32), ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32),
        ("block_n", C.c_int32), ("k_splits", C.c_int32),
        ("out_f32", _P), ("ld_f32", C.c_int32), ("out_bf16", _P), ("ld_bf16", C.c_int32),
        ("bias", _P), ("dact_ref", _P), ("ld_ref", C.c_int32), ("colsum", _P), ("flags", C.c_int32),
        ("A_lo", _P), ("B_lo", _P), ("out_bf16_lo", _P), ("dact_ref_lo", _P), ("max_ctas", C.c_int32),
        ("out_fx", _P), ("colsum_fx", _P), ("out_part", _P),
    ]


GEMM_ELU, GEMM_ATOMIC = 1, 2
PEER_MAX_WORLD = 16


class PeerExchange(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("bufs", _P), ("flags", _P), ("count", C.c_int64),
                ("epoch", _P), ("ticket", _P), ("mode", C.c_int32), ("reserved", C.c_int32)]
MAX_LAYERS = 8


class MlpDesc(C.Structure):
    _fields_ = [("num_layers", C.c_int32), ("dims", C.c_int32 * (MAX_LAYERS + 1)),
                ("w_off", C.c_int64 * MAX_LAYERS), ("b_off", C.c_int64 * MAX_LAYERS)]


class PolicyDesc(C.Structure):
    _fields_ = [("recurrent", C.c_int32), ("num_actor_obs", C.c_int32), ("num_critic_obs", C.c_int32),
                ("num_actions", C.c_int32), ("rnn_hidden", C.c_int32),
                ("actor", MlpDesc), ("critic", MlpDesc), ("std_off", C.c_int64),
                ("lstm_a_off", C.c_int64 * 4), ("lstm_c_off", C.c_int64 * 4), ("num_params", C.c_int64),
                ("max_rows", C.c_int32), ("max_seq", C.c_int32), ("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32),
                ("env_offset", C.c_int64), ("precision", C.c_int32)]


class Storage(C.Structure):
    _fields_ = [("T", C.c_int32), ("N", C.c_int32), ("num_obs", C.c_int32), ("num_priv", C.c_int32),
                ("num_actions", C.c_int32), ("rnn_hidden", C.c_int32),
                ("observations", _P), ("privileged_observations", _P), ("actions", _P), ("mu", _P), ("sigma", _P),
                ("rewards", _P), ("values", _P), ("returns", _P), ("advantages", _P), ("actions_log_prob", _P),
                ("dones", _P), ("hid_a_h", _P), ("hid_a_c", _P), ("hid_c_h", _P), ("hid_c_c", _P)]


class PpoHparams(C.Structure):
    _fields_ = [("clip_param", C.c_float), ("value_loss_coef", C.c_float), ("entropy_coef", C.c_float),
                ("max_grad_norm", C.c_float), ("desired_kl", C.c_float),
                ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("lr_min", C.c_float), ("lr_max", C.c_float),
                ("use_clipped_value_loss", C.c_int32), ("adaptive_lr", C.c_int32),
                ("grad_scale", C.c_float), ("kl_count_scale", C.c_float)]


(STAT_KL_SUM, STAT_SURR_SUM, STAT_VLOSS_SUM, STAT_ENT_SUM, STAT_LR, STAT_STEP, STAT_GNORM_SQ, STAT_CLIP_COEF,
 STAT_ACC_VLOSS, STAT_ACC_SURR, STAT_ACC_N, STAT_KL_MEAN, STAT_BC1, STAT_BC2) = range(14)
STAT_COUNT = 16

_lib = None

# symbol -> (restype, argtypes); every symbol declared in include/urlg_b200.h appears here
_I64P = C.c_void_p
SIGNATURES = {
    "urlg_version": (C.c_int, []),
    "urlg_last_error": (C.c_char_p, []),
    "urlg_launch_count": (C.c_int64, []),
    "urlg_env_create": (C.c_int, [C.POINTER(EnvDesc), C.POINTER(C.c_void_p)]),
    "urlg_env_bind": (C.c_int, [C.c_void_p, C.POINTER(EnvBuffers)]),
    "urlg_env_destroy": (None, [C.c_void_p]),
    "urlg_env_num_rng_slots": (C.c_int, [C.c_void_p]),
    "urlg_pd_torques": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "urlg_post_physics": (C.c_int, [C.c_void_p, _I64P, C.c_int64, C.c_int, C.c_void_p]),
    "urlg_env_bind_frame_pool": (C.c_int, [C.c_void_p, C.POINTER(FramePool)]),
    "urlg_pd_torques_at": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "urlg_post_physics_at": (C.c_int, [C.c_void_p, _I64P, C.c_int64, C.c_int, C.c_int64, C.c_void_p]),
    "urlg_scatter_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "urlg_env_reset_idx": (C.c_int, [C.c_void_p, _I64P, C.c_int64, _I64P, C.c_int64, C.c_void_p]),
    "urlg_copy_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "urlg_synth_refresh": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, _I64P,
                                     C.c_int64, C.c_void_p]),
    "urlg_tc_gemm": (C.c_int, [C.POINTER(GemmDesc), C.c_void_p]),
    "urlg_tc_gemm_group": (C.c_int, [C.POINTER(GemmDesc), C.c_int, C.c_void_p]),
    "urlg_tc_gemm_splits": (C.c_int, [C.c_int, C.c_int]),
    "urlg_peer_allreduce": (C.c_int, [C.POINTER(PeerExchange), C.c_void_p, C.c_int64, C.c_void_p]),
    "urlg_policy_bind_peer_exchange": (C.c_int, [C.c_void_p, C.POINTER(PeerExchange)]),
    "urlg_gemm_timing": (C.c_int, [C.c_int]),
    "urlg_gemm_timing_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_double),
                                       C.POINTER(C.c_double)]),
    "urlg_policy_workspace_bytes": (C.c_int64, [C.POINTER(PolicyDesc)]),
    "urlg_policy_create": (C.c_int, [C.POINTER(PolicyDesc), _P, _P, _P, _P, _P, _P, C.c_int64, C.POINTER(C.c_void_p)]),
    "urlg_policy_destroy": (None, [C.c_void_p]),
    "urlg_policy_sync_weights": (C.c_int, [C.c_void_p, C.c_void_p]),
    "urlg_policy_act": (C.c_int, [C.c_void_p, _P, _P, C.POINTER(Storage), C.c_int, _P, _P, _P, _I64P, C.c_int64,
                                  C.c_void_p]),
    "urlg_policy_forward": (C.c_int, [C.c_void_p, C.c_int, _P, C.c_int64, _P, _P, _P, C.c_void_p]),
    "urlg_process_env_step": (C.c_int, [C.POINTER(Storage), C.c_int, _P, _P, _P, C.c_float, _P, C.c_void_p]),
    "urlg_gae": (C.c_int, [C.POINTER(Storage), _P, C.c_float, C.c_float, _P, C.c_int, C.c_void_p]),
    "urlg_gae_noalloc": (C.c_int, [C.POINTER(Storage), _P, C.c_float, C.c_float, C.c_void_p]),
    "urlg_ppo_minibatch_grads": (C.c_int, [C.c_void_p, C.POINTER(Storage), _I64P, C.c_int64, C.c_int32, C.c_int32,
                                           C.POINTER(PpoHparams), C.c_void_p]),
    "urlg_ppo_apply": (C.c_int, [C.c_void_p, C.POINTER(PpoHparams), C.c_void_p]),
    "urlg_philox_uniform": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_uint32, C.c_uint32, C.c_uint32,
                                      C.c_int64, C.c_int64, C.c_void_p]),
}


def build(verbose=False):
    """Compile csrc/*.cu -> csrc/liburlg_b200.so (nvcc, -gencode arch=compute_100a,code=sm_100a -lineinfo)."""
    out = subprocess.run(["bash", os.path.join(CSRC, "build.sh")], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("liburlg_b200 build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)
    global _lib
    _lib = None
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback for the hot path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        if L.urlg_version() != ABI_VERSION:
            raise RuntimeError("liburlg_b200 ABI version mismatch")
        _lib = L
    return _lib


def check(code):
    if code != 0:
        msg = lib().urlg_last_error().decode()
        if code == -1:
            if msg == "Rollout buffer overflow":
                raise AssertionError(msg)  # same as [UPSTREAM] RolloutStorage.add_transitions
            if msg.startswith("Unknown controller type"):
                raise NameError(msg)       # same exception type as legged_robot.py:670
            if msg.startswith("Expected at least 4 feet"):
                raise AssertionError(msg)  # go2_handstand_env.py:19
            raise ValueError(msg)
        raise RuntimeError(f"urlg error {code}: {msg}")


def ptr(t):
    """Device pointer of a CUDA tensor (or None)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("urlg_b200: the hot path runs on CUDA tensors only (no CPU fallback)")
    return C.c_void_p(t.data_ptr())


class CopyBatch:
    """A fixed list of (dst, src) tensor copies enqueued on a stream by ONE urlg_copy_batch call (host tensors pinned):
    the per-env-step frame upload / reward read-back of a host-side physics engine without a Python call per tensor."""

    def __init__(self, pairs):
        n = len(pairs)
        for d, s_ in pairs:
            if d.numel() * d.element_size() != s_.numel() * s_.element_size() or not (d.is_contiguous() and s_.is_contiguous()):
                raise ValueError("CopyBatch: each pair must be two contiguous tensors of equal byte size")
        self.n = n
        self.dst = (C.c_void_p * n)(*[d.data_ptr() for d, _ in pairs])
        self.src = (C.c_void_p * n)(*[s_.data_ptr() for _, s_ in pairs])
        self.bytes = (C.c_int64 * n)(*[d.numel() * d.element_size() for d, _ in pairs])
        self.keep = pairs

    def enqueue(self, stream):
        check(lib().urlg_copy_batch(self.dst, self.src, self.bytes, self.n, C.c_void_p(stream.cuda_stream)))


def cur_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
