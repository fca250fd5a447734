/* urlg_b200.h — C ABI of the B200-native hot path of unitree-rl-gym.
 *
 * The reference has no FFI: its seams are Python duck-typed APIs above (LeggedRobot /
 * OnPolicyRunner / PPO / ActorCritic[Recurrent] / RolloutStorage) and the Isaac Gym tensor
 * API below.  This ABI sits between them; each entry point cites the reference code
 * (file:line under /root/reference, or [UPSTREAM rsl_rl v1.0.2] where the submodule is
 * absent) that it replaces.  The reference-side binding is the ctypes stub shown in
 * INTEGRATION.md.
 *
 * Conventions
 *  - plain C types only; every large buffer is BORROWED device memory (allocated by the
 *    caller, e.g. PyTorch); the library owns only its handle structs and small scratch.
 *  - every call enqueues on the given cudaStream_t (passed as void*), never synchronises,
 *    never reads device memory on the host: a whole rollout is CUDA-graph capturable.
 *  - return 0 on success, negative urlg_status otherwise; urlg_last_error() has the text.
 *  - "step" arguments come as (const int64_t* base_dev, int64_t offset): the effective
 *    value is (*base_dev if non-NULL else 0) + offset, so a captured graph can advance.
 *  - bool tensors are uint8 0/1 (torch.bool layout); quaternions are xyzw.
 */
#ifndef URLG_B200_H
#define URLG_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define URLG_ABI_VERSION 3
#define URLG_MAX_DOF 16
#define URLG_MAX_IDX 16
#define URLG_MAX_TERMS 32
#define URLG_MAX_OBS 128
#define URLG_MAX_FEET 4

typedef enum {
  URLG_OK = 0,
  URLG_EINVAL = -1,   /* bad argument / unsupported shape */
  URLG_ECUDA = -2,    /* a CUDA runtime call failed */
  URLG_ENOMEM = -3
} urlg_status;

/* Reward-term ids: alphabetical by reward NAME = the reference's execution order (helpers.py:15 dir()).  A name whose
 * formula a subclass overrides has two ids (base / _HS = go2_handstand_env.py); a config activates at most one of them. */
typedef enum {
  URLG_T_ACTION_RATE = 0, URLG_T_ALIVE, URLG_T_ANG_VEL_XY, URLG_T_BASE_HEIGHT, URLG_T_BASE_HEIGHT_HS, URLG_T_COLLISION,
  URLG_T_CONTACT, URLG_T_CONTACT_NO_VEL, URLG_T_DOF_ACC, URLG_T_DOF_POS_LIMITS, URLG_T_DOF_VEL, URLG_T_DOF_VEL_LIMITS,
  URLG_T_ENERGY, URLG_T_FEET_AIR_TIME, URLG_T_FEET_CONTACT_FORCES, URLG_T_FEET_SWING_HEIGHT, URLG_T_FRONT_FEET_CONTACT,
  URLG_T_FRONT_FEET_TOGETHER, URLG_T_FRONT_HIP_NEUTRAL, URLG_T_HIND_FEET_NO_CONTACT, URLG_T_HIP_POS, URLG_T_LIN_VEL_XY,
  URLG_T_LIN_VEL_Z, URLG_T_ORIENTATION, URLG_T_ORIENTATION_HS, URLG_T_POSE, URLG_T_STABILITY, URLG_T_STAND_STILL,
  URLG_T_STAY_STILL, URLG_T_STUMBLE, URLG_T_TORQUE_LIMITS, URLG_T_TORQUES, URLG_T_TRACKING_ANG_VEL,
  URLG_T_TRACKING_LIN_VEL,
  URLG_T_COUNT
} urlg_reward_term;

/* Env class variants (what the reference expresses by subclassing LeggedRobot). */
enum { URLG_FAMILY_BASE = 0,        /* LeggedRobot: 12+3A obs (legged_robot.py:789-811) */
       URLG_FAMILY_H1 = 1,          /* H1Robot / H1_2Robot / G1Robot: gait phase, 9+3A+2 obs + privileged obs (h1_env.py) */
       URLG_FAMILY_HANDSTAND = 2 }; /* GO2HandstandEnv: 6+3A+4 obs with contact flags (go2_handstand_env.py:140-176) */
enum { URLG_TERMINATION_RPY = 0,      /* contact | |pitch|>1 | |roll|>0.8 | time-out  (legged_robot.py:711-721) */
       URLG_TERMINATION_GRAVITY = 1 };/* contact | -0.2 < projected_gravity.z < 0.2 | time-out (go2_handstand_env.py:178-219) */
enum { URLG_CALLBACK_BASE = 0,      /* resample commands + heading command (legged_robot.py:488-517) */
       URLG_CALLBACK_NONE = 1 };    /* an override that never calls the base callback (go2_handstand_env.py:84-96) */

/* Init-time constants of one env instance: what LeggedRobot._parse_cfg / _init_buffers /
 * _process_dof_props / _prepare_reward_function / _get_noise_scale_vec compute once
 * (legged_robot.py:52-67, 69-186, 442-470, 817-840, 188-219; h1_env.py:10-32). */
typedef struct urlg_env_desc {
  int32_t family;            /* URLG_FAMILY_* */
  int32_t num_envs, num_dof, num_bodies, num_obs, num_priv;
  int32_t num_feet, num_pen, num_term_bodies, num_hip;
  int32_t decimation, control_type;          /* control_type: 'P','V','T' as ASCII */
  int32_t resample_steps, push_interval;
  int32_t push_robots, heading_command, add_noise, only_positive, has_termination, send_timeouts;
  int32_t num_terms;                          /* active terms, in order */
  int32_t num_sums;                           /* episode_sums rows (num_terms [+1 for termination]) */
  int32_t termination_sum_row;                /* row of 'termination' in episode_sums, or -1 */
  int32_t term_ids[URLG_MAX_TERMS];
  int32_t term_sum_row[URLG_MAX_TERMS];       /* episode_sums row of each active term */
  float term_scales[URLG_MAX_TERMS];          /* scale * dt */
  float termination_scale;
  float dt, sim_dt, max_episode_length, max_episode_length_s, max_push_vel;
  float action_scale, clip_actions, clip_obs;
  float tracking_sigma, base_height_target, soft_dof_vel_limit, soft_torque_limit, max_contact_force;
  float obs_scale_lin_vel, obs_scale_ang_vel, obs_scale_dof_pos, obs_scale_dof_vel;
  float cmd_x[2], cmd_y[2], cmd_yaw[2], cmd_heading[2];
  float p_gains[URLG_MAX_DOF], d_gains[URLG_MAX_DOF], default_dof_pos[URLG_MAX_DOF];
  float torque_limits[URLG_MAX_DOF], dof_vel_limits[URLG_MAX_DOF];
  float dof_pos_lo[URLG_MAX_DOF], dof_pos_hi[URLG_MAX_DOF];
  int32_t feet_idx[URLG_MAX_IDX], pen_idx[URLG_MAX_IDX], term_idx[URLG_MAX_IDX], hip_idx[URLG_MAX_IDX];
  float noise_scale_vec[URLG_MAX_OBS];
  float base_init_state[13];
  uint32_t seed_lo, seed_hi;
  int64_t env_offset;                         /* global id of local env 0 (multi-GPU shard) */
  /* ---- variant tables (ABI v2): what go2_handstand_env.py changes by overriding methods ---- */
  int32_t termination_mode;                   /* URLG_TERMINATION_* */
  int32_t callback_mode;                      /* URLG_CALLBACK_* */
  int32_t clip_torques;                       /* 1: clip to +-torque_limits (legged_robot.py:671); 0: none (go2_handstand_env.py:349-350) */
  int32_t num_front_dof;                      /* handstand pose term: DOFs [0, num_front_dof) weigh 1.0, the rest 0.3 (:277-307) */
  float pd_target_lo[URLG_MAX_DOF];           /* P control: target = clamp(action*scale + default, lo, hi); +-inf = no clamp */
  float pd_target_hi[URLG_MAX_DOF];           /*   (rear-leg clamp of go2_handstand_env.py:336-346) */
  float pose_targets[URLG_MAX_DOF];           /* handstand pose (go2_handstand_config.py handstand_pose) */
} urlg_env_desc;

/* Device pointers of every public tensor of the env (legged_robot.py:104-186, base_task.py:41-49,
 * h1_env.py:37-42).  All float32 unless noted; shapes in comments; N = num_envs. */
typedef struct urlg_env_buffers {
  float* root_states;              /* [N,13]   physics seam */
  float* dof_state;                /* [N*A,2]  physics seam, interleaved (pos,vel) */
  const float* contact_forces;     /* [N,B,3]  physics seam */
  const float* rigid_body_states;  /* [N,B,13] physics seam, H1 family only (else NULL) */
  float* obs_buf[2];               /* [N,num_obs] ping-pong (reference re-binds obs_buf every step) */
  float* priv_obs_buf[2];          /* [N,num_priv] or NULL */
  float* rew_buf;                  /* [N] */
  uint8_t* reset_buf;              /* [N] bool */
  uint8_t* time_out_buf;           /* [N] bool */
  int64_t* episode_length_buf;     /* [N] int64 */
  float* torques;                  /* [N,A] */
  float* actions;                  /* [N,A] */
  float* last_actions;             /* [N,A] */
  float* last_dof_vel;             /* [N,A] */
  float* last_root_vel;            /* [N,6] */
  float* commands;                 /* [N,4] */
  float* feet_air_time;            /* [N,F] */
  uint8_t* last_contacts;          /* [N,F] bool */
  float* rpy;                      /* [N,3] */
  float* base_lin_vel;             /* [N,3] */
  float* base_ang_vel;             /* [N,3] */
  float* projected_gravity;        /* [N,3] */
  float* phase;                    /* [N]   H1 family (else NULL) */
  float* leg_phase;                /* [N,2] H1 family (else NULL) */
  float* episode_sums;             /* [num_sums,N] one row per dict entry of episode_sums */
  const float* env_origins;        /* [N,3] */
  float* extras_episode;           /* [num_sums] extras["episode"]["rew_*"], refreshed only when >=1 env reset */
  uint8_t* extras_time_outs;       /* [N] bool, extras["time_outs"], same staleness rule (legged_robot.py:742,767) */
  int32_t* reset_ids;              /* [N] compacted ids of envs reset this step: argument of set_dof_state_tensor_indexed
                                      (legged_robot.py:569-572); order is not defined; may be NULL */
  int32_t* reset_count;            /* [1] */
  int32_t* root_ids;               /* [N] envs whose root state was rewritten this step = reset envs + push-due envs:
                                      argument of set_actor_root_state_tensor_indexed (:552-555, 591-594); may be NULL */
  int32_t* root_count;             /* [1] */
  const float* rng_override;       /* optional [N,num_rng_slots] uniforms replacing Philox (debug), else NULL */
} urlg_env_buffers;

typedef struct urlg_env urlg_env;

int urlg_version(void);
const char* urlg_last_error(void);
/* Number of kernels this library has launched (or recorded into a CUDA graph under capture) since load:
 * bench.py's gpu_launches evidence. */
int64_t urlg_launch_count(void);

/* ---- env half: replaces LeggedRobot.step minus PhysX (legged_robot.py:615-811) ---- */
int urlg_env_create(const urlg_env_desc* desc, urlg_env** out);
int urlg_env_bind(urlg_env* env, const urlg_env_buffers* bufs);
void urlg_env_destroy(urlg_env* env);
int urlg_env_num_rng_slots(const urlg_env* env);

/* K1: _compute_torques (legged_robot.py:649-671).  If raw_actions != NULL this is the first
 * substep: actions = clip(raw_actions, +-clip_actions) is stored first (legged_robot.py:623-624). */
int urlg_pd_torques(urlg_env* env, const float* raw_actions, void* stream);

/* K2 (+finalize): post_physics_step (legged_robot.py:673-709) + the obs clips of step()
 * (:643-646).  step = common_step_counter AFTER the increment at :682.  obs_slot selects the
 * ping-pong buffer written. */
int urlg_post_physics(urlg_env* env, const int64_t* step_base, int64_t step_offset, int obs_slot, void* stream);

/* ---- pool-indexed physics seam (bench / rollout graphs) ----
 * Instead of copying the next synthetic frame into root_states / dof_state / contact_forces / rigid_body_states before
 * every kernel (urlg_synth_refresh = the stand-in for PhysX filling its tensors), the kernels can read the current frame
 * straight out of a device-resident pool: frame index = ((*calls_base + calls_offset - 1) / substeps) % num_frames,
 * dof substep = (*calls_base + calls_offset - 1) % substeps, calls = number of simulate() calls so far (<= 0: the bound
 * tensors are read).  Resets / pushes are still written to the bound root_states / dof_state tensors.  Results are
 * identical to the copy path (tests/test_env_cuda.py::test_pool_seam_matches_copy_seam). */
typedef struct urlg_frame_pool {
  const float* root;          /* [P, N, 13] */
  const float* dof;           /* [P, substeps, N*A, 2] */
  const float* contact;       /* [P, N*B, 3] */
  const float* rigid_body;    /* [P, N*B, 13] or NULL */
  int64_t num_frames, substeps;
  const int64_t* calls_base;  /* device int64 [1] */
  /* 0: every tensor is its own dense [P, ...] array (above).  > 0: PACKED frames -- frame f of every tensor starts
   * frame_stride floats after frame f - 1 (root / dof / contact / rigid_body then point at frame 0's sections of one
   * [P, frame_stride] buffer), so a host-side physics engine uploads a whole frame with ONE contiguous copy. */
  int64_t frame_stride;
} urlg_frame_pool;
int urlg_env_bind_frame_pool(urlg_env* env, const urlg_frame_pool* pool /* NULL: unbind */);
int urlg_pd_torques_at(urlg_env* env, const float* raw_actions, int64_t calls_offset, void* stream);
int urlg_post_physics_at(urlg_env* env, const int64_t* step_base, int64_t step_offset, int obs_slot,
                         int64_t calls_offset, void* stream);

/* reset_idx(env_ids) called from outside step() (base_task.py:82-86, legged_robot.py:723-768).
 * env_ids: device int64 [n], or NULL for "all envs". */
int urlg_env_reset_idx(urlg_env* env, const int64_t* env_ids, int64_t n, const int64_t* step_base,
                       int64_t step_offset, void* stream);

/* ---- synthetic physics stand-in for the Isaac Gym tensor seam (bench / tests) ---- */
/* dst[i] = pool[frame * elems + i], frame = ((*base + offset) / divisor) % num_frames * frames_per + ((*base+offset) % divisor) * sub
 * i.e. refresh_*_tensor (legged_robot.py:639,678-679; h1_env.py:49) served from a device-resident pool. */
int urlg_synth_refresh(float* dst, const float* pool, int64_t elems, int64_t num_frames, int64_t substeps,
                       int per_substep, const int64_t* calls_base, int64_t calls_offset, void* stream);

/* dst rows ids[0 .. *count) <- the same rows of src (row_floats floats each; *count read on the device, at most max_rows):
 * what set_dof_state_tensor_indexed / set_actor_root_state_tensor_indexed do inside a simulator whose state lives apart
 * from the exposed tensors (legged_robot.py:552-555, 569-572, 591-594), without a host read of the count. */
int urlg_scatter_rows(float* dst, const float* src, const int32_t* ids, const int32_t* count, int row_floats,
                      int max_rows, void* stream);

/* n asynchronous copies (cudaMemcpyDefault) enqueued on `stream` in one call: the per-env-step host <-> device traffic
 * of a physics engine that lives on the host (state frame in, rewards / dones out) without n trips through the caller's
 * language runtime.  Host pointers should be pinned. */
int urlg_copy_batch(void* const* dst, const void* const* src, const int64_t* bytes, int n, void* stream);

/* Philox4x32-10 uniforms exactly as the kernels draw them (for tests): out[N,num_slots]. */
int urlg_philox_uniform(float* out, int64_t n, int64_t num_slots, uint32_t seed_lo, uint32_t seed_hi,
                        uint32_t stream_id, int64_t env_offset, int64_t step, void* stream);

/* ---- tensor-core building block of the PPO half ----
 * D[M,N] = sum_k A[m,k]*B[n,k] on tcgen05 (bf16 operands via TMA, fp32 accumulate in TMEM) with a fused
 * epilogue.  Replaces the cuBLAS GEMMs behind nn.Linear forward/backward in ActorCritic
 * ([UPSTREAM rsl_rl v1.0.2] modules/actor_critic.py; reference call sites task_registry.py:118-119).
 * a_mn/b_mn = 0: operand is row-major [rows, K] (K-major); = 1: row-major [K, rows] (MN-major). */
#define URLG_GEMM_ELU 1      /* out = ELU(acc + bias) */
#define URLG_GEMM_ATOMIC 2   /* out_f32 += acc (split-K accumulation; caller zeroes out_f32) */
typedef struct urlg_gemm_desc {
  const void* A; int64_t lda;     /* bf16, ld in elements (row pitch must be a multiple of 16 bytes) */
  const void* B; int64_t ldb;
  int32_t a_mn, b_mn;
  int32_t M, N, K;
  int32_t block_n;                /* 0 = auto */
  int32_t k_splits;               /* 0/1 = none */
  float* out_f32; int32_t ld_f32; /* optional fp32 output [M,N] */
  void* out_bf16; int32_t ld_bf16;/* optional bf16 output [M,N] */
  const float* bias;              /* optional [N] */
  const void* dact_ref; int32_t ld_ref;   /* optional bf16 [M,N]: out *= ELU'(y), y = post-activation */
  float* colsum;                  /* optional [N]: += column sums of the stored values */
  int32_t flags;
  /* split-bf16 ("bf16x3") mode, enabled when A_lo and B_lo are non-NULL: every operand is a (hi, lo) pair with
   * x ~= hi + lo (lo = bf16(x - hi), same layout / ld as hi) and D = A_hi*B_hi + A_lo*B_hi + A_hi*B_lo, i.e. about
   * 16 mantissa bits per operand on the bf16 tensor pipe (fp32 accumulate): what the 1e-3 PPO parity needs. */
  const void* A_lo; const void* B_lo;
  void* out_bf16_lo;              /* optional low part of out_bf16 */
  const void* dact_ref_lo;        /* optional low part of dact_ref */
  int32_t max_ctas;               /* 0 = one persistent CTA per SM; else cap the grid (two GEMMs side by side on two streams) */
  /* Order-independent accumulation (with URLG_GEMM_ATOMIC): when out_fx / colsum_fx are set the values are added as
   * 64-bit fixed point (value * 2^URLG_FX_SHIFT, rounded to nearest) with integer atomics instead of fp32 atomics into
   * out_f32 / colsum: integer addition is associative, so split-K sums are bit-reproducible run to run.  Same shapes /
   * leading dimension as out_f32 / colsum. */
  int64_t* out_fx;
  int64_t* colsum_fx;
  /* Split-K without atomics: when out_part is set, split z of the K range stores its fp32 partial tile into
   * out_part[z][M][ld_f32] (plain stores; k_splits > 1 allowed without URLG_GEMM_ATOMIC) and the caller sums the
   * urlg_tc_gemm_splits(K, k_splits) slabs in split order -- bit-reproducible like out_fx, without the 64-bit atomics.
   * Takes precedence over out_f32 / out_fx. */
  float* out_part;
} urlg_gemm_desc;
#define URLG_FX_SHIFT 40          /* gradient accumulators: resolution 2^-40 = 9.1e-13, range +-8.4e6 */
#define URLG_FX_STAT_SHIFT 30     /* loss / KL sums: resolution 9.3e-10, range +-8.6e9 */
int urlg_tc_gemm(const urlg_gemm_desc* desc, void* stream);
/* The number of K splits a launch with this K and requested k_splits really uses (splits cover whole 64-wide k-blocks). */
int urlg_tc_gemm_splits(int K, int k_splits);
/* n (1..4) independent problems in ONE launch: their tiles form one persistent tile space, so e.g. the same layer of the
 * actor and of the critic, or a layer's data and weight gradients, share a launch's fixed cost (~4 us) and fill each
 * other's wave tails.  All problems must use the same operand mode (hi+lo pairs or plain bf16); results are those of n
 * separate urlg_tc_gemm calls. */
int urlg_tc_gemm_group(const urlg_gemm_desc* descs, int n, void* stream);
/* ---- data-parallel gradient exchange over peer memory (NVLink 5 / NVSwitch) ----
 * No counterpart in the reference (it is single-GPU); this is the path's one real exchange step when it is sharded over
 * GPUs: per minibatch the flat gradient + the four loss / KL sums become their sum over ranks before clip_grad_norm_ / Adam
 * ([UPSTREAM rsl_rl v1.0.2] algorithms/ppo.py update).  One-shot all-reduce: every rank owns an exchange
 * buffer that all peers have mapped (symmetric memory); a call publishes `data` into it, signals the peers' flag rows, waits
 * for theirs and replaces data[i] by the sum over ranks in rank order (bit-identical on every rank, reproducible).
 * Two launches on `stream`, no host synchronisation, capturable in a CUDA graph.  Every rank must make the same sequence
 * of calls.  Layout of one rank's symmetric allocation, in floats: [2 * count exchange halves | URLG_PEER_MAX_WORLD flag
 * words] (mode 1: see below), zero-filled before the first call (and a barrier between the fill and the first call). */
#define URLG_PEER_MAX_WORLD 16
typedef struct urlg_peer_exchange {
  int32_t world, rank;
  float* const* bufs;        /* DEVICE array [world]: peer-mapped pointer to rank r's exchange halves */
  uint32_t* const* flags;    /* DEVICE array [world]: peer-mapped pointer to rank r's flag row */
  int64_t count;             /* floats per half (multiple of 4, >= n) */
  uint32_t* epoch;           /* local device word, zero at start: exchanges completed */
  uint32_t* ticket;          /* local device words [2], zero at start */
  /* 0: one-shot (every rank reads all halves: world x n floats in, best at 2 ranks).  1: two-shot, fused form only
   * (urlg_policy_bind_peer_exchange): rank r sums slice r of the halves and stores it into every rank's result area, then
   * every rank reads its own result area -- 2 (world - 1) / world x n floats per rank, what 4 and 8 ranks need.  Needs the
   * larger allocation: [2 * count halves | count result area | 2 * URLG_PEER_MAX_WORLD flag words]; `flags[r]` points at
   * the first flag word either way. */
  int32_t mode;
  int32_t reserved;
} urlg_peer_exchange;
int urlg_peer_allreduce(const urlg_peer_exchange* x, float* data, int64_t n, void* stream);
/* The same exchange fused into the update: with an exchange bound to the policy (halves of >= num_params + 4 floats; the
 * stats block must lie directly behind the flat gradients), urlg_ppo_minibatch_grads publishes [gradient | KL, surrogate,
 * value, entropy sums] from the kernel that finalises the gradient, and urlg_ppo_apply sums the ranks' halves inside its
 * gradient-norm kernel: the all-reduce adds NO launch to a minibatch step.  Call urlg_ppo_minibatch_grads and urlg_ppo_apply
 * strictly alternately on every rank while bound.  NULL unbinds. */
struct urlg_policy;
int urlg_policy_bind_peer_exchange(struct urlg_policy* policy, const urlg_peer_exchange* x);

/* Measurement hook (bench.py roofline): while enabled, every tensor-core launch of this library (GEMM or fused forward)
 * is bracketed by CUDA events on its stream (do not enable under graph capture); read returns the summed duration, the
 * launch count, the algorithmic flops (2*M*N*K, counted once) and the algorithmic HBM bytes (every operand and output
 * element once).  One launch per bracket, outputs untouched: a timed pass computes exactly what an untimed one does. */
int urlg_gemm_timing(int enable);
int urlg_gemm_timing_read(double* total_ms, int64_t* launches, double* flops, double* bytes);

/* ======================================================================================
 * PPO half: rsl_rl v1.0.2 [UPSTREAM, submodule absent from /root/reference] — reference call
 * sites legged_gym/utils/task_registry.py:8-9,118-126; scripts/train.py:14; play.py:34,39;
 * utils/helpers.py:150-189; config keys envs/base/legged_robot_config.py:244-286.
 * ====================================================================================== */
#define URLG_MAX_LAYERS 8

/* One MLP: Linear(dims[0],dims[1]) ELU ... Linear(dims[L-1],dims[L]) (no activation after the last).
 * w_off/b_off: element offsets of weight [out,in] (row-major, nn.Linear layout) and bias [out]
 * inside the flat fp32 parameter buffer (same offsets in the flat gradient / Adam buffers). */
typedef struct urlg_mlp_desc {
  int32_t num_layers;
  int32_t dims[URLG_MAX_LAYERS + 1];
  int64_t w_off[URLG_MAX_LAYERS], b_off[URLG_MAX_LAYERS];
} urlg_mlp_desc;

/* ActorCritic / ActorCriticRecurrent ([UPSTREAM] modules/actor_critic.py, actor_critic_recurrent.py;
 * shapes from legged_robot_config.py:247-255, h1_config.py:103-115, g1_config.py:91-105). */
typedef struct urlg_policy_desc {
  int32_t recurrent;                 /* 0: MLP on obs; 1: LSTM(obs->rnn_hidden) then MLP on the LSTM output */
  int32_t num_actor_obs, num_critic_obs, num_actions, rnn_hidden;
  urlg_mlp_desc actor, critic;
  int64_t std_off;
  int64_t lstm_a_off[4], lstm_c_off[4];   /* weight_ih_l0 [4H,in], weight_hh_l0 [4H,H], bias_ih_l0, bias_hh_l0 */
  int64_t num_params;
  int32_t max_rows;                  /* largest row count of any call (rollout N or minibatch rows) */
  int32_t max_seq;                   /* recurrent: largest T of an update call */
  uint32_t seed_lo, seed_hi;         /* Philox key of the action-sampling stream */
  int64_t env_offset;                /* global id of local env 0 */
  int32_t precision;                 /* 0: bf16 operands; 1: split bf16 (hi+lo, 3 MMAs, ~fp32 operand precision) */
} urlg_policy_desc;

/* RolloutStorage tensors ([UPSTREAM] storage/rollout_storage.py), time-major [T,N,...], fp32 unless noted. */
typedef struct urlg_storage {
  int32_t T, N, num_obs, num_priv, num_actions, rnn_hidden;
  float* observations;               /* [T,N,num_obs] */
  float* privileged_observations;    /* [T,N,num_priv] or NULL (critic then reads observations) */
  float* actions; float* mu; float* sigma;          /* [T,N,A] */
  float* rewards; float* values; float* returns; float* advantages; float* actions_log_prob;   /* [T,N,1] */
  uint8_t* dones;                    /* [T,N,1] uint8 */
  float* hid_a_h; float* hid_a_c; float* hid_c_h; float* hid_c_c;   /* saved hidden states [T,1,N,H] or NULL */
} urlg_storage;

typedef struct urlg_ppo_hparams {
  float clip_param, value_loss_coef, entropy_coef, max_grad_norm, desired_kl;
  float beta1, beta2, eps;
  float lr_min, lr_max;
  int32_t use_clipped_value_loss, adaptive_lr;
  float grad_scale;                  /* 1/world_size: applied to the (all-reduced) gradient sum */
  float kl_count_scale;              /* world_size: rows contributing to the all-reduced KL sum = rows * this */
} urlg_ppo_hparams;

/* Device-resident scalars of the update (float[URLG_STAT_COUNT]); nothing is read back per step. */
enum {
  /* the four per-minibatch sums come first: a data-parallel caller that lays the stats block directly behind the flat
   * gradient buffer all-reduces [grads | these four] in ONE collective per minibatch */
  URLG_STAT_KL_SUM = 0,    /* this minibatch: sum over rows of KL */
  URLG_STAT_SURR_SUM, URLG_STAT_VLOSS_SUM, URLG_STAT_ENT_SUM,   /* this minibatch: sums over rows */
  URLG_STAT_LR,            /* current learning rate (adaptive schedule runs on the device) */
  URLG_STAT_STEP,          /* Adam step count */
  URLG_STAT_GNORM_SQ,      /* this minibatch: squared gradient norm (after grad_scale) */
  URLG_STAT_CLIP_COEF,
  URLG_STAT_ACC_VLOSS, URLG_STAT_ACC_SURR, URLG_STAT_ACC_N,      /* accumulated minibatch means since last reset */
  URLG_STAT_KL_MEAN,       /* last minibatch's KL mean (what the LR rule saw) */
  URLG_STAT_BC1, URLG_STAT_BC2,   /* Adam bias corrections 1-beta^step of the current step */
  URLG_STAT_COUNT = 16
};

typedef struct urlg_policy urlg_policy;

/* The library borrows params/grads/exp_avg/exp_avg_sq (flat fp32 [num_params]), stats
 * (float[URLG_STAT_COUNT]) and a workspace of urlg_policy_workspace_bytes() bytes. */
int64_t urlg_policy_workspace_bytes(const urlg_policy_desc* desc);
int urlg_policy_create(const urlg_policy_desc* desc, float* params, float* grads, float* exp_avg, float* exp_avg_sq,
                       float* stats, void* workspace, int64_t workspace_bytes, urlg_policy** out);
void urlg_policy_destroy(urlg_policy* p);
/* Refresh the bf16 operand images of the weights from the fp32 parameters (after load_state_dict etc.). */
int urlg_policy_sync_weights(urlg_policy* p, void* stream);

/* PPO.act ([UPSTREAM] algorithms/ppo.py act): actor forward, Normal(mu, std).sample() with Philox
 * (or mu + std*noise when noise != NULL: parity injection), critic forward, log-prob; writes the transition
 * straight into storage slabs [t] (observations, privileged_observations, actions, mu, sigma, values,
 * actions_log_prob, and for a recurrent policy the pre-step hidden states) and actions_out [N,A]. */
int urlg_policy_act(urlg_policy* p, const float* obs, const float* critic_obs, const urlg_storage* s, int t,
                    float* hidden /* recurrent: [4,N,H] = a_h,a_c,c_h,c_c, updated in place */,
                    const float* noise, float* actions_out, const int64_t* step_base, int64_t step_offset,
                    void* stream);
/* act_inference / evaluate on their own: out = actor(obs) [rows,A] (which=0) or critic(obs) [rows,1] (which=1). */
int urlg_policy_forward(urlg_policy* p, int which, const float* x, int64_t rows, float* hidden_h, float* hidden_c,
                        float* out, void* stream);

/* PPO.process_env_step: rewards[t] = rew + gamma * values[t] * time_outs (if non-NULL), dones[t] = dones,
 * recurrent: hidden[..., dones, :] = 0. */
int urlg_process_env_step(const urlg_storage* s, int t, const float* rew, const uint8_t* dones,
                          const uint8_t* time_outs, float gamma, float* hidden, void* stream);

/* RolloutStorage.compute_returns: GAE reverse scan, advantages = returns - values, then
 * (adv - mean) / (std_unbiased + 1e-8) over all T*N.  moments: double[3] device scratch (sum, sumsq, count);
 * phase 1 = scan + local moments, phase 2 = normalise (an all-reduce of `moments` may go between), 3 = both. */
int urlg_gae(const urlg_storage* s, const float* last_values, float gamma, float lam, double* moments, int phase,
             void* stream);

/* Allocation-free variant of urlg_gae (phase 3: scan + normalise) for callers that cannot call cudaMalloc or hand in
 * scratch (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It has NO counterpart
 * in the reference.  The moments scratch (per-block partial sums
 * + the folded moments) is a file-scope `static __device__` array of the library, 64.25 KB, held by every GPU that loads
 * the library until exit; its device address is resolved with cudaGetSymbolAddress once per device on the first call
 * made with that device current.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 * Limit: s->N <= URLG_GAE_NOALLOC_MAX_ENVS, otherwise URLG_EINVAL with a message naming the limit.  Inputs and outputs
 * are the caller's, as in urlg_gae; results equal urlg_gae's up to the order of the double-precision moment sums.
 * All calls on a device share the one array (it keeps the last call's contents; every slot a call reads it has
 * written first, so nothing needs clearing): calls that overlap on the GPU (different streams, no ordering) race --
 * serialise them. */
#define URLG_GAE_NOALLOC_MAX_ENVS 1048576
int urlg_gae_noalloc(const urlg_storage* s, const float* last_values, float gamma, float lam, void* stream);

/* One PPO minibatch: gather rows `indices` (flattened t*N+env, device int64 [rows]), forward, loss,
 * backward -> flat grads (+ KL / loss sums in stats).  Recurrent policies take an env slice instead:
 * envs [env_begin, env_begin+env_count), all T steps, hidden state re-seeded from the saved states at
 * trajectory starts (equivalent to split_and_pad_trajectories + masks). */
int urlg_ppo_minibatch_grads(urlg_policy* p, const urlg_storage* s, const int64_t* indices, int64_t rows,
                             int32_t env_begin, int32_t env_count, const urlg_ppo_hparams* h, void* stream);
/* Adaptive-KL learning-rate rule, clip_grad_norm_, Adam, refresh of the bf16 weight images ([UPSTREAM rsl_rl v1.0.2]
 * algorithms/ppo.py update: the schedule block, nn.utils.clip_grad_norm_, optimizer.step()).  Call once after every
 * urlg_ppo_minibatch_grads with the same hparams.  Single rank (grad_scale == kl_count_scale == 1, no peer exchange
 * bound): the norm and the scalar logic (STAT_LR, STAT_STEP, STAT_CLIP_COEF, STAT_BC1/2, loss accumulators) are already
 * produced by the last kernel of urlg_ppo_minibatch_grads, and this call is the Adam launch alone; data-parallel callers
 * reduce the [grads | sums] block in between (or let this call do it: urlg_policy_bind_peer_exchange). */
int urlg_ppo_apply(urlg_policy* p, const urlg_ppo_hparams* h, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* URLG_B200_H */
