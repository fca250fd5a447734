// Env half of the hot path: LeggedRobot.step minus the PhysX call, as three kernels.
//   K1  pd_torques_kernel      _compute_torques                 legged_robot.py:649-671 (+ :623-624)
//   K2  post_physics_kernel    post_physics_step + obs clips    legged_robot.py:673-709, 643-646; h1_env.py:55-123
//   K2b finalize_step_kernel   _push_robots' all-env rewrite, last_root_vel, extras publication
//                                                               legged_robot.py:540-555, 709, 742-768
// Data layout: one warp owns 32 consecutive envs.  Row tensors ([N,13] root, [N,2A] dof_state,
// [N,3B] contacts, [N,A] action-like) are staged through shared memory with coalesced loads and
// odd row strides (bank-conflict-free thread-per-env access); [N] and [N,<=4] tensors are accessed
// directly (already coalesced).  Compiled with -fmad=false: expressions evaluate as written.
#include "common.cuh"
#include "philox.cuh"
#include <new>

namespace urlg {

thread_local char g_err[512] = "";
long long g_launches = 0;

constexpr int kMaxFeet = 8;
constexpr float kPi = 3.14159274101257324f;      // float32(np.pi)
constexpr float kTwoPi = 6.28318548202636719f;   // float32(2*np.pi)
constexpr float kHalfPi = 1.57079637050628662f;  // float32(np.pi/2)

struct StepScratch {   // library-owned device scratch (zeroed at create)
  float ep_acc[3][URLG_MAX_TERMS + 1];   // [slot][row] sums over reset envs, [slot][URLG_MAX_TERMS] = count
  int push_flag[3];
  int reset_cnt[3];
};

}  // namespace urlg

struct urlg_env {
  urlg_env_desc desc;
  urlg_env_buffers bufs;
  urlg::StepScratch* scratch;
  int bound;
  int smem_bytes;
  int vec_ok;      // every row tensor K2 stages / stores in bulk is 16-byte aligned (set at bind)
};

namespace urlg {

// ---------------------------------------------------------------- shared-memory layout (K2)
// One block = 32 consecutive envs x 4 lanes per env (128 threads).  The row tensors of those 32 envs are
// contiguous in global memory, so each is ONE 1-D bulk-TMA copy (cp.async.bulk) into an unpadded shared image;
// outputs with row structure (obs, privileged obs, last_actions, last_dof_vel) are assembled in shared memory and
// leave as flat 16-byte vector stores.
constexpr int kEnvsPerBlock = 32, kLanesPerEnv = 4, kK2Threads = kEnvsPerBlock * kLanesPerEnv;

struct Layout {
  int oRoot, oDof, oCon, oAct, oLact, oLdv, oTq, oRb, oObs, oPriv, oTerm, oBar, total;   // float offsets
};

__host__ __device__ inline Layout make_layout(const urlg_env_desc& c) {
  Layout L;
  const int A = c.num_dof, R = kEnvsPerBlock;
  int o = 0;
  L.oRoot = o; o += R * 13;
  L.oDof = o; o += R * 2 * A;
  L.oCon = o; o += R * 3 * c.num_bodies;
  L.oAct = o; o += R * A;
  L.oLact = o; o += R * A;
  L.oLdv = o; o += R * A;
  L.oTq = o; o += R * A;
  L.oRb = o; o += c.family == 1 ? R * c.num_feet * 13 : 0;
  o = (o + 3) & ~3;
  L.oObs = o; o += R * c.num_obs;
  L.oPriv = o; o += R * c.num_priv;
  o = (o + 3) & ~3;
  L.oTerm = o; o += R * URLG_MAX_TERMS;      // this step's scaled term values per env (episode_sums increments)
  L.oBar = o; o += 4;
  L.total = o;
  return L;
}

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_g2s(float* dst, const float* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_addr(dst)), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(float* dst, const float* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(reinterpret_cast<uint64_t>(dst)), "r"(smem_addr(src)), "r"(bytes) : "memory");
}
// Flat copy of `n` floats (tail / unaligned fallback).
static __device__ __noinline__ void flat_in(float* __restrict__ s, const float* __restrict__ g, int n, int n_pad, int tid) {
#pragma unroll 1
  for (int i = tid; i < n_pad; i += kK2Threads) s[i] = i < n ? g[i] : 0.0f;
}
static __device__ __noinline__ void flat_out(float* __restrict__ g, const float* __restrict__ s, int n, int tid) {
#pragma unroll 1
  for (int i = tid; i < n; i += kK2Threads) g[i] = s[i];
}
// Sum over the 4 lanes of an env; every lane gets the same bits (commutative pairwise adds).
__device__ __forceinline__ float qsum(float x) {
  x += __shfl_xor_sync(0xffffffffu, x, 1);
  x += __shfl_xor_sync(0xffffffffu, x, 2);
  return x;
}
// episode_length_buf % period (period > 0); the counters are small, so avoid the 64-bit division.
__device__ __forceinline__ bool divisible(int64_t len, int period) {
  if ((uint64_t)len <= 0x7fffffffull) return ((uint32_t)len % (uint32_t)period) == 0u;
  return (len % period) == 0;
}

// torch.remainder for floats: fmod then fix the sign (ATen BinaryOpsKernel remainder_kernel).
static __device__ __noinline__ float py_remainder(float a, float b) {
  float m = fmodf(a, b);
  if (m != 0.0f && ((b < 0.0f) != (m < 0.0f))) m += b;
  return m;
}

// One out-of-line copy each of the libm-heavy helpers: K2 is latency-bound at 4096 envs and an I-cache miss costs more
// than a call (ncu: 45% of stall samples were no_instructions with everything inlined, 10k SASS instructions).
static __device__ __noinline__ float atan2_ni(float y, float x) { return atan2f(y, x); }
__device__ __forceinline__ float norm3(float x, float y, float z) { return sqrtf(x * x + y * y + z * z); }

// [UPSTREAM isaacgym.torch_utils.quat_rotate_inverse], q = xyzw
__device__ __forceinline__ void rot_inv(float qx, float qy, float qz, float qw, float vx, float vy, float vz,
                                        float& ox, float& oy, float& oz) {
  const float s = 2.0f * (qw * qw) - 1.0f;
  const float cx = qy * vz - qz * vy, cy = qz * vx - qx * vz, cz = qx * vy - qy * vx;
  const float d = qx * vx + qy * vy + qz * vz;
  ox = vx * s - cx * qw * 2.0f + qx * d * 2.0f;
  oy = vy * s - cy * qw * 2.0f + qy * d * 2.0f;
  oz = vz * s - cz * qw * 2.0f + qz * d * 2.0f;
}

// ---------------------------------------------------------------- K1: PD torques
// One element: clip + store the action (first substep), PD law, torque-limit clip (legged_robot.py:649-671).
__device__ __forceinline__ float pd_one(const urlg_env_desc& c, int j, float a, float q, float qd, float last_qd) {
  const float as = a * c.action_scale;
  float tq;
  if (c.control_type == 'P') {
    tq = c.p_gains[j] * (as + c.default_dof_pos[j] - q) - c.d_gains[j] * qd;
  } else if (c.control_type == 'V') {
    tq = c.p_gains[j] * (as - qd) - c.d_gains[j] * (qd - last_qd) / c.sim_dt;
  } else {
    tq = as;
  }
  const float lim = c.torque_limits[j];
  return fminf(fmaxf(tq, -lim), lim);
}

// Vector form (num_dof % 4 == 0, 16-byte aligned tensors): one thread per 4 consecutive DOFs of an env -- one float4 of
// actions, two float4 of interleaved (pos, vel) pairs in, one float4 of torques out; one element per thread, no loop,
// so every load of the grid is in flight at once.
__global__ void __launch_bounds__(256) pd_torques_vec4_kernel(const __grid_constant__ urlg_env_desc c,
                                                              const __grid_constant__ urlg_env_buffers b,
                                                              const float* __restrict__ raw_actions, uint32_t total4) {
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= total4) return;
  const uint32_t quads = (uint32_t)c.num_dof >> 2;
  const int j0 = (int)(i % quads) * 4;
  float4 a;
  if (raw_actions) {
    a = __ldg(reinterpret_cast<const float4*>(raw_actions) + i);
    const float cl = c.clip_actions;
    a.x = fminf(fmaxf(a.x, -cl), cl); a.y = fminf(fmaxf(a.y, -cl), cl);
    a.z = fminf(fmaxf(a.z, -cl), cl); a.w = fminf(fmaxf(a.w, -cl), cl);
    reinterpret_cast<float4*>(b.actions)[i] = a;
  } else {
    a = reinterpret_cast<const float4*>(b.actions)[i];
  }
  const float4 s0 = reinterpret_cast<const float4*>(b.dof_state)[2 * (size_t)i];
  const float4 s1 = reinterpret_cast<const float4*>(b.dof_state)[2 * (size_t)i + 1];
  float4 l = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  if (c.control_type == 'V') l = reinterpret_cast<const float4*>(b.last_dof_vel)[i];
  float4 t;
  t.x = pd_one(c, j0, a.x, s0.x, s0.y, l.x);
  t.y = pd_one(c, j0 + 1, a.y, s0.z, s0.w, l.y);
  t.z = pd_one(c, j0 + 2, a.z, s1.x, s1.y, l.z);
  t.w = pd_one(c, j0 + 3, a.w, s1.z, s1.w, l.w);
  reinterpret_cast<float4*>(b.torques)[i] = t;
}

// Scalar form: one thread per (env, dof); actions / torques coalesced, dof_state read as float2.
__global__ void __launch_bounds__(256) pd_torques_kernel(const __grid_constant__ urlg_env_desc c,
                                                         const __grid_constant__ urlg_env_buffers b,
                                                         const float* __restrict__ raw_actions) {
  const uint32_t total = (uint32_t)c.num_envs * (uint32_t)c.num_dof;      // host checks num_envs * num_dof < 2^31
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int j = (int)(i % (uint32_t)c.num_dof);
    float a;
    if (raw_actions) {
      a = fminf(fmaxf(raw_actions[i], -c.clip_actions), c.clip_actions);
      b.actions[i] = a;
    } else {
      a = b.actions[i];
    }
    const float2 pv = reinterpret_cast<const float2*>(b.dof_state)[i];
    b.torques[i] = pd_one(c, j, a, pv.x, pv.y, c.control_type == 'V' ? b.last_dof_vel[i] : 0.0f);
  }
}

// ---------------------------------------------------------------- command resampling (E6)
static __device__ __noinline__ void resample_commands(const urlg_env_desc& c, SlotRng& rng, int slot, float* cmd) {
  cmd[0] = (c.cmd_x[1] - c.cmd_x[0]) * rng.uniform(slot) + c.cmd_x[0];
  cmd[1] = (c.cmd_y[1] - c.cmd_y[0]) * rng.uniform(slot + 1) + c.cmd_y[0];
  if (c.heading_command)
    cmd[3] = (c.cmd_heading[1] - c.cmd_heading[0]) * rng.uniform(slot + 2) + c.cmd_heading[0];
  else
    cmd[2] = (c.cmd_yaw[1] - c.cmd_yaw[0]) * rng.uniform(slot + 2) + c.cmd_yaw[0];
  const float keep = sqrtf(cmd[0] * cmd[0] + cmd[1] * cmd[1]) > 0.2f ? 1.0f : 0.0f;
  cmd[0] *= keep;
  cmd[1] *= keep;
}

}  // namespace urlg

namespace urlg {

// ---------------------------------------------------------------- K2: fused post-physics step
// Thread layout: tid = 4 * local_env + g.  Lane g of an env owns DOFs / bodies / feet / episode-sum rows
// g, g+4, g+8, ... and the observation columns [g*ceil(num_obs/4), ...); per-env scalars (frame transforms,
// commands, termination) are computed by all four lanes on identical inputs, so every lane holds identical bits
// and the quad never diverges on an env-level decision.
__global__ void __launch_bounds__(kK2Threads, 6) post_physics_kernel(const __grid_constant__ urlg_env_desc c,
                                                                 const __grid_constant__ urlg_env_buffers b,
                                                                 StepScratch* __restrict__ scratch,
                                                                 const int64_t* __restrict__ step_base, int64_t step_off,
                                                                 int obs_slot, int num_slots, int vec_ok) {
  extern __shared__ __align__(128) float sm[];
  const Layout L = make_layout(c);
  const int tid = threadIdx.x;
  const int env0 = blockIdx.x * kEnvsPerBlock;
  const int nrows = min(kEnvsPerBlock, c.num_envs - env0);
  const int A = c.num_dof, B = c.num_bodies, F = c.num_feet, N = c.num_envs;
  const int64_t step = eff_step(step_base, step_off);
  const int acc = (int)(step & 1);
  const bool bulk = vec_ok && nrows == kEnvsPerBlock;

  // ---- stage the row tensors of this block's 32 envs ----
  const uint32_t bar = smem_addr(sm + L.oBar);
  if (bulk) {
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      const uint32_t bytes = (uint32_t)(kEnvsPerBlock * (13 + 2 * A + 3 * B + 4 * A) * sizeof(float));
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
      const uint32_t rA = kEnvsPerBlock * A * sizeof(float);
      bulk_g2s(sm + L.oRoot, b.root_states + (size_t)env0 * 13, kEnvsPerBlock * 13 * sizeof(float), bar);
      bulk_g2s(sm + L.oDof, b.dof_state + (size_t)env0 * 2 * A, 2 * rA, bar);
      bulk_g2s(sm + L.oCon, b.contact_forces + (size_t)env0 * 3 * B, kEnvsPerBlock * 3 * B * sizeof(float), bar);
      bulk_g2s(sm + L.oAct, b.actions + (size_t)env0 * A, rA, bar);
      bulk_g2s(sm + L.oLact, b.last_actions + (size_t)env0 * A, rA, bar);
      bulk_g2s(sm + L.oLdv, b.last_dof_vel + (size_t)env0 * A, rA, bar);
      bulk_g2s(sm + L.oTq, b.torques + (size_t)env0 * A, rA, bar);
    }
  } else {
    flat_in(sm + L.oRoot, b.root_states + (size_t)env0 * 13, nrows * 13, kEnvsPerBlock * 13, tid);
    flat_in(sm + L.oDof, b.dof_state + (size_t)env0 * 2 * A, nrows * 2 * A, kEnvsPerBlock * 2 * A, tid);
    flat_in(sm + L.oCon, b.contact_forces + (size_t)env0 * 3 * B, nrows * 3 * B, kEnvsPerBlock * 3 * B, tid);
    flat_in(sm + L.oAct, b.actions + (size_t)env0 * A, nrows * A, kEnvsPerBlock * A, tid);
    flat_in(sm + L.oLact, b.last_actions + (size_t)env0 * A, nrows * A, kEnvsPerBlock * A, tid);
    flat_in(sm + L.oLdv, b.last_dof_vel + (size_t)env0 * A, nrows * A, kEnvsPerBlock * A, tid);
    flat_in(sm + L.oTq, b.torques + (size_t)env0 * A, nrows * A, kEnvsPerBlock * A, tid);
  }
  if (c.family == 1) {   // only the feet rows of rigid_body_states (h1_env.py:51-53): 52-byte rows, plain loads
    const int w = F * 13;
    for (int i = tid; i < kEnvsPerBlock * w; i += kK2Threads) {
      const int r = i / w, rem = i - r * w, f = rem / 13, k = rem - f * 13;
      sm[L.oRb + i] = r < nrows ? b.rigid_body_states[((size_t)(env0 + r) * B + c.feet_idx[f]) * 13 + k] : 0.0f;
    }
  }
  const int le = tid >> 2, g = tid & 3;
  const int e = env0 + le;
  const bool live = le < nrows;            // tail block: dead quads compute on zeros and store nothing
  // ---- per-env scalars straight from global, issued before the wait so their latency hides under the bulk copy ----
  int64_t len = (live ? b.episode_length_buf[e] : 0) + 1;              // :681
  float cmd[4] = {0.0f, 0.0f, 0.0f, 0.0f};
  if (live) {
    const float4 v = reinterpret_cast<const float4*>(b.commands)[e];
    cmd[0] = v.x; cmd[1] = v.y; cmd[2] = v.z; cmd[3] = v.w;
  }
  float air[2] = {0.0f, 0.0f};
  uint8_t lastc[2] = {0, 0};
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int f = g + 4 * q;
    if (f < F && live) {
      air[q] = b.feet_air_time[(size_t)e * F + f];
      lastc[q] = b.last_contacts[(size_t)e * F + f];
    }
  }
  float sums[URLG_MAX_TERMS / kLanesPerEnv];      // episode_sums rows g, g+4, ... of this env
#pragma unroll
  for (int q = 0; q < URLG_MAX_TERMS / kLanesPerEnv; ++q) {
    const int k = g + kLanesPerEnv * q;
    sums[q] = (live && k < c.num_sums) ? b.episode_sums[(size_t)k * N + e] : 0.0f;
  }
  __syncthreads();
  if (bulk) {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(bar) : "memory");
    }
  }

  float* root = sm + L.oRoot + le * 13;
  float* dof = sm + L.oDof + le * 2 * A;
  const float* con = sm + L.oCon + le * 3 * B;
  float* act = sm + L.oAct + le * A;
  float* lact = sm + L.oLact + le * A;
  float* ldv = sm + L.oLdv + le * A;
  const float* tq = sm + L.oTq + le * A;
  const float* rb = sm + L.oRb + le * F * 13;
  float* obs = sm + L.oObs + le * c.num_obs;
  float* priv = sm + L.oPriv + le * c.num_priv;
  float* term = sm + L.oTerm + le * URLG_MAX_TERMS;

  SlotRng rng;
  rng.init(c.seed_lo, c.seed_hi, kStreamEnv, c.env_offset + e, step,
           (b.rng_override && live) ? b.rng_override + (size_t)e * num_slots : nullptr);
  const float flen = (float)len;

  // ---- E4 frame transforms :685-690, isaacgym_utils.py:11-30 ----
  const float qx = root[3], qy = root[4], qz = root[5], qw = root[6];
  const float roll = atan2_ni(2.0f * (qw * qx + qy * qz), qw * qw - qx * qx - qy * qy + qz * qz);
  const float sinp = 2.0f * (qw * qy - qz * qx);
  const float pitch = fabsf(sinp) >= 1.0f ? copysignf(kHalfPi, sinp) : asinf(sinp);
  const float yaw = atan2_ni(2.0f * (qw * qz + qx * qy), qw * qw + qx * qx - qy * qy - qz * qz);
  float lvx, lvy, lvz, avx, avy, avz, gx, gy, gz;
  rot_inv(qx, qy, qz, qw, root[7], root[8], root[9], lvx, lvy, lvz);
  rot_inv(qx, qy, qz, qw, root[10], root[11], root[12], avx, avy, avz);
  rot_inv(qx, qy, qz, qw, 0.0f, 0.0f, -1.0f, gx, gy, gz);
  const float root_z = root[2];

  // ---- E5 callback: gait phase (h1_env.py:58-63), resample (:511-512), heading (:513-516) ----
  float phase = 0.0f, leg_phase[2] = {0.0f, 0.0f};
  if (c.family == 1) {
    phase = py_remainder(flen * c.dt, 0.8f) / 0.8f;
    leg_phase[0] = phase;
    leg_phase[1] = py_remainder(phase + 0.5f, 1.0f);
  }
  if (divisible(len, c.resample_steps)) resample_commands(c, rng, 0, cmd);
  if (c.heading_command) {
    // forward = quat_apply(q, (1,0,0)): t = 2 (u x v) = (0, 2 qz, -2 qy)
    const float ty = qz * 2.0f, tz = -qy * 2.0f;
    const float fx = 1.0f + qw * 0.0f + (qy * tz - qz * ty);
    const float fy = 0.0f + qw * ty + (qz * 0.0f - qx * tz);
    const float heading = atan2_ni(fy, fx);
    float a = py_remainder(cmd[3] - heading, kTwoPi);
    a -= kTwoPi * (a > kPi ? 1.0f : 0.0f);
    cmd[2] = fminf(fmaxf(0.5f * a, -1.0f), 1.0f);
  }

  // ---- E7 termination :715-721 ----
  float hits = 0.0f;
  _Pragma("unroll 1") for (int k = g; k < c.num_term_bodies; k += kLanesPerEnv) {
    const float* f3 = con + 3 * c.term_idx[k];
    hits += norm3(f3[0], f3[1], f3[2]) > 1.0f ? 1.0f : 0.0f;
  }
  bool reset = qsum(hits) > 0.0f;
  reset |= (fabsf(pitch) > 1.0f) || (fabsf(roll) > 0.8f);
  const bool time_out = flen > c.max_episode_length;
  reset |= time_out;

  // ---- E8/E9 rewards, alphabetical, :775-787.  Lane g owns feet g and g+4. ----
  const float cmd_xy = sqrtf(cmd[0] * cmd[0] + cmd[1] * cmd[1]);
  float rew = 0.0f;
  for (int t = 0; t < c.num_terms; ++t) {
    float r = 0.0f;
    switch (c.term_ids[t]) {
      case URLG_T_ACTION_RATE:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) { const float d = lact[j] - act[j]; r += d * d; }
        r = qsum(r);
        break;
      case URLG_T_ALIVE: r = 1.0f; break;
      case URLG_T_ANG_VEL_XY: r = avx * avx + avy * avy; break;
      case URLG_T_BASE_HEIGHT: { const float d = root_z - c.base_height_target; r = d * d; } break;
      case URLG_T_COLLISION:
        _Pragma("unroll 1") for (int k = g; k < c.num_pen; k += kLanesPerEnv) {
          const float* f3 = con + 3 * c.pen_idx[k];
          r += norm3(f3[0], f3[1], f3[2]) > 0.1f ? 1.0f : 0.0f;
        }
        r = qsum(r);
        break;
      case URLG_T_CONTACT:
        if (g < F && g < 2) {
          const bool stance = leg_phase[g] < 0.55f;
          const bool contact = con[3 * c.feet_idx[g] + 2] > 1.0f;
          r = (contact != stance) ? 0.0f : 1.0f;
        }
        r = qsum(r);
        break;
      case URLG_T_CONTACT_NO_VEL:
        _Pragma("unroll 1") for (int f = g; f < F; f += kLanesPerEnv) {
          const float* f3 = con + 3 * c.feet_idx[f];
          const float m = norm3(f3[0], f3[1], f3[2]) > 1.0f ? 1.0f : 0.0f;
          for (int k = 0; k < 3; ++k) { const float v = rb[f * 13 + 7 + k] * m; r += v * v; }
        }
        r = qsum(r);
        break;
      case URLG_T_DOF_ACC:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) { const float d = (ldv[j] - dof[2 * j + 1]) / c.dt; r += d * d; }
        r = qsum(r);
        break;
      case URLG_T_DOF_POS_LIMITS:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) {
          const float q = dof[2 * j];
          r += -fminf(q - c.dof_pos_lo[j], 0.0f) + fmaxf(q - c.dof_pos_hi[j], 0.0f);
        }
        r = qsum(r);
        break;
      case URLG_T_DOF_VEL:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) { const float v = dof[2 * j + 1]; r += v * v; }
        r = qsum(r);
        break;
      case URLG_T_DOF_VEL_LIMITS:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv)
          r += fminf(fmaxf(fabsf(dof[2 * j + 1]) - c.dof_vel_limits[j] * c.soft_dof_vel_limit, 0.0f), 1.0f);
        r = qsum(r);
        break;
      case URLG_T_FEET_AIR_TIME: {   // stateful, :912-923
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int f = g + 4 * q;
          if (f < F) {
            const bool contact = con[3 * c.feet_idx[f] + 2] > 1.0f;
            const bool filt = contact || lastc[q];
            lastc[q] = contact;
            const float first = (air[q] > 0.0f && filt) ? 1.0f : 0.0f;
            air[q] += c.dt;
            r += (air[q] - 0.5f) * first;
            air[q] *= filt ? 0.0f : 1.0f;
          }
        }
        r = qsum(r);
        r *= cmd_xy > 0.1f ? 1.0f : 0.0f;
      } break;
      case URLG_T_FEET_CONTACT_FORCES:
        _Pragma("unroll 1") for (int f = g; f < F; f += kLanesPerEnv) {
          const float* f3 = con + 3 * c.feet_idx[f];
          r += fmaxf(norm3(f3[0], f3[1], f3[2]) - c.max_contact_force, 0.0f);
        }
        r = qsum(r);
        break;
      case URLG_T_FEET_SWING_HEIGHT:
        _Pragma("unroll 1") for (int f = g; f < F; f += kLanesPerEnv) {
          const float* f3 = con + 3 * c.feet_idx[f];
          const float m = norm3(f3[0], f3[1], f3[2]) > 1.0f ? 0.0f : 1.0f;
          const float d = rb[f * 13 + 2] - 0.08f;
          r += d * d * m;
        }
        r = qsum(r);
        break;
      case URLG_T_HIP_POS:
        _Pragma("unroll 1") for (int k = g; k < c.num_hip; k += kLanesPerEnv) { const float q = dof[2 * c.hip_idx[k]]; r += q * q; }
        r = qsum(r);
        break;
      case URLG_T_LIN_VEL_Z: r = lvz * lvz; break;
      case URLG_T_ORIENTATION: r = gx * gx + gy * gy; break;
      case URLG_T_STAND_STILL:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) r += fabsf(dof[2 * j] - c.default_dof_pos[j]);
        r = qsum(r);
        r *= cmd_xy < 0.1f ? 1.0f : 0.0f;
        break;
      case URLG_T_STUMBLE: {
        _Pragma("unroll 1") for (int f = g; f < F; f += kLanesPerEnv) {
          const float* f3 = con + 3 * c.feet_idx[f];
          r += sqrtf(f3[0] * f3[0] + f3[1] * f3[1]) > 5.0f * fabsf(f3[2]) ? 1.0f : 0.0f;
        }
        r = qsum(r) > 0.0f ? 1.0f : 0.0f;
      } break;
      case URLG_T_TORQUE_LIMITS:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) r += fmaxf(fabsf(tq[j]) - c.torque_limits[j] * c.soft_torque_limit, 0.0f);
        r = qsum(r);
        break;
      case URLG_T_TORQUES:
        _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) r += tq[j] * tq[j];
        r = qsum(r);
        break;
      case URLG_T_TRACKING_ANG_VEL: { const float d = cmd[2] - avz; r = expf(-(d * d) / c.tracking_sigma); } break;
      case URLG_T_TRACKING_LIN_VEL: {
        const float dx = cmd[0] - lvx, dy = cmd[1] - lvy;
        r = expf(-(dx * dx + dy * dy) / c.tracking_sigma);
      } break;
      default: break;
    }
    r *= c.term_scales[t];
    rew += r;
    if (g == 0) term[c.term_sum_row[t]] = r;
  }
  if (c.only_positive) rew = fmaxf(rew, 0.0f);
  if (c.has_termination) {
    const float r = ((reset && !time_out) ? 1.0f : 0.0f) * c.termination_scale;
    rew += r;
    if (g == 0) term[c.termination_sum_row] = r;
  }

  // ---- E10 reset_idx for this env :723-768 (quad-uniform branch) ----
  if (reset) {
    if (live) {
      if (g == 0) {
        atomicAdd(&scratch->ep_acc[acc][URLG_MAX_TERMS], 1.0f);
        const int pos = atomicAdd(&scratch->reset_cnt[acc], 1);
        if (b.reset_ids) b.reset_ids[pos] = e;
      }
    }
    float* gdof = b.dof_state + (size_t)e * 2 * A;
    float* gact = b.actions + (size_t)e * A;
    _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) {                        // _reset_dofs :566-567, actions :752
      const float p = c.default_dof_pos[j] * ((1.5f - 0.5f) * rng.uniform(3 + j) + 0.5f);
      dof[2 * j] = p; dof[2 * j + 1] = 0.0f;
      act[j] = 0.0f;
      if (live) { gdof[2 * j] = p; gdof[2 * j + 1] = 0.0f; gact[j] = 0.0f; }
    }
    float* groot = b.root_states + (size_t)e * 13;                     // _reset_root_states :587-590
    _Pragma("unroll 1") for (int k = g; k < 13; k += kLanesPerEnv) {
      float v = c.base_init_state[k];
      if (k < 3 && live) v += b.env_origins[(size_t)e * 3 + k];
      if (k >= 7) v = (0.5f - (-0.5f)) * rng.uniform(3 + A + (k - 7)) + (-0.5f);
      if (live) groot[k] = v;
    }
    resample_commands(c, rng, 9 + A, cmd);                             // :749
    air[0] = 0.0f; air[1] = 0.0f;                                      // :755
    len = 0;                                                            // :756
  }
  const bool push_due = live && c.push_robots && divisible(len, c.push_interval);   // :545, evaluated after the reset

  // ---- episode_sums rows g, g+4, ...: += this step's term (:779-780, 786); reset envs publish and clear (:759-763) ----
  __syncwarp();
#pragma unroll
  for (int q = 0; q < URLG_MAX_TERMS / kLanesPerEnv; ++q) {
    const int k = g + kLanesPerEnv * q;
    if (live && k < c.num_sums) {
      float v = sums[q] + term[k];
      if (reset) { atomicAdd(&scratch->ep_acc[acc][k], v); v = 0.0f; }
      b.episode_sums[(size_t)k * N + e] = v;
    }
  }

  // ---- E12 observations :789-811 / h1_env.py:68-95 into the shared rows (raw), then noise + the clips of
  //      step() :643-646 over this lane's contiguous column range ----
  {
    const int o0 = c.family == 1 ? 0 : 3;          // column of ang_vel
    if (g == 0) {
      if (c.family == 0) {
        obs[0] = lvx * c.obs_scale_lin_vel; obs[1] = lvy * c.obs_scale_lin_vel; obs[2] = lvz * c.obs_scale_lin_vel;
      }
      obs[o0 + 0] = avx * c.obs_scale_ang_vel; obs[o0 + 1] = avy * c.obs_scale_ang_vel; obs[o0 + 2] = avz * c.obs_scale_ang_vel;
      obs[o0 + 3] = gx; obs[o0 + 4] = gy; obs[o0 + 5] = gz;
      obs[o0 + 6] = cmd[0] * c.obs_scale_lin_vel; obs[o0 + 7] = cmd[1] * c.obs_scale_lin_vel;
      obs[o0 + 8] = cmd[2] * c.obs_scale_ang_vel;
      if (c.family == 1) {
        obs[9 + 3 * A] = sinf(kTwoPi * phase);
        obs[10 + 3 * A] = cosf(kTwoPi * phase);
        const float cl = c.clip_obs;
        priv[0] = fminf(fmaxf(lvx * c.obs_scale_lin_vel, -cl), cl);
        priv[1] = fminf(fmaxf(lvy * c.obs_scale_lin_vel, -cl), cl);
        priv[2] = fminf(fmaxf(lvz * c.obs_scale_lin_vel, -cl), cl);
      }
    }
    _Pragma("unroll 1") for (int j = g; j < A; j += kLanesPerEnv) {
      const float qv = dof[2 * j + 1];
      obs[o0 + 9 + j] = (dof[2 * j] - c.default_dof_pos[j]) * c.obs_scale_dof_pos;
      obs[o0 + 9 + A + j] = qv * c.obs_scale_dof_vel;
      obs[o0 + 9 + 2 * A + j] = act[j];
      lact[j] = act[j];                                                 // tail copies :707-708
      ldv[j] = qv;
    }
    __syncwarp();
    const int chunk = (c.num_obs + kLanesPerEnv - 1) / kLanesPerEnv;
    const int i1 = min(c.num_obs, (g + 1) * chunk);
    const float cl = c.clip_obs;
    for (int i = g * chunk; i < i1; ++i) {
      float v = obs[i];
      if (c.family == 1) priv[3 + i] = fminf(fmaxf(v, -cl), cl);       // privileged obs carry no noise
      const float ns = c.noise_scale_vec[i];
      if (c.add_noise && ns != 0.0f) v += (2.0f * rng.uniform(14 + A + i) - 1.0f) * ns;
      obs[i] = fminf(fmaxf(v, -cl), cl);
    }
  }

  // ---- per-env scalars and short rows ----
  if (live) {
    if (g == 0) {
      b.episode_length_buf[e] = len;
      b.rew_buf[e] = rew;
      b.reset_buf[e] = reset ? 1 : 0;
      b.time_out_buf[e] = time_out ? 1 : 0;
      reinterpret_cast<float4*>(b.commands)[e] = make_float4(cmd[0], cmd[1], cmd[2], cmd[3]);
      if (c.family == 1) {
        b.phase[e] = phase;
        b.leg_phase[(size_t)e * 2] = leg_phase[0];
        b.leg_phase[(size_t)e * 2 + 1] = leg_phase[1];
      }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int f = g + 4 * q;
      if (f < F) {
        b.feet_air_time[(size_t)e * F + f] = air[q];
        b.last_contacts[(size_t)e * F + f] = lastc[q];
      }
    }
    // lane g stores the g-th of (rpy, base_lin_vel, base_ang_vel, projected_gravity)
    float* o3 = (g == 0 ? b.rpy : g == 1 ? b.base_lin_vel : g == 2 ? b.base_ang_vel : b.projected_gravity) + (size_t)e * 3;
    o3[0] = g == 0 ? roll : g == 1 ? lvx : g == 2 ? avx : gx;
    o3[1] = g == 0 ? pitch : g == 1 ? lvy : g == 2 ? avy : gy;
    o3[2] = g == 0 ? yaw : g == 1 ? lvz : g == 2 ? avz : gz;
  }
  if (__any_sync(0xffffffffu, push_due) && (tid & 31) == 0) atomicOr(&scratch->push_flag[acc], 1);
  if (bulk) {
    // ---- row outputs leave as 1-D bulk-TMA stores (shared -> global), issued by one thread ----
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      bulk_s2g(b.obs_buf[obs_slot] + (size_t)env0 * c.num_obs, sm + L.oObs, kEnvsPerBlock * c.num_obs * sizeof(float));
      if (c.num_priv)
        bulk_s2g(b.priv_obs_buf[obs_slot] + (size_t)env0 * c.num_priv, sm + L.oPriv, kEnvsPerBlock * c.num_priv * sizeof(float));
      bulk_s2g(b.last_actions + (size_t)env0 * A, sm + L.oLact, kEnvsPerBlock * A * sizeof(float));
      bulk_s2g(b.last_dof_vel + (size_t)env0 * A, sm + L.oLdv, kEnvsPerBlock * A * sizeof(float));
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    return;
  }
  __syncthreads();

  // ---- tail block / unaligned fallback: flat stores of the row outputs ----
  flat_out(b.obs_buf[obs_slot] + (size_t)env0 * c.num_obs, sm + L.oObs, nrows * c.num_obs, tid);
  if (c.num_priv)
    flat_out(b.priv_obs_buf[obs_slot] + (size_t)env0 * c.num_priv, sm + L.oPriv, nrows * c.num_priv, tid);
  flat_out(b.last_actions + (size_t)env0 * A, sm + L.oLact, nrows * A, tid);
  flat_out(b.last_dof_vel + (size_t)env0 * A, sm + L.oLdv, nrows * A, tid);
}

// ---------------------------------------------------------------- K2b: finalize
// Runs after K2 of the same step (stream order).  Reads this step's accumulators, applies the
// all-env push (quirk: any due env => every env's xy velocity is redrawn, :544-550), copies
// last_root_vel (:709), publishes extras only if >= 1 env reset (:742-743, 759-768), and clears the
// other parity's accumulators for the next step.
__global__ void __launch_bounds__(256) finalize_step_kernel(const __grid_constant__ urlg_env_desc c,
                                                            const __grid_constant__ urlg_env_buffers b,
                                                            StepScratch* __restrict__ scratch,
                                                            const int64_t* __restrict__ step_base, int64_t step_off,
                                                            int num_slots) {
  const int64_t step = eff_step(step_base, step_off);
  const int acc = (int)(step & 1);
  const bool push = scratch->push_flag[acc] != 0;
  const float cnt = scratch->ep_acc[acc][URLG_MAX_TERMS];
  const int A = c.num_dof;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < c.num_envs; e += gridDim.x * blockDim.x) {
    float* root = b.root_states + (size_t)e * 13;
    if (push) {
      SlotRng rng;
      rng.init(c.seed_lo, c.seed_hi, kStreamEnv, c.env_offset + e, step,
               b.rng_override ? b.rng_override + (size_t)e * num_slots : nullptr);
      const float m = c.max_push_vel;
      root[7] = (m - (-m)) * rng.uniform(12 + A) + (-m);
      root[8] = (m - (-m)) * rng.uniform(13 + A) + (-m);
    }
    float* lrv = b.last_root_vel + (size_t)e * 6;
    for (int k = 0; k < 6; ++k) lrv[k] = root[7 + k];
    if (cnt > 0.0f && c.send_timeouts) b.extras_time_outs[e] = b.time_out_buf[e];
  }
  if (blockIdx.x == 0 && threadIdx.x < URLG_MAX_TERMS + 1) {
    const int k = threadIdx.x;
    if (cnt > 0.0f && k < c.num_sums) b.extras_episode[k] = scratch->ep_acc[acc][k] / cnt / c.max_episode_length_s;
    if (k == 0 && b.reset_count) *b.reset_count = scratch->reset_cnt[acc];
    // clear the other parity for step+1 (its last reader was finalize of step-1)
    scratch->ep_acc[acc ^ 1][k] = 0.0f;
    if (k == 0) { scratch->push_flag[acc ^ 1] = 0; scratch->reset_cnt[acc ^ 1] = 0; }
  }
}

}  // namespace urlg

namespace urlg {

// ---------------------------------------------------------------- reset_idx(env_ids) outside step()
// base_task.py:82-86 -> legged_robot.py:723-768.  One thread per listed env, straight to global memory
// (construction-time path, not bandwidth critical).  Accumulator slot 2 is private to this path.
__global__ void __launch_bounds__(128) reset_idx_kernel(const __grid_constant__ urlg_env_desc c,
                                                        const __grid_constant__ urlg_env_buffers b,
                                                        StepScratch* __restrict__ scratch,
                                                        const int64_t* __restrict__ env_ids, int64_t n,
                                                        const int64_t* __restrict__ step_base, int64_t step_off,
                                                        int num_slots) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int e = env_ids ? (int)env_ids[i] : (int)i;
  const int A = c.num_dof, F = c.num_feet, N = c.num_envs;
  const int64_t step = eff_step(step_base, step_off);
  SlotRng rng;
  rng.init(c.seed_lo, c.seed_hi, kStreamEnv, c.env_offset + e, step,
           b.rng_override ? b.rng_override + (size_t)e * num_slots : nullptr);
  float* gdof = b.dof_state + (size_t)e * 2 * A;
  for (int j = 0; j < A; ++j) {
    gdof[2 * j] = c.default_dof_pos[j] * ((1.5f - 0.5f) * rng.uniform(3 + j) + 0.5f);
    gdof[2 * j + 1] = 0.0f;
  }
  float* groot = b.root_states + (size_t)e * 13;
  for (int k = 0; k < 13; ++k) groot[k] = c.base_init_state[k] + (k < 3 ? b.env_origins[(size_t)e * 3 + k] : 0.0f);
  for (int k = 0; k < 6; ++k) groot[7 + k] = (0.5f - (-0.5f)) * rng.uniform(3 + A + k) + (-0.5f);
  float cmd[4];
  for (int k = 0; k < 4; ++k) cmd[k] = b.commands[(size_t)e * 4 + k];
  resample_commands(c, rng, 9 + A, cmd);
  for (int k = 0; k < 4; ++k) b.commands[(size_t)e * 4 + k] = cmd[k];
  for (int j = 0; j < A; ++j) {
    b.actions[(size_t)e * A + j] = 0.0f;
    b.last_actions[(size_t)e * A + j] = 0.0f;
    b.last_dof_vel[(size_t)e * A + j] = 0.0f;
  }
  for (int f = 0; f < F; ++f) b.feet_air_time[(size_t)e * F + f] = 0.0f;
  b.episode_length_buf[e] = 0;
  b.reset_buf[e] = 1;
  for (int k = 0; k < c.num_sums; ++k) {
    float* s = b.episode_sums + (size_t)k * N + e;
    atomicAdd(&scratch->ep_acc[2][k], *s);
    *s = 0.0f;
  }
  atomicAdd(&scratch->ep_acc[2][URLG_MAX_TERMS], 1.0f);
}

__global__ void __launch_bounds__(256) finalize_reset_kernel(const __grid_constant__ urlg_env_desc c,
                                                             const __grid_constant__ urlg_env_buffers b,
                                                             StepScratch* __restrict__ scratch) {
  // single block: publish, then clear slot 2
  const float cnt = scratch->ep_acc[2][URLG_MAX_TERMS];
  if (cnt > 0.0f) {
    if (c.send_timeouts)
      for (int e = threadIdx.x; e < c.num_envs; e += blockDim.x) b.extras_time_outs[e] = b.time_out_buf[e];
    if (threadIdx.x < c.num_sums)
      b.extras_episode[threadIdx.x] = scratch->ep_acc[2][threadIdx.x] / cnt / c.max_episode_length_s;
  }
  __syncthreads();
  if (threadIdx.x < URLG_MAX_TERMS + 1) scratch->ep_acc[2][threadIdx.x] = 0.0f;
}

// ---------------------------------------------------------------- synthetic physics seam
__global__ void __launch_bounds__(256) synth_refresh_kernel(float* __restrict__ dst, const float* __restrict__ pool,
                                                            int64_t elems, int64_t num_frames, int64_t substeps,
                                                            int per_substep, const int64_t* __restrict__ calls_base,
                                                            int64_t calls_off) {
  const int64_t calls = eff_step(calls_base, calls_off);
  if (calls <= 0) return;
  const int64_t cidx = calls - 1;
  const int64_t frame = (cidx / substeps) % num_frames;
  const int64_t sub = cidx % substeps;
  const float* src = pool + (per_substep ? (frame * substeps + sub) : frame) * elems;
  const int64_t n4 = elems >> 2;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
  if (((elems & 3) == 0) && ((((uintptr_t)src | (uintptr_t)dst) & 15) == 0)) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int64_t i = tid; i < n4; i += nt) d4[i] = s4[i];
  } else {
    for (int64_t i = tid; i < elems; i += nt) dst[i] = src[i];
  }
}

__global__ void philox_uniform_kernel(float* out, int64_t n, int64_t num_slots, uint32_t seed_lo, uint32_t seed_hi,
                                      uint32_t stream_id, int64_t env_offset, int64_t step) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * num_slots) return;
  const int64_t e = i / num_slots;
  const int s = (int)(i - e * num_slots);
  SlotRng rng;
  rng.init(seed_lo, seed_hi, stream_id, env_offset + e, step, nullptr);
  out[i] = rng.uniform(s);
}

static int grid_for(int64_t work, int block, int max_waves = 8) {
  int64_t g = (work + block - 1) / block;
  const int64_t cap = (int64_t)kNumSMs * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace urlg

// ================================================================= C ABI
using namespace urlg;

extern "C" int urlg_version(void) { return URLG_ABI_VERSION; }
extern "C" const char* urlg_last_error(void) { return g_err; }
extern "C" int64_t urlg_launch_count(void) { return (int64_t)g_launches; }

extern "C" int urlg_env_create(const urlg_env_desc* d, urlg_env** out) {
  if (!d || !out) return fail(URLG_EINVAL, "urlg_env_create: null argument");
  if (d->num_envs <= 0 || d->num_dof <= 0 || d->num_dof > URLG_MAX_DOF)
    return fail(URLG_EINVAL, "urlg_env_create: num_envs=%d num_dof=%d out of range (max dof %d)", d->num_envs,
                d->num_dof, URLG_MAX_DOF);
  if (d->num_obs > URLG_MAX_OBS || d->num_priv > URLG_MAX_OBS || d->num_terms > URLG_MAX_TERMS ||
      d->num_sums > URLG_MAX_TERMS || d->num_feet > kMaxFeet || d->num_pen > URLG_MAX_IDX ||
      d->num_term_bodies > URLG_MAX_IDX || d->num_hip > URLG_MAX_IDX)
    return fail(URLG_EINVAL, "urlg_env_create: a table exceeds its ABI bound");
  if (d->control_type != 'P' && d->control_type != 'V' && d->control_type != 'T')
    return fail(URLG_EINVAL, "Unknown controller type: %c", (char)d->control_type);   // NameError at legged_robot.py:670
  const int expect_obs = d->family == 0 ? 12 + 3 * d->num_dof : 9 + 3 * d->num_dof + 2;
  if (d->num_obs != expect_obs || (d->family == 1 && d->num_priv != d->num_obs + 3) || (d->family == 0 && d->num_priv != 0))
    return fail(URLG_EINVAL, "urlg_env_create: num_obs=%d num_priv=%d inconsistent with family %d, num_dof %d",
                d->num_obs, d->num_priv, d->family, d->num_dof);
  if (d->resample_steps <= 0 || d->push_interval <= 0)
    return fail(URLG_EINVAL, "urlg_env_create: resample_steps / push_interval must be positive");
  urlg_env* e = new (std::nothrow) urlg_env();
  if (!e) return fail(URLG_ENOMEM, "urlg_env_create: host allocation failed");
  e->desc = *d;
  e->bound = 0;
  e->scratch = nullptr;
  URLG_CUDA(cudaMalloc(&e->scratch, sizeof(StepScratch)));
  URLG_CUDA(cudaMemset(e->scratch, 0, sizeof(StepScratch)));
  e->smem_bytes = make_layout(e->desc).total * (int)sizeof(float);
  if (e->smem_bytes > 48 * 1024)
    URLG_CUDA(cudaFuncSetAttribute(post_physics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_bytes));
  *out = e;
  return URLG_OK;
}

extern "C" int urlg_env_bind(urlg_env* env, const urlg_env_buffers* b) {
  if (!env || !b) return fail(URLG_EINVAL, "urlg_env_bind: null argument");
  const void* need[] = {b->root_states, b->dof_state, b->contact_forces, b->obs_buf[0], b->obs_buf[1], b->rew_buf,
                        b->reset_buf, b->time_out_buf, b->episode_length_buf, b->torques, b->actions, b->last_actions,
                        b->last_dof_vel, b->last_root_vel, b->commands, b->feet_air_time, b->last_contacts, b->rpy,
                        b->base_lin_vel, b->base_ang_vel, b->projected_gravity, b->episode_sums, b->env_origins,
                        b->extras_episode, b->extras_time_outs};
  for (size_t i = 0; i < sizeof(need) / sizeof(need[0]); ++i)
    if (!need[i]) return fail(URLG_EINVAL, "urlg_env_bind: required buffer #%zu is NULL", i);
  if (env->desc.family == 1 && (!b->rigid_body_states || !b->priv_obs_buf[0] || !b->priv_obs_buf[1] || !b->phase || !b->leg_phase))
    return fail(URLG_EINVAL, "urlg_env_bind: H1-family env needs rigid_body_states, priv_obs_buf, phase, leg_phase");
  if (((uintptr_t)b->commands & 15) || ((uintptr_t)b->dof_state & 7))
    return fail(URLG_EINVAL, "urlg_env_bind: commands must be 16-byte and dof_state 8-byte aligned");
  env->bufs = *b;
  {
    const void* al[] = {b->root_states, b->dof_state, b->contact_forces, b->actions, b->last_actions, b->last_dof_vel,
                        b->torques, b->obs_buf[0], b->obs_buf[1], b->priv_obs_buf[0], b->priv_obs_buf[1]};
    uintptr_t bits = 0;
    for (size_t i = 0; i < sizeof(al) / sizeof(al[0]); ++i) bits |= (uintptr_t)al[i];
    env->vec_ok = (bits & 15) == 0;
  }
  env->bound = 1;
  return URLG_OK;
}

extern "C" void urlg_env_destroy(urlg_env* env) {
  if (!env) return;
  if (env->scratch) cudaFree(env->scratch);
  delete env;
}

extern "C" int urlg_env_num_rng_slots(const urlg_env* env) {
  return env ? 14 + env->desc.num_dof + env->desc.num_obs : 0;
}

extern "C" int urlg_pd_torques(urlg_env* env, const float* raw_actions, void* stream) {
  if (!env || !env->bound) return fail(URLG_EINVAL, "urlg_pd_torques: env not bound");
  const int64_t total = (int64_t)env->desc.num_envs * env->desc.num_dof;
  if (total >= 0x7fffffffll) return fail(URLG_EINVAL, "urlg_pd_torques: num_envs * num_dof must be < 2^31");
  const urlg_env_buffers& eb = env->bufs;
  const uintptr_t bits = (uintptr_t)raw_actions | (uintptr_t)eb.actions | (uintptr_t)eb.dof_state | (uintptr_t)eb.torques |
                         (uintptr_t)eb.last_dof_vel;
  if ((env->desc.num_dof & 3) == 0 && (bits & 15) == 0) {
    const uint32_t total4 = (uint32_t)(total / 4);
    pd_torques_vec4_kernel<<<(total4 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(env->desc, env->bufs, raw_actions, total4);
  } else {
    pd_torques_kernel<<<grid_for(total, 256, 64), 256, 0, (cudaStream_t)stream>>>(env->desc, env->bufs, raw_actions);
  }
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_post_physics(urlg_env* env, const int64_t* step_base, int64_t step_offset, int obs_slot,
                                 void* stream) {
  if (!env || !env->bound) return fail(URLG_EINVAL, "urlg_post_physics: env not bound");
  if (obs_slot != 0 && obs_slot != 1) return fail(URLG_EINVAL, "urlg_post_physics: obs_slot must be 0 or 1");
  const int N = env->desc.num_envs;
  const int slots = urlg_env_num_rng_slots(env);
  cudaStream_t st = (cudaStream_t)stream;
  post_physics_kernel<<<(N + kEnvsPerBlock - 1) / kEnvsPerBlock, kK2Threads, env->smem_bytes, st>>>(
      env->desc, env->bufs, env->scratch, step_base, step_offset, obs_slot, slots, env->vec_ok);
  URLG_LAUNCH_CHECK();
  finalize_step_kernel<<<grid_for(N, 256), 256, 0, st>>>(env->desc, env->bufs, env->scratch, step_base, step_offset,
                                                        slots);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_env_reset_idx(urlg_env* env, const int64_t* env_ids, int64_t n, const int64_t* step_base,
                                  int64_t step_offset, void* stream) {
  if (!env || !env->bound) return fail(URLG_EINVAL, "urlg_env_reset_idx: env not bound");
  if (!env_ids) n = env->desc.num_envs;
  if (n <= 0) return URLG_OK;   // len(env_ids)==0 early-out, legged_robot.py:742-743
  cudaStream_t st = (cudaStream_t)stream;
  reset_idx_kernel<<<(int)((n + 127) / 128), 128, 0, st>>>(env->desc, env->bufs, env->scratch, env_ids, n, step_base,
                                                          step_offset, urlg_env_num_rng_slots(env));
  URLG_LAUNCH_CHECK();
  finalize_reset_kernel<<<1, 256, 0, st>>>(env->desc, env->bufs, env->scratch);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_copy_batch(void* const* dst, const void* const* src, const int64_t* bytes, int n, void* stream) {
  if (n < 0 || (n > 0 && (!dst || !src || !bytes))) return fail(URLG_EINVAL, "urlg_copy_batch: bad argument");
  for (int i = 0; i < n; ++i) {
    if (bytes[i] < 0 || (bytes[i] > 0 && (!dst[i] || !src[i]))) return fail(URLG_EINVAL, "urlg_copy_batch: bad entry %d", i);
    if (bytes[i]) URLG_CUDA(cudaMemcpyAsync(dst[i], src[i], (size_t)bytes[i], cudaMemcpyDefault, (cudaStream_t)stream));
  }
  return URLG_OK;
}

extern "C" int urlg_synth_refresh(float* dst, const float* pool, int64_t elems, int64_t num_frames, int64_t substeps,
                                  int per_substep, const int64_t* calls_base, int64_t calls_offset, void* stream) {
  if (!dst || !pool || elems <= 0 || num_frames <= 0 || substeps <= 0)
    return fail(URLG_EINVAL, "urlg_synth_refresh: bad argument");
  synth_refresh_kernel<<<grid_for((elems + 3) / 4, 256), 256, 0, (cudaStream_t)stream>>>(
      dst, pool, elems, num_frames, substeps, per_substep, calls_base, calls_offset);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_philox_uniform(float* out, int64_t n, int64_t num_slots, uint32_t seed_lo, uint32_t seed_hi,
                                   uint32_t stream_id, int64_t env_offset, int64_t step, void* stream) {
  if (!out || n <= 0 || num_slots <= 0) return fail(URLG_EINVAL, "urlg_philox_uniform: bad argument");
  const int64_t total = n * num_slots;
  philox_uniform_kernel<<<(int)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(out, n, num_slots, seed_lo,
                                                                                      seed_hi, stream_id, env_offset, step);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}
