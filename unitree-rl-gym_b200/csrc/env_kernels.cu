// Env half of the hot path: LeggedRobot.step minus the PhysX call, as three kernels.
//   K1  pd_torques_kernel      _compute_torques                 legged_robot.py:649-671 (+ :623-624), go2_handstand_env.py:326-350
//   K2  post_physics_kernel    post_physics_step + obs clips    legged_robot.py:673-709, 643-646; h1_env.py:55-123;
//                                                               go2_handstand_env.py:84-96, 140-219, 224-383
//   K2b finalize_step_kernel   _push_robots' all-env rewrite, last_root_vel, extras publication
//                                                               legged_robot.py:540-555, 709, 742-768
// K2 layout (v4): ONE THREAD PER ENV, one warp (32 consecutive envs) per block.  Every per-env quantity is computed once
// (v3 split an env over 4 lanes and recomputed the frame transforms / termination / scalar terms on all four: 394
// warp-instructions per env, instruction-bound at 0.24 of HBM).  Rows whose width is a multiple of 16 bytes (actions,
// last_actions, last_dof_vel, torques, dof_state) are read with vector loads straight into registers; the odd-width
// rows (root_states 13, contact_forces 3B floats) are staged by 1-D bulk TMA into shared memory, where the odd row
// stride makes thread-per-row access bank-conflict-free; observations are assembled in registers, written to a shared
// row image (noise is added there by a rolled loop over Philox blocks so the instruction footprint stays small) and
// leave as one bulk-TMA store.  The rare reset path (2-3 % of envs) runs warp-cooperatively: one lane per random draw.
// Floating point: every expression that DECIDES a mask (termination norms and angles, contact / stumble / command-norm
// thresholds, the gait-phase comparison, the heading wrap) is written with __fmul_rn / __fadd_rn / __fsub_rn, which are
// never contracted, so it evaluates exactly like the reference's separate eager ops and masks stay bit-exact; the
// value-only arithmetic (reward sums, observation scaling) may contract to FMA (1e-7 relative, tolerance 1e-5).
#include "common.cuh"
#include "philox.cuh"
#include <new>
#include <cmath>

namespace urlg {

thread_local char g_err[512] = "";
long long g_launches = 0;

constexpr float kPi = 3.14159274101257324f;      // float32(np.pi)
constexpr float kTwoPi = 6.28318548202636719f;   // float32(2*np.pi)
constexpr float kHalfPi = 1.57079637050628662f;  // float32(np.pi/2)
constexpr int kAccCopies = 16;                   // episode-sum accumulators are spread over 16 copies (atomic contention)

struct StepScratch {   // library-owned device scratch (zeroed at create)
  float ep_acc[3][kAccCopies][URLG_MAX_TERMS + 1];   // [slot][copy][row] sums over reset envs, [..][URLG_MAX_TERMS] = count
  int push_flag[3];
  int reset_cnt[3];
  int root_cnt[3];
};

// Reward table by term id (built from the desc's active list at create): kernels test `mask >> id & 1`.
struct TermTable {
  unsigned long long mask;
  float scale[URLG_T_COUNT];
  int row[URLG_T_COUNT];
  // sqrtf(s) > t  <=>  s > sq_thr(t): the largest float whose correctly-rounded root is still <= t (found on the host
  // at create; sqrt is monotonic), so mask-deciding norm comparisons need no square root and stay bit-exact
  float sq_thr_1p0, sq_thr_0p1;
  // Philox round keys of the env stream (key + r * Weyl constant, r = 0..9): constant-bank operands of the inlined
  // rounds instead of two integer adds per round; and which Philox blocks of the observation-noise slot range carry
  // a non-zero noise scale (bit = block - first block)
  uint32_t pk0[10], pk1[10];
  uint32_t noise_blocks;
};

struct PoolArgs {
  urlg_frame_pool pool;
  int64_t calls_off;
  int enabled;
};

}  // namespace urlg

struct urlg_env {
  urlg_env_desc desc;
  urlg_env_buffers bufs;
  urlg::TermTable tt;
  urlg::PoolArgs pool;
  urlg::StepScratch* scratch;
  int bound;
  int smem_bytes;
};

namespace urlg {

struct Seam {
  const float* root;
  const float* dof;
  const float* contact;
  const float* rigid;
};

// Physics tensors of the current (sub)step: the bound tensors, or a frame of the bound pool (include/urlg_b200.h).
__device__ __forceinline__ Seam resolve_seam(const urlg_env_desc& c, const urlg_env_buffers& b, const PoolArgs& pa) {
  Seam s{b.root_states, b.dof_state, b.contact_forces, b.rigid_body_states};
  if (pa.enabled) {
    const int64_t calls = *pa.pool.calls_base + pa.calls_off;
    if (calls > 0) {
      const uint64_t cidx = (uint64_t)(calls - 1);
      uint64_t frame, sub;
      const uint32_t D = (uint32_t)pa.pool.substeps, P = (uint32_t)pa.pool.num_frames;
      if (cidx <= 0xffffffffull) {
        const uint32_t q = (uint32_t)cidx / D;
        sub = (uint32_t)cidx - q * D;
        frame = q % P;
      } else {
        frame = (cidx / D) % P;
        sub = cidx % D;
      }
      const size_t N = (size_t)c.num_envs, fs = (size_t)pa.pool.frame_stride;
      const size_t dof_sub = N * (size_t)c.num_dof * 2;
      s.root = pa.pool.root + frame * (fs ? fs : N * 13);
      s.dof = pa.pool.dof + frame * (fs ? fs : D * dof_sub) + sub * dof_sub;
      s.contact = pa.pool.contact + frame * (fs ? fs : N * (size_t)c.num_bodies * 3);
      if (pa.pool.rigid_body) s.rigid = pa.pool.rigid_body + frame * (fs ? fs : N * (size_t)c.num_bodies * 13);
    }
  }
  return s;
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
// episode_length_buf % period (period > 0); the counters are small, so avoid the 64-bit division.
__device__ __forceinline__ bool divisible(int64_t len, int period) {
  if ((uint64_t)len <= 0x7fffffffull) return ((uint32_t)len % (uint32_t)period) == 0u;
  return (len % period) == 0;
}

#define MUL __fmul_rn
#define ADD __fadd_rn
#define SUB __fsub_rn
// x*x + y*y + z*z and x*x + y*y evaluated left to right without contraction (mask-deciding norms)
__device__ __forceinline__ float sq3(float x, float y, float z) { return ADD(ADD(MUL(x, x), MUL(y, y)), MUL(z, z)); }
__device__ __forceinline__ float sq2(float x, float y) { return ADD(MUL(x, x), MUL(y, y)); }

// torch.remainder(a, b) for b > 0 (ATen: fmod, then += b if the sign differs).  fmod is exact; for |a| < 2b it is a or
// a -+ b (exact by Sterbenz), for 0 <= a/b < 2^22 it is fma(-trunc(a/b), b, a) with a one-step correction (the exact
// remainder is representable, so the fma is exact); anything else takes libm's fmodf.
static __device__ __noinline__ float py_remainder(float a, float b) {
  float m;
  const float aa = fabsf(a);
  if (aa < b) {
    m = a;
  } else if (aa < MUL(2.0f, b)) {
    m = a < 0.0f ? ADD(a, b) : SUB(a, b);
  } else if (a > 0.0f && a < b * 4194304.0f) {
    const float q = truncf(a / b);
    m = fmaf(-q, b, a);
    if (m < 0.0f) m = ADD(m, b);
    else if (m >= b) m = SUB(m, b);
  } else {
    m = fmodf(a, b);
  }
  if (m != 0.0f && m < 0.0f) m = ADD(m, b);
  return m;
}

// One out-of-line copy of atan2f: K2 is latency-bound at 4096 envs and a cold instruction fetch costs more than a call.
static __device__ __noinline__ float atan2_ni(float y, float x) { return atan2f(y, x); }
__device__ __forceinline__ float norm3(float x, float y, float z) { return sqrtf(sq3(x, y, z)); }
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// [UPSTREAM isaacgym.torch_utils.quat_rotate_inverse], q = xyzw: v*(2w^2-1) - 2w(u x v) + 2u(u.v), uncontracted (the
// handstand termination compares projected gravity with a band; the velocities share the code)
__device__ __forceinline__ void rot_inv(float qx, float qy, float qz, float qw, float vx, float vy, float vz,
                                        float& ox, float& oy, float& oz) {
  const float s = SUB(MUL(2.0f, MUL(qw, qw)), 1.0f);
  const float cx = SUB(MUL(qy, vz), MUL(qz, vy)), cy = SUB(MUL(qz, vx), MUL(qx, vz)), cz = SUB(MUL(qx, vy), MUL(qy, vx));
  const float d = ADD(ADD(MUL(qx, vx), MUL(qy, vy)), MUL(qz, vz));
  ox = ADD(SUB(MUL(vx, s), MUL(MUL(cx, qw), 2.0f)), MUL(MUL(qx, d), 2.0f));
  oy = ADD(SUB(MUL(vy, s), MUL(MUL(cy, qw), 2.0f)), MUL(MUL(qy, d), 2.0f));
  oz = ADD(SUB(MUL(vz, s), MUL(MUL(cz, qw), 2.0f)), MUL(MUL(qz, d), 2.0f));
}

// Uniform of (global env, step, slot): Philox word slot%4 of block slot/4, or the injected table.
__device__ __forceinline__ float slot_uniform(const urlg_env_desc& c, const float* ovr, int64_t genv, int64_t step,
                                              int slot, int num_slots, int local_env) {
  if (ovr) return ovr[(size_t)local_env * num_slots + slot];
  const uint4 r = philox4x32_10(make_uint4((uint32_t)genv, (uint32_t)(step & 0xffffffff), (uint32_t)((uint64_t)step >> 32),
                                           (uint32_t)(slot >> 2)), c.seed_lo, c.seed_hi ^ kStreamEnv);
  const int l = slot & 3;
  return u01(l == 0 ? r.x : l == 1 ? r.y : l == 2 ? r.z : r.w);
}

// ---------------------------------------------------------------- K1: PD torques
// One element: PD law (legged_robot.py:649-671).  P control clamps the target to [pd_target_lo, pd_target_hi]
// (+-inf unless a task narrows it: go2_handstand_env.py:336-346); the torque clip is skipped when the task's override
// has none (:349-350).
__device__ __forceinline__ float pd_one(const urlg_env_desc& c, int j, float a, float q, float qd, float last_qd) {
  const float as = a * c.action_scale;
  float tq;
  if (c.control_type == 'P') {
    const float target = clampf(as + c.default_dof_pos[j], c.pd_target_lo[j], c.pd_target_hi[j]);
    tq = c.p_gains[j] * (target - q) - c.d_gains[j] * qd;
  } else if (c.control_type == 'V') {
    tq = c.p_gains[j] * (as - qd) - c.d_gains[j] * (qd - last_qd) / c.sim_dt;
  } else {
    tq = as;
  }
  if (!c.clip_torques) return tq;
  const float lim = c.torque_limits[j];
  return fminf(fmaxf(tq, -lim), lim);
}

// Vector form (num_dof % 4 == 0, 16-byte aligned tensors): one thread per 4 consecutive DOFs of an env -- one float4 of
// actions, two float4 of interleaved (pos, vel) pairs in, one float4 of torques out; one element per thread, no loop,
// so every load of the grid is in flight at once.
__global__ void __launch_bounds__(256) pd_torques_vec4_kernel(const __grid_constant__ urlg_env_desc c,
                                                              const __grid_constant__ urlg_env_buffers b,
                                                              const __grid_constant__ PoolArgs pa,
                                                              const float* __restrict__ raw_actions, uint32_t total4) {
  const uint32_t i = blockIdx.x * 256u + threadIdx.x;
  if (i >= total4) return;
  const Seam seam = resolve_seam(c, b, pa);
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
  const float4 s0 = reinterpret_cast<const float4*>(seam.dof)[2 * (size_t)i];
  const float4 s1 = reinterpret_cast<const float4*>(seam.dof)[2 * (size_t)i + 1];
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
                                                         const __grid_constant__ PoolArgs pa,
                                                         const float* __restrict__ raw_actions) {
  const Seam seam = resolve_seam(c, b, pa);
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
    const float2 pv = reinterpret_cast<const float2*>(seam.dof)[i];
    b.torques[i] = pd_one(c, j, a, pv.x, pv.y, c.control_type == 'V' ? b.last_dof_vel[i] : 0.0f);
  }
}

// ---------------------------------------------------------------- row helpers (thread per env)
// W floats of one row into registers.  Rows of a multiple of 16 / 8 bytes use 128- / 64-bit loads (the tensors are
// 16-byte aligned, checked at bind).
template <int W>
__device__ __forceinline__ void ld_row(float (&dst)[W], const float* __restrict__ src) {
  if constexpr (W % 4 == 0) {
#pragma unroll
    for (int i = 0; i < W / 4; ++i) {
      const float4 v = reinterpret_cast<const float4*>(src)[i];
      dst[4 * i] = v.x; dst[4 * i + 1] = v.y; dst[4 * i + 2] = v.z; dst[4 * i + 3] = v.w;
    }
  } else if constexpr (W % 2 == 0) {
#pragma unroll
    for (int i = 0; i < W / 2; ++i) {
      const float2 v = reinterpret_cast<const float2*>(src)[i];
      dst[2 * i] = v.x; dst[2 * i + 1] = v.y;
    }
  } else {
#pragma unroll
    for (int i = 0; i < W; ++i) dst[i] = src[i];
  }
}
template <int W>
__device__ __forceinline__ void st_row(float* __restrict__ dst, const float (&src)[W]) {
  if constexpr (W % 4 == 0) {
#pragma unroll
    for (int i = 0; i < W / 4; ++i)
      reinterpret_cast<float4*>(dst)[i] = make_float4(src[4 * i], src[4 * i + 1], src[4 * i + 2], src[4 * i + 3]);
  } else if constexpr (W % 2 == 0) {
#pragma unroll
    for (int i = 0; i < W / 2; ++i) reinterpret_cast<float2*>(dst)[i] = make_float2(src[2 * i], src[2 * i + 1]);
  } else {
#pragma unroll
    for (int i = 0; i < W; ++i) dst[i] = src[i];
  }
}

// ---------------------------------------------------------------- shared-memory layout (K2): 32 envs per block
// Row stride of the shared observation image: a stride that is a multiple of 8 floats would put the 32 rows of a warp
// on 4 banks (48-dim Go2 obs: 16-way conflicts), so such rows are padded by one float and leave through a
// warp-cooperative float4 copy; every other width is stored densely and leaves as one bulk-TMA store.
__host__ __device__ constexpr int obs_stride(int num_obs) { return num_obs % 8 == 0 ? num_obs + 1 : num_obs; }

struct Layout {
  int oRoot, oCon, oObs, oPriv, oTerm, oBar, total;   // float offsets
};
__host__ __device__ inline Layout make_layout(const urlg_env_desc& c) {
  Layout L;
  int o = 0;
  L.oRoot = o; o += 32 * 13;
  // the observation image is assembled after the last read of the contact rows and overlays them
  const int con = 32 * 3 * c.num_bodies, ob = 32 * obs_stride(c.num_obs);
  L.oCon = o; L.oObs = o; o += con > ob ? con : ob;
  o = (o + 3) & ~3;
  L.oPriv = o; o += c.family == URLG_FAMILY_H1 ? 32 * c.num_priv : 0;
  o = (o + 3) & ~3;
  L.oTerm = o; o += 32 * (c.num_sums > 0 ? c.num_sums : 1);
  o = (o + 3) & ~3;
  L.oBar = o; o += 4;
  L.total = o;
  return L;
}

template <int A, int FAM>
struct ObsDims {
  static constexpr int kObs = FAM == URLG_FAMILY_BASE ? 12 + 3 * A : FAM == URLG_FAMILY_H1 ? 11 + 3 * A : 10 + 3 * A;
  static constexpr int kPriv = FAM == URLG_FAMILY_H1 ? kObs + 3 : 0;
};

// command resampling (E6, legged_robot.py:519-538) from three uniforms
__device__ __forceinline__ void resample_commands(const urlg_env_desc& c, float ux, float uy, float uh, float (&cmd)[4]) {
  cmd[0] = ADD(MUL(SUB(c.cmd_x[1], c.cmd_x[0]), ux), c.cmd_x[0]);
  cmd[1] = ADD(MUL(SUB(c.cmd_y[1], c.cmd_y[0]), uy), c.cmd_y[0]);
  if (c.heading_command)
    cmd[3] = ADD(MUL(SUB(c.cmd_heading[1], c.cmd_heading[0]), uh), c.cmd_heading[0]);
  else
    cmd[2] = ADD(MUL(SUB(c.cmd_yaw[1], c.cmd_yaw[0]), uh), c.cmd_yaw[0]);
  const float keep = sqrtf(sq2(cmd[0], cmd[1])) > 0.2f ? 1.0f : 0.0f;
  cmd[0] *= keep;
  cmd[1] *= keep;
}

// ---------------------------------------------------------------- K2: fused post-physics step
// RNG slot map per (env, step) (shared with oracle/refharness/run_reference.py):
//   0..2 callback resample | 3..3+A-1 _reset_dofs | 3+A..8+A root velocities | 9+A..11+A reset resample |
//   12+A,13+A push | 14+A..14+A+O-1 observation noise | handstand: 14+A+O..+3 contact-flag flips
#ifndef URLG_K2_MIN_BLOCKS
#define URLG_K2_MIN_BLOCKS 16
#endif
// MASK: the active reward terms at compile time (0 = read tt.mask at run time).  The shipped task configs get their own
// instantiation: the 34 per-term mask tests disappear and the reward stage becomes one schedulable block.
template <int A, int FAM, unsigned long long MASK>
__global__ void __launch_bounds__(32, URLG_K2_MIN_BLOCKS) post_physics_kernel(const __grid_constant__ urlg_env_desc c,
                                                          const __grid_constant__ urlg_env_buffers b,
                                                          const __grid_constant__ TermTable tt,
                                                          const __grid_constant__ PoolArgs pa,
                                                          StepScratch* __restrict__ scratch,
                                                          const int64_t* __restrict__ step_base, int64_t step_off,
                                                          int obs_slot, int num_slots) {
  constexpr int O = ObsDims<A, FAM>::kObs, PV = ObsDims<A, FAM>::kPriv, OS = obs_stride(O);
  constexpr bool kPadObs = OS != O;
  constexpr unsigned kFull = 0xffffffffu;
  extern __shared__ __align__(128) float sm[];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // lets K2b's blocks start their Philox draws early
  const Layout L = make_layout(c);
  const int lane = threadIdx.x;
  const int env0 = blockIdx.x * 32;
  const int N = c.num_envs, B = c.num_bodies, F = c.num_feet;
  const int nrows = min(32, N - env0);
  const bool live = lane < nrows;
  const int e = live ? env0 + lane : env0;       // dead lanes of the tail block shadow env0 and store nothing
  const int64_t step = eff_step(step_base, step_off);
  const int acc = (int)(step & 1);
  const bool bulk = nrows == 32;
  const Seam seam = resolve_seam(c, b, pa);
  const uint64_t mask = MASK ? MASK : tt.mask;
#define HAS(ID) ((mask >> (ID)) & 1ull)

  // ---- stage the odd-width rows (root 13, contacts 3B floats per env) ----
  const uint32_t bar = smem_addr(sm + L.oBar);
  if (bulk) {
    if (lane == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      const uint32_t bytes = (uint32_t)(32 * (13 + 3 * B) * sizeof(float));
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
      bulk_g2s(sm + L.oRoot, seam.root + (size_t)env0 * 13, 32 * 13 * sizeof(float), bar);
      bulk_g2s(sm + L.oCon, seam.contact + (size_t)env0 * 3 * B, 32 * 3 * B * sizeof(float), bar);
    }
  } else {
    for (int i = lane; i < 32 * 13; i += 32) sm[L.oRoot + i] = i < nrows * 13 ? seam.root[(size_t)env0 * 13 + i] : 0.0f;
    for (int i = lane; i < 32 * 3 * B; i += 32)
      sm[L.oCon + i] = i < nrows * 3 * B ? seam.contact[(size_t)env0 * 3 * B + i] : 0.0f;
  }

  // ---- everything else straight into registers, issued before the wait ----
  int64_t len = b.episode_length_buf[e] + 1;                             // :681
  float cmd[4];
  {
    const float4 v = reinterpret_cast<const float4*>(b.commands)[e];
    cmd[0] = v.x; cmd[1] = v.y; cmd[2] = v.z; cmd[3] = v.w;
  }
  float act[A], lact[A], ldv[A], tq[A], dofs[2 * A];
  ld_row<A>(act, b.actions + (size_t)e * A);
  ld_row<A>(lact, b.last_actions + (size_t)e * A);
  ld_row<A>(ldv, b.last_dof_vel + (size_t)e * A);
  ld_row<A>(tq, b.torques + (size_t)e * A);
  ld_row<2 * A>(dofs, seam.dof + (size_t)e * 2 * A);
  float air[URLG_MAX_FEET];
  uint8_t lastc[URLG_MAX_FEET];
#pragma unroll
  for (int f = 0; f < URLG_MAX_FEET; ++f) {
    air[f] = 0.0f;
    lastc[f] = 0;
    if (f < F) {
      air[f] = b.feet_air_time[(size_t)e * F + f];
      lastc[f] = b.last_contacts[(size_t)e * F + f];
    }
  }
  // feet rows of rigid_body_states (h1_env.py:51-53; go2_handstand_env.py:73-76): only the columns the terms read
  float foot_z[2] = {0.0f, 0.0f}, foot_v[2][3] = {{0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}}, foot_xy[2][2] = {{0.0f, 0.0f}, {0.0f, 0.0f}};
  if constexpr (FAM == URLG_FAMILY_H1) {
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      if (f < F) {
        const float* rb = seam.rigid + ((size_t)e * B + c.feet_idx[f]) * 13;
        foot_z[f] = rb[2];
        foot_v[f][0] = rb[7]; foot_v[f][1] = rb[8]; foot_v[f][2] = rb[9];
      }
    }
  }
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    if (HAS(URLG_T_FRONT_FEET_TOGETHER)) {
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        const float* rb = seam.rigid + ((size_t)e * B + c.feet_idx[f]) * 13;
        foot_xy[f][0] = rb[0]; foot_xy[f][1] = rb[1];
      }
    }
  }
  // episode_sums rows of this env (read-modify-write at the end of the reward stage): in flight before the wait
  constexpr int kSumsReg = 16;
  float es[kSumsReg];
  float* const esp = b.episode_sums + e;         // row k at esp[k * N]; num_sums * N < 2^31 (checked at create)
#pragma unroll
  for (int k = 0; k < kSumsReg; ++k) {
    es[k] = 0.0f;
    if (k < c.num_sums) es[k] = esp[(uint32_t)k * (uint32_t)N];
  }
  const float* ovr = b.rng_override;
  const int64_t genv = c.env_offset + e;

  if (bulk) {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok) : "r"(bar) : "memory");
    }
  } else {
    __syncwarp();
  }

  float root[13];
#pragma unroll
  for (int k = 0; k < 13; ++k) root[k] = sm[L.oRoot + lane * 13 + k];
  const float* con = sm + L.oCon + lane * 3 * B;
  float* term = sm + L.oTerm + lane;           // term[row * 32]
  const float flen = (float)len;

  // ---- E4 frame transforms :685-690, isaacgym_utils.py:11-30 ----
  const float qx = root[3], qy = root[4], qz = root[5], qw = root[6];
  const float ww = MUL(qw, qw), xx = MUL(qx, qx), yy = MUL(qy, qy), zz = MUL(qz, qz);
  const float roll = atan2_ni(MUL(2.0f, ADD(MUL(qw, qx), MUL(qy, qz))), ADD(SUB(SUB(ww, xx), yy), zz));
  const float sinp = MUL(2.0f, SUB(MUL(qw, qy), MUL(qz, qx)));
  const float pitch = fabsf(sinp) >= 1.0f ? copysignf(kHalfPi, sinp) : asinf(sinp);
  const float yaw = atan2_ni(MUL(2.0f, ADD(MUL(qw, qz), MUL(qx, qy))), SUB(SUB(ADD(ww, xx), yy), zz));
  float lvx, lvy, lvz, avx, avy, avz, gx, gy, gz;
  rot_inv(qx, qy, qz, qw, root[7], root[8], root[9], lvx, lvy, lvz);
  rot_inv(qx, qy, qz, qw, root[10], root[11], root[12], avx, avy, avz);
  rot_inv(qx, qy, qz, qw, 0.0f, 0.0f, -1.0f, gx, gy, gz);
  const float root_z = root[2];

  // ---- E5 callback: gait phase (h1_env.py:58-63), resample (:511-512), heading (:513-516) ----
  float phase = 0.0f, leg_phase[2] = {0.0f, 0.0f};
  if constexpr (FAM == URLG_FAMILY_H1) {
    phase = py_remainder(MUL(flen, c.dt), 0.8f) / 0.8f;
    leg_phase[0] = phase;
    leg_phase[1] = py_remainder(ADD(phase, 0.5f), 1.0f);
  }
  if (c.callback_mode == URLG_CALLBACK_BASE) {
    if (divisible(len, c.resample_steps)) {
      const float ux = slot_uniform(c, ovr, genv, step, 0, num_slots, e);
      const float uy = slot_uniform(c, ovr, genv, step, 1, num_slots, e);
      const float uh = slot_uniform(c, ovr, genv, step, 2, num_slots, e);
      resample_commands(c, ux, uy, uh, cmd);
    }
    if (c.heading_command) {
      // forward = quat_apply(q, (1,0,0)): t = 2 (u x v) = (0, 2 qz, -2 qy)
      const float ty = MUL(qz, 2.0f), tz = MUL(-qy, 2.0f);
      const float fx = ADD(ADD(1.0f, MUL(qw, 0.0f)), SUB(MUL(qy, tz), MUL(qz, ty)));
      const float fy = ADD(ADD(0.0f, MUL(qw, ty)), SUB(MUL(qz, 0.0f), MUL(qx, tz)));
      const float heading = atan2_ni(fy, fx);
      float a = py_remainder(SUB(cmd[3], heading), kTwoPi);
      a = SUB(a, MUL(kTwoPi, a > kPi ? 1.0f : 0.0f));
      cmd[2] = fminf(fmaxf(0.5f * a, -1.0f), 1.0f);
    }
  }

  // ---- E7 termination :715-721 / go2_handstand_env.py:178-219 ----
  bool reset = false;
  _Pragma("unroll 1") for (int k = 0; k < c.num_term_bodies; ++k) {
    const float* f3 = con + 3 * c.term_idx[k];
    reset |= sq3(f3[0], f3[1], f3[2]) > tt.sq_thr_1p0;          // ||F|| > 1.0
  }
  if (c.termination_mode == URLG_TERMINATION_RPY)
    reset |= (fabsf(pitch) > 1.0f) || (fabsf(roll) > 0.8f);
  else
    reset |= (gz > -0.2f) && (gz < 0.2f);
  const bool time_out = flen > c.max_episode_length;
  reset |= time_out;

  // ---- feet contact state shared by several terms ----
  float foot_fz[URLG_MAX_FEET], foot_fn[URLG_MAX_FEET], foot_fxy[URLG_MAX_FEET];
#pragma unroll
  for (int f = 0; f < URLG_MAX_FEET; ++f) {
    foot_fz[f] = 0.0f; foot_fn[f] = 0.0f; foot_fxy[f] = 0.0f;
    if (f < F) {
      const float* f3 = con + 3 * c.feet_idx[f];
      const float x = f3[0], y = f3[1], z = f3[2];
      foot_fz[f] = z;
      foot_fn[f] = norm3(x, y, z);
      foot_fxy[f] = sqrtf(sq2(x, y));
    }
  }

  // ---- E8/E9 rewards in alphabetical order of the reward names (:775-787) ----
  // tensor / python-scalar: ATen's CUDA kernel multiplies by the reciprocal (BinaryDivTrueKernel.cu, CPU-scalar divisor),
  // i.e. what the reference computes on its own device; the CPU fixtures (true division) agree to 1.2e-7 relative
  const float inv_dt = 1.0f / c.dt, inv_sigma = 1.0f / c.tracking_sigma;
  const float cmd_xy = sqrtf(sq2(cmd[0], cmd[1]));
  float rew = 0.0f;
#define TERM_END(ID) r *= tt.scale[ID]; rew += r; term[tt.row[ID] * 32] = r; }
#define TERM_BEGIN(ID) if (HAS(ID)) { float r = 0.0f;
  TERM_BEGIN(URLG_T_ACTION_RATE)
#pragma unroll
    for (int j = 0; j < A; ++j) { const float d = lact[j] - act[j]; r += d * d; }
  TERM_END(URLG_T_ACTION_RATE)
  TERM_BEGIN(URLG_T_ALIVE) r = 1.0f; TERM_END(URLG_T_ALIVE)
  TERM_BEGIN(URLG_T_ANG_VEL_XY) r = avx * avx + avy * avy; TERM_END(URLG_T_ANG_VEL_XY)
  TERM_BEGIN(URLG_T_BASE_HEIGHT) { const float d = root_z - c.base_height_target; r = d * d; } TERM_END(URLG_T_BASE_HEIGHT)
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    TERM_BEGIN(URLG_T_BASE_HEIGHT_HS)           // go2_handstand_env.py:237-246
      const float err = fmaxf(c.base_height_target - root_z, 0.0f);
      r = expf(-8.0f * err * err);
    TERM_END(URLG_T_BASE_HEIGHT_HS)
  }
  TERM_BEGIN(URLG_T_COLLISION)
    _Pragma("unroll 1") for (int k = 0; k < c.num_pen; ++k) {
      const float* f3 = con + 3 * c.pen_idx[k];
      r += sq3(f3[0], f3[1], f3[2]) > tt.sq_thr_0p1 ? 1.0f : 0.0f;   // ||F|| > 0.1
    }
  TERM_END(URLG_T_COLLISION)
  if constexpr (FAM == URLG_FAMILY_H1) {
    TERM_BEGIN(URLG_T_CONTACT)                  // h1_env.py:98-104
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        if (f < F) {
          const bool stance = leg_phase[f] < 0.55f;
          const bool contact = foot_fz[f] > 1.0f;
          r += (contact != stance) ? 0.0f : 1.0f;
        }
      }
    TERM_END(URLG_T_CONTACT)
    TERM_BEGIN(URLG_T_CONTACT_NO_VEL)           // h1_env.py:115-120
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        const float m = foot_fn[f] > 1.0f ? 1.0f : 0.0f;
#pragma unroll
        for (int k = 0; k < 3; ++k) { const float v = foot_v[f][k] * m; r += v * v; }
      }
    TERM_END(URLG_T_CONTACT_NO_VEL)
  }
  TERM_BEGIN(URLG_T_DOF_ACC)
#pragma unroll
    for (int j = 0; j < A; ++j) { const float d = (ldv[j] - dofs[2 * j + 1]) * inv_dt; r += d * d; }
  TERM_END(URLG_T_DOF_ACC)
  TERM_BEGIN(URLG_T_DOF_POS_LIMITS)
#pragma unroll
    for (int j = 0; j < A; ++j) {
      const float q = dofs[2 * j];
      r += -fminf(q - c.dof_pos_lo[j], 0.0f) + fmaxf(q - c.dof_pos_hi[j], 0.0f);
    }
  TERM_END(URLG_T_DOF_POS_LIMITS)
  TERM_BEGIN(URLG_T_DOF_VEL)
#pragma unroll
    for (int j = 0; j < A; ++j) { const float v = dofs[2 * j + 1]; r += v * v; }
  TERM_END(URLG_T_DOF_VEL)
  TERM_BEGIN(URLG_T_DOF_VEL_LIMITS)
#pragma unroll
    for (int j = 0; j < A; ++j)
      r += fminf(fmaxf(fabsf(dofs[2 * j + 1]) - c.dof_vel_limits[j] * c.soft_dof_vel_limit, 0.0f), 1.0f);
  TERM_END(URLG_T_DOF_VEL_LIMITS)
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    TERM_BEGIN(URLG_T_ENERGY)                   // go2_handstand_env.py:319-324
#pragma unroll
      for (int j = 0; j < A; ++j) r += fabsf(tq[j]) * fabsf(dofs[2 * j + 1]);
    TERM_END(URLG_T_ENERGY)
  }
  TERM_BEGIN(URLG_T_FEET_AIR_TIME)              // stateful, :912-923
#pragma unroll
    for (int f = 0; f < URLG_MAX_FEET; ++f) {
      if (f < F) {
        const bool contact = foot_fz[f] > 1.0f;
        const bool filt = contact || lastc[f];
        lastc[f] = contact;
        const float first = (air[f] > 0.0f && filt) ? 1.0f : 0.0f;
        air[f] += c.dt;
        r += (air[f] - 0.5f) * first;
        air[f] *= filt ? 0.0f : 1.0f;
      }
    }
    r *= cmd_xy > 0.1f ? 1.0f : 0.0f;
  TERM_END(URLG_T_FEET_AIR_TIME)
  TERM_BEGIN(URLG_T_FEET_CONTACT_FORCES)
#pragma unroll
    for (int f = 0; f < URLG_MAX_FEET; ++f)
      if (f < F) r += fmaxf(foot_fn[f] - c.max_contact_force, 0.0f);
  TERM_END(URLG_T_FEET_CONTACT_FORCES)
  if constexpr (FAM == URLG_FAMILY_H1) {
    TERM_BEGIN(URLG_T_FEET_SWING_HEIGHT)        // h1_env.py:106-109
#pragma unroll
      for (int f = 0; f < 2; ++f) {
        if (f < F) {
          const float m = foot_fn[f] > 1.0f ? 0.0f : 1.0f;
          const float d = foot_z[f] - 0.08f;
          r += d * d * m;
        }
      }
    TERM_END(URLG_T_FEET_SWING_HEIGHT)
  }
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    TERM_BEGIN(URLG_T_FRONT_FEET_CONTACT)       // :248-256, feet 0,1 = front
      r = ((foot_fz[0] > 5.0f ? 1.0f : 0.0f) + (foot_fz[1] > 5.0f ? 1.0f : 0.0f)) / 2.0f;
    TERM_END(URLG_T_FRONT_FEET_CONTACT)
    TERM_BEGIN(URLG_T_FRONT_FEET_TOGETHER)      // :360-373
      const float dx = foot_xy[0][0] - foot_xy[1][0], dy = foot_xy[0][1] - foot_xy[1][1];
      const float err = fmaxf(sqrtf(dx * dx + dy * dy) - 0.12f, 0.0f);
      r = expf(-20.0f * err * err);
    TERM_END(URLG_T_FRONT_FEET_TOGETHER)
    TERM_BEGIN(URLG_T_FRONT_HIP_NEUTRAL)        // :352-358, hip_idx = (FL_hip, FR_hip)
      float s = 0.0f;
      _Pragma("unroll 1") for (int k = 0; k < c.num_hip; ++k) {
        float q = 0.0f;
#pragma unroll
        for (int j = 0; j < A; ++j) q = c.hip_idx[k] == j ? dofs[2 * j] : q;
        s += q * q;
      }
      r = expf(-4.0f * s);
    TERM_END(URLG_T_FRONT_HIP_NEUTRAL)
    TERM_BEGIN(URLG_T_HIND_FEET_NO_CONTACT)     // :258-265, feet 2,3 = hind
      r = 1.0f - ((foot_fz[2] > 5.0f ? 1.0f : 0.0f) + (foot_fz[3] > 5.0f ? 1.0f : 0.0f)) / 2.0f;
    TERM_END(URLG_T_HIND_FEET_NO_CONTACT)
  }
  if constexpr (FAM == URLG_FAMILY_H1) {
    TERM_BEGIN(URLG_T_HIP_POS)                  // h1_env.py:122-123
      _Pragma("unroll 1") for (int k = 0; k < c.num_hip; ++k) {
        float q = 0.0f;
#pragma unroll
        for (int j = 0; j < A; ++j) q = c.hip_idx[k] == j ? dofs[2 * j] : q;
        r += q * q;
      }
    TERM_END(URLG_T_HIP_POS)
  }
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    TERM_BEGIN(URLG_T_LIN_VEL_XY) r = sqrtf(lvx * lvx + lvy * lvy); TERM_END(URLG_T_LIN_VEL_XY)   // :382-383
  }
  TERM_BEGIN(URLG_T_LIN_VEL_Z) r = lvz * lvz; TERM_END(URLG_T_LIN_VEL_Z)
  TERM_BEGIN(URLG_T_ORIENTATION) r = gx * gx + gy * gy; TERM_END(URLG_T_ORIENTATION)
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    TERM_BEGIN(URLG_T_ORIENTATION_HS)           // :224-235
      const float err = 1.0f - clampf(gz, -1.0f, 1.0f);
      r = expf(-3.0f * err * err);
    TERM_END(URLG_T_ORIENTATION_HS)
    TERM_BEGIN(URLG_T_POSE)                     // :277-307
      float sf = 0.0f, sh = 0.0f;
#pragma unroll
      for (int j = 0; j < A; ++j) {
        const float d = dofs[2 * j] - c.pose_targets[j];
        if (j < c.num_front_dof) sf += d * d; else sh += d * d;
      }
      r = expf(-4.0f * (sf + 0.3f * sh));
    TERM_END(URLG_T_POSE)
    TERM_BEGIN(URLG_T_STABILITY)                // :309-317
      r = expf(-2.0f * (sqrtf(lvx * lvx + lvy * lvy) + 0.5f * sqrtf(avx * avx + avy * avy)));
    TERM_END(URLG_T_STABILITY)
  }
  TERM_BEGIN(URLG_T_STAND_STILL)
#pragma unroll
    for (int j = 0; j < A; ++j) r += fabsf(dofs[2 * j] - c.default_dof_pos[j]);
    r *= cmd_xy < 0.1f ? 1.0f : 0.0f;
  TERM_END(URLG_T_STAND_STILL)
  if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
    TERM_BEGIN(URLG_T_STAY_STILL)               // :375-380
      const float a = sqrtf(lvx * lvx + lvy * lvy) * (1.0f / 0.05f), w = fabsf(avz) * (1.0f / 0.2f);
      r = expf(-(a * a) - w * w);
    TERM_END(URLG_T_STAY_STILL)
  }
  TERM_BEGIN(URLG_T_STUMBLE)
    bool any = false;
#pragma unroll
    for (int f = 0; f < URLG_MAX_FEET; ++f)
      if (f < F) any |= foot_fxy[f] > MUL(5.0f, fabsf(foot_fz[f]));
    r = any ? 1.0f : 0.0f;
  TERM_END(URLG_T_STUMBLE)
  TERM_BEGIN(URLG_T_TORQUE_LIMITS)
#pragma unroll
    for (int j = 0; j < A; ++j) r += fmaxf(fabsf(tq[j]) - c.torque_limits[j] * c.soft_torque_limit, 0.0f);
  TERM_END(URLG_T_TORQUE_LIMITS)
  TERM_BEGIN(URLG_T_TORQUES)
#pragma unroll
    for (int j = 0; j < A; ++j) r += tq[j] * tq[j];
  TERM_END(URLG_T_TORQUES)
  TERM_BEGIN(URLG_T_TRACKING_ANG_VEL) { const float d = cmd[2] - avz; r = expf(-(d * d) * inv_sigma); } TERM_END(URLG_T_TRACKING_ANG_VEL)
  TERM_BEGIN(URLG_T_TRACKING_LIN_VEL)
    const float dx = cmd[0] - lvx, dy = cmd[1] - lvy;
    r = expf(-(dx * dx + dy * dy) * inv_sigma);
  TERM_END(URLG_T_TRACKING_LIN_VEL)
#undef TERM_BEGIN
#undef TERM_END
  if (c.only_positive) rew = fmaxf(rew, 0.0f);
  if (c.has_termination) {
    const float r = ((reset && !time_out) ? 1.0f : 0.0f) * c.termination_scale;
    rew += r;
    term[c.termination_sum_row * 32] = r;
  }

  // ---- episode_sums += this step's terms (:779-780, 786); reset envs publish their sums and clear (:759-763) ----
  {
    const bool pub = reset && live;
    float* accp = scratch->ep_acc[acc][blockIdx.x & (kAccCopies - 1)];
#pragma unroll
    for (int k = 0; k < kSumsReg; ++k) {
      if (k < c.num_sums) {
        float v = es[k] + term[k * 32];
        if (pub) { atomicAdd(accp + k, v); v = 0.0f; }
        if (live) esp[(uint32_t)k * (uint32_t)N] = v;
      }
    }
    _Pragma("unroll 1") for (int k = kSumsReg; k < c.num_sums; ++k) {
      float v = b.episode_sums[(size_t)k * N + e] + term[k * 32];
      if (pub) { atomicAdd(accp + k, v); v = 0.0f; }
      if (live) b.episode_sums[(size_t)k * N + e] = v;
    }
    if (pub) atomicAdd(accp + URLG_MAX_TERMS, 1.0f);
  }

  // ---- E10 reset_idx :723-768, warp-cooperative: for every reset env of the warp, lane l draws slot 3 + l
  //      (l < A: _reset_dofs :566-567; A <= l < A+6: root velocities :590; A+6 <= l < A+9: _resample_commands :749)
  //      and writes its element of the physics tensors; the owner lane picks the new values up by shuffle ----
  {
    unsigned rmask = __ballot_sync(kFull, reset && live);
    while (rmask) {
      const int src = __ffs(rmask) - 1;
      rmask &= rmask - 1;
      const int er = env0 + src;
      float u = 0.0f;
      if (lane < A + 9) u = slot_uniform(c, ovr, c.env_offset + er, step, 3 + lane, num_slots, er);
      float p = 0.0f;
      if (lane < A) {
        p = c.default_dof_pos[lane] * ((1.5f - 0.5f) * u + 0.5f);
        reinterpret_cast<float2*>(b.dof_state)[(size_t)er * A + lane] = make_float2(p, 0.0f);
        b.actions[(size_t)er * A + lane] = 0.0f;                             // :752
      }
      const float uv = __shfl_sync(kFull, u, (lane >= 7 && lane < 13) ? A + (lane - 7) : 0);
      if (lane < 13) {                                                      // _reset_root_states :587-590
        float v = c.base_init_state[lane];
        if (lane < 3) v += b.env_origins[(size_t)er * 3 + lane];
        if (lane >= 7) v = (0.5f - (-0.5f)) * uv + (-0.5f);
        b.root_states[(size_t)er * 13 + lane] = v;
      }
      const float ux = __shfl_sync(kFull, u, A + 6), uy = __shfl_sync(kFull, u, A + 7), uh = __shfl_sync(kFull, u, A + 8);
      const bool mine = lane == src;
#pragma unroll
      for (int j = 0; j < A; ++j) {
        const float pj = __shfl_sync(kFull, p, j);
        if (mine) { dofs[2 * j] = pj; dofs[2 * j + 1] = 0.0f; act[j] = 0.0f; }
      }
      if (mine) {
        resample_commands(c, ux, uy, uh, cmd);
#pragma unroll
        for (int f = 0; f < URLG_MAX_FEET; ++f) air[f] = 0.0f;              // :755
        len = 0;                                                             // :756
      }
    }
  }
  const bool push_due = live && c.push_robots && divisible(len, c.push_interval);   // :545, evaluated after the reset

  // ---- compacted id lists for the indexed setters of the physics seam (:552-555, 569-572, 591-594) ----
  {
    const unsigned rm = __ballot_sync(kFull, reset && live);
    const unsigned pm = __ballot_sync(kFull, (reset && live) || push_due);
    const unsigned dm = __ballot_sync(kFull, push_due);
    int rbase = 0, pbase = 0;
    if (lane == 0) {
      if (rm) rbase = atomicAdd(&scratch->reset_cnt[acc], __popc(rm));
      if (pm) pbase = atomicAdd(&scratch->root_cnt[acc], __popc(pm));
      if (dm) atomicOr(&scratch->push_flag[acc], 1);
    }
    rbase = __shfl_sync(kFull, rbase, 0);
    pbase = __shfl_sync(kFull, pbase, 0);
    const unsigned below = (1u << lane) - 1u;
    if (reset && live && b.reset_ids) b.reset_ids[rbase + __popc(rm & below)] = e;
    if (((reset && live) || push_due) && b.root_ids) b.root_ids[pbase + __popc(pm & below)] = e;
  }

  // ---- E12 observations :789-811 / h1_env.py:68-95 / go2_handstand_env.py:140-176, assembled in registers ----
  float* obs_s = sm + L.oObs + lane * OS;
  {
    float obs[O];
    const float cl = c.clip_obs;
    constexpr int o0 = FAM == URLG_FAMILY_BASE ? 3 : 0;          // column of ang_vel
    constexpr int oq = FAM == URLG_FAMILY_HANDSTAND ? o0 + 6 : o0 + 9;   // column of dof_pos (handstand has no commands)
    if constexpr (FAM == URLG_FAMILY_BASE) {
      obs[0] = lvx * c.obs_scale_lin_vel; obs[1] = lvy * c.obs_scale_lin_vel; obs[2] = lvz * c.obs_scale_lin_vel;
    }
    obs[o0 + 0] = avx * c.obs_scale_ang_vel; obs[o0 + 1] = avy * c.obs_scale_ang_vel; obs[o0 + 2] = avz * c.obs_scale_ang_vel;
    obs[o0 + 3] = gx; obs[o0 + 4] = gy; obs[o0 + 5] = gz;
    if constexpr (FAM != URLG_FAMILY_HANDSTAND) {
      obs[o0 + 6] = cmd[0] * c.obs_scale_lin_vel; obs[o0 + 7] = cmd[1] * c.obs_scale_lin_vel;
      obs[o0 + 8] = cmd[2] * c.obs_scale_ang_vel;
    }
#pragma unroll
    for (int j = 0; j < A; ++j) {
      obs[oq + j] = (dofs[2 * j] - c.default_dof_pos[j]) * c.obs_scale_dof_pos;
      obs[oq + A + j] = dofs[2 * j + 1] * c.obs_scale_dof_vel;
      obs[oq + 2 * A + j] = act[j];
    }
    if constexpr (FAM == URLG_FAMILY_H1) {
      obs[9 + 3 * A] = sinf(kTwoPi * phase);
      obs[10 + 3 * A] = cosf(kTwoPi * phase);
      float* priv_s = sm + L.oPriv + lane * PV;                  // privileged obs carry no noise
      priv_s[0] = clampf(lvx * c.obs_scale_lin_vel, -cl, cl);
      priv_s[1] = clampf(lvy * c.obs_scale_lin_vel, -cl, cl);
      priv_s[2] = clampf(lvz * c.obs_scale_lin_vel, -cl, cl);
#pragma unroll
      for (int i = 0; i < O; ++i) priv_s[3 + i] = clampf(obs[i], -cl, cl);
    }
    if constexpr (FAM == URLG_FAMILY_HANDSTAND) {
      // contact flags of the front (feet 0,1) and hind (feet 2,3) feet; with noise on, re-binarised and flipped with
      // probability 0.01 (:160-169) -- their noise scale is 0, so the flip can be applied here
      float uf[4] = {1.0f, 1.0f, 1.0f, 1.0f};
      if (c.add_noise) {
        constexpr int sf = 14 + A + O;
        if (ovr) {
#pragma unroll
          for (int k = 0; k < 4; ++k) uf[k] = ovr[(size_t)e * num_slots + sf + k];
        } else if constexpr ((sf & 3) == 0) {          // the four flip slots are one Philox block
          const uint4 r = philox4x32_10(make_uint4((uint32_t)genv, (uint32_t)(step & 0xffffffff),
                                                   (uint32_t)((uint64_t)step >> 32), (uint32_t)(sf >> 2)), c.seed_lo,
                                        c.seed_hi ^ kStreamEnv);
          uf[0] = u01(r.x); uf[1] = u01(r.y); uf[2] = u01(r.z); uf[3] = u01(r.w);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) uf[k] = slot_uniform(c, ovr, genv, step, sf + k, num_slots, e);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool flag = (foot_fz[k] > 5.0f) != (uf[k] < 0.01f);
        obs[oq + 3 * A + k] = flag ? 1.0f : 0.0f;
      }
    }
    // observation noise :811: obs += (2u - 1) * noise_scale_vec, one Philox block (4 uniforms) per 4 consecutive slots;
    // unrolled so that columns and noise scales are compile-time indexed; blocks whose scales are all zero are skipped
    if (c.add_noise) {
      constexpr int s0 = 14 + A, b0 = s0 >> 2, b1 = (s0 + O - 1) >> 2;
#pragma unroll
      for (int blk = b0; blk <= b1; ++blk) {
        const int i0 = 4 * blk - s0;               // obs column of this block's word 0 (may be < 0 / run past O)
        if ((tt.noise_blocks >> (blk - b0)) & 1u) {
          float u[4];
          if (ovr) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int i = i0 + k;
              u[k] = (i >= 0 && i < O) ? ovr[(size_t)e * num_slots + s0 + (i >= 0 && i < O ? i : 0)] : 0.0f;
            }
          } else {
            const uint4 r = philox4x32_10_keys(make_uint4((uint32_t)genv, (uint32_t)(step & 0xffffffff),
                                                          (uint32_t)((uint64_t)step >> 32), (uint32_t)blk), tt.pk0, tt.pk1);
            u[0] = u01(r.x); u[1] = u01(r.y); u[2] = u01(r.z); u[3] = u01(r.w);
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int i = i0 + k;      // a zero scale adds (2u - 1) * 0 = +-0: the value is unchanged
            if (i >= 0 && i < O) obs[i] = obs[i] + (2.0f * u[k] - 1.0f) * c.noise_scale_vec[i >= 0 && i < O ? i : 0];
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < O; ++i) obs[i] = clampf(obs[i], -cl, cl);        // the clips of step() :643-646
    __syncwarp();                                                        // every lane is done with the contact rows
    if constexpr (kPadObs) {
#pragma unroll
      for (int i = 0; i < O; ++i) obs_s[i] = obs[i];
    } else {
      st_row<O>(obs_s, obs);
    }
  }

  // ---- per-env stores ----
  if (live) {
    b.episode_length_buf[e] = len;
    b.rew_buf[e] = rew;
    b.reset_buf[e] = reset ? 1 : 0;
    b.time_out_buf[e] = time_out ? 1 : 0;
    reinterpret_cast<float4*>(b.commands)[e] = make_float4(cmd[0], cmd[1], cmd[2], cmd[3]);
    if constexpr (FAM == URLG_FAMILY_H1) {
      b.phase[e] = phase;
      reinterpret_cast<float2*>(b.leg_phase)[e] = make_float2(leg_phase[0], leg_phase[1]);
    }
#pragma unroll
    for (int f = 0; f < URLG_MAX_FEET; ++f) {
      if (f < F) {
        b.feet_air_time[(size_t)e * F + f] = air[f];
        b.last_contacts[(size_t)e * F + f] = lastc[f];
      }
    }
    float* o3 = b.rpy + (size_t)e * 3;
    o3[0] = roll; o3[1] = pitch; o3[2] = yaw;
    o3 = b.base_lin_vel + (size_t)e * 3;
    o3[0] = lvx; o3[1] = lvy; o3[2] = lvz;
    o3 = b.base_ang_vel + (size_t)e * 3;
    o3[0] = avx; o3[1] = avy; o3[2] = avz;
    o3 = b.projected_gravity + (size_t)e * 3;
    o3[0] = gx; o3[1] = gy; o3[2] = gz;
    // tail copies :707-708 (after a reset: zeros)
    float dv[A];
#pragma unroll
    for (int j = 0; j < A; ++j) dv[j] = dofs[2 * j + 1];
    st_row<A>(b.last_actions + (size_t)e * A, act);
    st_row<A>(b.last_dof_vel + (size_t)e * A, dv);
  }

  // ---- observation rows leave as one bulk-TMA store per tensor (flat stores for the tail block) ----
  float* gobs = b.obs_buf[obs_slot] + (size_t)env0 * O;
  float* gpriv = c.num_priv ? b.priv_obs_buf[obs_slot] + (size_t)env0 * c.num_priv : nullptr;
  if constexpr (kPadObs) {
    static_assert(!kPadObs || (O % 4 == 0 && PV == 0), "padded observation rows leave as float4");
    __syncwarp();
#pragma unroll
    for (int it = 0; it < O / 4; ++it) {
      const int q = it * 32 + lane, row = q / (O / 4), c4 = q - row * (O / 4);
      const float* sp = sm + L.oObs + row * OS + 4 * c4;
      const float4 v = make_float4(sp[0], sp[1], sp[2], sp[3]);
      if (row < nrows) reinterpret_cast<float4*>(gobs)[q] = v;
    }
  } else if (bulk) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(gobs, sm + L.oObs, 32 * O * sizeof(float));
      if constexpr (FAM == URLG_FAMILY_H1) bulk_s2g(gpriv, sm + L.oPriv, 32 * PV * sizeof(float));
      if constexpr (FAM == URLG_FAMILY_HANDSTAND)                         // privileged_obs_buf = obs.clone() (:172)
        if (gpriv) bulk_s2g(gpriv, sm + L.oObs, 32 * O * sizeof(float));
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
  } else {
    __syncwarp();
    for (int i = lane; i < nrows * O; i += 32) gobs[i] = sm[L.oObs + i];
    if constexpr (FAM == URLG_FAMILY_H1)
      for (int i = lane; i < nrows * PV; i += 32) gpriv[i] = sm[L.oPriv + i];
    if constexpr (FAM == URLG_FAMILY_HANDSTAND)
      if (gpriv) for (int i = lane; i < nrows * O; i += 32) gpriv[i] = sm[L.oObs + i];
  }
#undef HAS
}

// ---------------------------------------------------------------- K2b: finalize
// Launched behind K2 with programmatic stream serialization: its blocks draw the push velocities of their envs (a
// Philox block each, independent of K2) while K2 drains, then wait for K2's completion.  Reads this step's accumulators,
// applies the all-env push (quirk: any due env => every env's xy velocity is redrawn, :544-550), copies last_root_vel
// (:709), publishes extras only if >= 1 env reset (:742-743, 759-768), and clears the other parity's accumulators.
__global__ void __launch_bounds__(128) finalize_step_kernel(const __grid_constant__ urlg_env_desc c,
                                                            const __grid_constant__ urlg_env_buffers b,
                                                            const __grid_constant__ PoolArgs pa,
                                                            StepScratch* __restrict__ scratch,
                                                            const int64_t* __restrict__ step_base, int64_t step_off,
                                                            int num_slots) {
  __shared__ float rs[128 * 13];            // root rows of the block's 128 envs: coalesced in, thread-per-row use (stride 13)
  __shared__ float lr[128 * 6];             // last_root_vel rows, coalesced out
  const int tid = threadIdx.x;
  const int env0 = blockIdx.x * 128;
  const int e = env0 + tid;
  const int A = c.num_dof;
  const int nrows = min(128, c.num_envs - env0);
  const bool live = tid < nrows;
  float px = 0.0f, py = 0.0f;
  const bool may_push = c.push_robots != 0;    // post_physics_step calls _push_robots itself (:700-701), not the callback
  // ---- independent of K2: step counter (advanced by an earlier graph node), push draws (slots 12+A, 13+A) ----
  const int64_t step = eff_step(step_base, step_off);
  if (live && may_push) {
    const float m = c.max_push_vel;
    const float* ovr = b.rng_override;
    float ux, uy;
    const int s0 = 12 + A;
    if (ovr || (s0 & 3) == 3) {
      ux = slot_uniform(c, ovr, c.env_offset + e, step, s0, num_slots, e);
      uy = slot_uniform(c, ovr, c.env_offset + e, step, s0 + 1, num_slots, e);
    } else {                                   // both slots in one Philox block
      const uint4 r = philox4x32_10(make_uint4((uint32_t)(c.env_offset + e), (uint32_t)(step & 0xffffffff),
                                               (uint32_t)((uint64_t)step >> 32), (uint32_t)(s0 >> 2)), c.seed_lo,
                                    c.seed_hi ^ kStreamEnv);
      const int l = s0 & 3;
      ux = u01(l == 0 ? r.x : l == 1 ? r.y : r.z);
      uy = u01(l == 0 ? r.y : l == 1 ? r.z : r.w);
    }
    px = (m - (-m)) * ux + (-m);
    py = (m - (-m)) * uy + (-m);
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");       // K2 of this step has completed and flushed
  const int acc = (int)(step & 1);
  const bool push = scratch->push_flag[acc] != 0;
  float cnt = 0.0f;
#pragma unroll
  for (int k = 0; k < kAccCopies; ++k) cnt += scratch->ep_acc[acc][k][URLG_MAX_TERMS];
  const Seam seam = resolve_seam(c, b, pa);
  for (int i = tid; i < nrows * 13; i += 128) rs[i] = seam.root[(size_t)env0 * 13 + i];
  __syncthreads();
  if (live) {
    float* row = rs + tid * 13;
    // in pool mode the frame row is current except for envs K2 reset this step: their row lives in the bound tensor
    if (pa.enabled && b.reset_buf[e]) {
#pragma unroll
      for (int k = 0; k < 13; ++k) row[k] = b.root_states[(size_t)e * 13 + k];
    }
    if (push) { row[7] = px; row[8] = py; }
#pragma unroll
    for (int k = 0; k < 6; ++k) lr[tid * 6 + k] = row[7 + k];
    if (cnt > 0.0f && c.send_timeouts) b.extras_time_outs[e] = b.time_out_buf[e];
  }
  __syncthreads();
  for (int i = tid; i < nrows * 6; i += 128) b.last_root_vel[(size_t)env0 * 6 + i] = lr[i];
  if (push)            // whole rows back, coalesced (the unpushed columns are rewritten with their own values)
    for (int i = tid; i < nrows * 13; i += 128) b.root_states[(size_t)env0 * 13 + i] = rs[i];
  if (blockIdx.x == 0 && tid < URLG_MAX_TERMS + 1) {
    const int k = tid;
    if (cnt > 0.0f && k < c.num_sums) {
      float sum = 0.0f;
#pragma unroll
      for (int q = 0; q < kAccCopies; ++q) sum += scratch->ep_acc[acc][q][k];
      b.extras_episode[k] = sum / cnt / c.max_episode_length_s;
    }
    if (k == 0) {
      if (b.reset_count) *b.reset_count = scratch->reset_cnt[acc];
      if (b.root_count) *b.root_count = scratch->root_cnt[acc];
    }
    // clear the other parity for step+1 (its last reader was finalize of step-1)
    for (int q = 0; q < kAccCopies; ++q) scratch->ep_acc[acc ^ 1][q][k] = 0.0f;
    if (k == 0) { scratch->push_flag[acc ^ 1] = 0; scratch->reset_cnt[acc ^ 1] = 0; scratch->root_cnt[acc ^ 1] = 0; }
  }
}

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
  resample_commands(c, rng.uniform(9 + A), rng.uniform(10 + A), rng.uniform(11 + A), cmd);
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
    atomicAdd(&scratch->ep_acc[2][0][k], *s);
    *s = 0.0f;
  }
  atomicAdd(&scratch->ep_acc[2][0][URLG_MAX_TERMS], 1.0f);
  if (b.reset_ids) b.reset_ids[i] = e;          // the explicit list is its own compaction (:569-572, 591-594)
  if (b.root_ids) b.root_ids[i] = e;
}

__global__ void __launch_bounds__(256) finalize_reset_kernel(const __grid_constant__ urlg_env_desc c,
                                                             const __grid_constant__ urlg_env_buffers b,
                                                             StepScratch* __restrict__ scratch, int64_t n) {
  // single block: publish, then clear slot 2
  const float cnt = scratch->ep_acc[2][0][URLG_MAX_TERMS];
  if (cnt > 0.0f) {
    if (c.send_timeouts)
      for (int e = threadIdx.x; e < c.num_envs; e += blockDim.x) b.extras_time_outs[e] = b.time_out_buf[e];
    if (threadIdx.x < c.num_sums)
      b.extras_episode[threadIdx.x] = scratch->ep_acc[2][0][threadIdx.x] / cnt / c.max_episode_length_s;
  }
  if (threadIdx.x == 0) {
    if (b.reset_count) *b.reset_count = (int32_t)n;
    if (b.root_count) *b.root_count = (int32_t)n;
  }
  __syncthreads();
  if (threadIdx.x < URLG_MAX_TERMS + 1) scratch->ep_acc[2][0][threadIdx.x] = 0.0f;
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

// rows[ids[i]] of dst <- rows[ids[i]] of src for i < *count: what set_*_state_tensor_indexed does inside a simulator
// whose state lives apart from the exposed tensors (tests' stateful fake gym), with the count on the device.
__global__ void __launch_bounds__(256) scatter_rows_kernel(float* __restrict__ dst, const float* __restrict__ src,
                                                           const int32_t* __restrict__ ids,
                                                           const int32_t* __restrict__ count, int row_floats, int max_n) {
  const int n = min(*count, max_n);
  const int64_t total = (int64_t)n * row_floats;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = ids[i / row_floats];
    const int k = (int)(i % row_floats);
    dst[(size_t)r * row_floats + k] = src[(size_t)r * row_floats + k];
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

using K2Fn = void (*)(const urlg_env_desc, const urlg_env_buffers, const TermTable, const PoolArgs, StepScratch*,
                      const int64_t*, int64_t, int, int);
#define T_(x) (1ull << URLG_T_##x)
// active terms of the five registered task configs (go2_handstand_config.py:62-133)
constexpr unsigned long long kMaskGo2 = T_(ACTION_RATE) | T_(ANG_VEL_XY) | T_(COLLISION) | T_(DOF_ACC) | T_(DOF_POS_LIMITS) |
                                        T_(FEET_AIR_TIME) | T_(LIN_VEL_Z) | T_(TORQUES) | T_(TRACKING_ANG_VEL) | T_(TRACKING_LIN_VEL);
constexpr unsigned long long kMaskHumanoid = T_(ACTION_RATE) | T_(ALIVE) | T_(ANG_VEL_XY) | T_(BASE_HEIGHT) | T_(CONTACT) |
                                             T_(CONTACT_NO_VEL) | T_(DOF_ACC) | T_(DOF_POS_LIMITS) | T_(FEET_SWING_HEIGHT) |
                                             T_(HIP_POS) | T_(LIN_VEL_Z) | T_(ORIENTATION) | T_(TRACKING_ANG_VEL) | T_(TRACKING_LIN_VEL);
constexpr unsigned long long kMaskH1 = kMaskHumanoid | T_(COLLISION);
constexpr unsigned long long kMaskG1 = kMaskHumanoid | T_(DOF_VEL) | T_(TORQUES);          // g1 and h1_2
constexpr unsigned long long kMaskHandstand = T_(ACTION_RATE) | T_(ANG_VEL_XY) | T_(BASE_HEIGHT_HS) | T_(COLLISION) |
    T_(DOF_POS_LIMITS) | T_(DOF_VEL) | T_(FRONT_FEET_CONTACT) | T_(FRONT_FEET_TOGETHER) | T_(FRONT_HIP_NEUTRAL) |
    T_(HIND_FEET_NO_CONTACT) | T_(LIN_VEL_XY) | T_(LIN_VEL_Z) | T_(ORIENTATION_HS) | T_(POSE) | T_(STABILITY) | T_(STAY_STILL) | T_(TORQUES);
#undef T_
static K2Fn k2_for(int num_dof, int family, unsigned long long mask = 0) {
  if (num_dof == 12) {
    if (family == URLG_FAMILY_BASE)
      return mask == kMaskGo2 ? post_physics_kernel<12, URLG_FAMILY_BASE, kMaskGo2> : post_physics_kernel<12, URLG_FAMILY_BASE, 0>;
    if (family == URLG_FAMILY_H1)
      return mask == kMaskG1 ? post_physics_kernel<12, URLG_FAMILY_H1, kMaskG1> : post_physics_kernel<12, URLG_FAMILY_H1, 0>;
    if (family == URLG_FAMILY_HANDSTAND)
      return mask == kMaskHandstand ? post_physics_kernel<12, URLG_FAMILY_HANDSTAND, kMaskHandstand>
                                    : post_physics_kernel<12, URLG_FAMILY_HANDSTAND, 0>;
  } else if (num_dof == 10) {
    if (family == URLG_FAMILY_BASE) return post_physics_kernel<10, URLG_FAMILY_BASE, 0>;
    if (family == URLG_FAMILY_H1)
      return mask == kMaskH1 ? post_physics_kernel<10, URLG_FAMILY_H1, kMaskH1> : post_physics_kernel<10, URLG_FAMILY_H1, 0>;
  }
  return nullptr;
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
      d->num_sums > URLG_MAX_TERMS || d->num_feet > URLG_MAX_FEET || d->num_pen > URLG_MAX_IDX ||
      d->num_term_bodies > URLG_MAX_IDX || d->num_hip > URLG_MAX_IDX)
    return fail(URLG_EINVAL, "urlg_env_create: a table exceeds its ABI bound");
  if (d->control_type != 'P' && d->control_type != 'V' && d->control_type != 'T')
    return fail(URLG_EINVAL, "Unknown controller type: %c", (char)d->control_type);   // NameError at legged_robot.py:670
  if (d->family < URLG_FAMILY_BASE || d->family > URLG_FAMILY_HANDSTAND)
    return fail(URLG_EINVAL, "urlg_env_create: unknown family %d", d->family);
  if (!k2_for(d->num_dof, d->family))
    return fail(URLG_EINVAL, "urlg_env_create: no fused step kernel for num_dof=%d family=%d (built: 12 x {base, h1, "
                "handstand}, 10 x {base, h1})", d->num_dof, d->family);
  const int A = d->num_dof;
  const int expect_obs = d->family == URLG_FAMILY_BASE ? 12 + 3 * A : d->family == URLG_FAMILY_H1 ? 11 + 3 * A : 10 + 3 * A;
  const bool priv_ok = d->family == URLG_FAMILY_BASE ? d->num_priv == 0
                     : d->family == URLG_FAMILY_H1 ? d->num_priv == d->num_obs + 3
                                                   : (d->num_priv == 0 || d->num_priv == d->num_obs);
  if (d->num_obs != expect_obs || !priv_ok)
    return fail(URLG_EINVAL, "urlg_env_create: num_obs=%d num_priv=%d inconsistent with family %d, num_dof %d",
                d->num_obs, d->num_priv, d->family, d->num_dof);
  if (d->family == URLG_FAMILY_HANDSTAND && d->num_feet < 4)
    return fail(URLG_EINVAL, "Expected at least 4 feet for Go2.");                    // go2_handstand_env.py:19
  if (d->family == URLG_FAMILY_H1 && d->num_feet > 2)
    return fail(URLG_EINVAL, "urlg_env_create: the H1-family terms read two feet (h1_env.py:98-123), got %d", d->num_feet);
  if ((int64_t)d->num_envs * URLG_MAX_TERMS >= 0x7fffffffll)
    return fail(URLG_EINVAL, "urlg_env_create: num_envs too large for 32-bit episode-sum offsets");
  if (d->resample_steps <= 0 || d->push_interval <= 0)
    return fail(URLG_EINVAL, "urlg_env_create: resample_steps / push_interval must be positive");
  urlg_env* e = new (std::nothrow) urlg_env();
  if (!e) return fail(URLG_ENOMEM, "urlg_env_create: host allocation failed");
  e->desc = *d;
  e->bound = 0;
  e->scratch = nullptr;
  e->pool.enabled = 0;
  e->pool.calls_off = 0;
  e->tt.mask = 0;
  for (int i = 0; i < URLG_T_COUNT; ++i) { e->tt.scale[i] = 0.0f; e->tt.row[i] = 0; }
  for (int t = 0; t < d->num_terms; ++t) {
    const int id = d->term_ids[t];
    const bool fam_ok = [&] {
      switch (id) {
        case URLG_T_ALIVE: case URLG_T_CONTACT: case URLG_T_CONTACT_NO_VEL: case URLG_T_FEET_SWING_HEIGHT: case URLG_T_HIP_POS:
          return d->family == URLG_FAMILY_H1;
        case URLG_T_BASE_HEIGHT_HS: case URLG_T_ENERGY: case URLG_T_FRONT_FEET_CONTACT: case URLG_T_FRONT_FEET_TOGETHER:
        case URLG_T_FRONT_HIP_NEUTRAL: case URLG_T_HIND_FEET_NO_CONTACT: case URLG_T_LIN_VEL_XY: case URLG_T_ORIENTATION_HS:
        case URLG_T_POSE: case URLG_T_STABILITY: case URLG_T_STAY_STILL:
          return d->family == URLG_FAMILY_HANDSTAND;
        default: return true;
      }
    }();
    if (id < 0 || id >= URLG_T_COUNT || !fam_ok || d->term_sum_row[t] < 0 || d->term_sum_row[t] >= d->num_sums ||
        (t > 0 && id <= d->term_ids[t - 1])) {
      delete e;
      return fail(URLG_EINVAL, "urlg_env_create: reward term #%d (id %d) is unknown, out of order, or not a term of "
                  "env family %d", t, id, d->family);
    }
    e->tt.mask |= 1ull << id;
    e->tt.scale[id] = d->term_scales[t];
    e->tt.row[id] = d->term_sum_row[t];
  }
  auto sq_thr = [](float t) {       // largest s with sqrtf(s) <= t
    float s2 = t * t;
    while (std::sqrt(s2) > t) s2 = std::nextafter(s2, 0.0f);
    while (std::sqrt(std::nextafter(s2, INFINITY)) <= t) s2 = std::nextafter(s2, INFINITY);
    return s2;
  };
  for (int r = 0; r < 10; ++r) {
    e->tt.pk0[r] = d->seed_lo + (uint32_t)r * kPhiloxW0;
    e->tt.pk1[r] = (d->seed_hi ^ kStreamEnv) + (uint32_t)r * kPhiloxW1;
  }
  {
    const int s0 = 14 + d->num_dof, b0 = s0 >> 2;
    e->tt.noise_blocks = 0;
    for (int i = 0; i < d->num_obs; ++i)
      if (d->noise_scale_vec[i] != 0.0f) e->tt.noise_blocks |= 1u << (((s0 + i) >> 2) - b0);
  }
  e->tt.sq_thr_1p0 = sq_thr(1.0f);
  e->tt.sq_thr_0p1 = sq_thr(0.1f);
  URLG_CUDA(cudaMalloc(&e->scratch, sizeof(StepScratch)));
  URLG_CUDA(cudaMemset(e->scratch, 0, sizeof(StepScratch)));
  e->smem_bytes = make_layout(e->desc).total * (int)sizeof(float);
  K2Fn fn = k2_for(d->num_dof, d->family, e->tt.mask);
  URLG_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, e->smem_bytes > 48 * 1024 ? e->smem_bytes : 48 * 1024));
  URLG_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
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
  const int fam = env->desc.family;
  if (fam == URLG_FAMILY_H1 && (!b->rigid_body_states || !b->priv_obs_buf[0] || !b->priv_obs_buf[1] || !b->phase || !b->leg_phase))
    return fail(URLG_EINVAL, "urlg_env_bind: H1-family env needs rigid_body_states, priv_obs_buf, phase, leg_phase");
  if (fam == URLG_FAMILY_HANDSTAND && !b->rigid_body_states)
    return fail(URLG_EINVAL, "urlg_env_bind: the handstand env needs rigid_body_states (feet positions)");
  if (env->desc.num_priv && (!b->priv_obs_buf[0] || !b->priv_obs_buf[1]))
    return fail(URLG_EINVAL, "urlg_env_bind: num_priv > 0 needs both priv_obs_buf slots");
  {
    const void* al[] = {b->root_states, b->dof_state, b->contact_forces, b->actions, b->last_actions, b->last_dof_vel,
                        b->torques, b->obs_buf[0], b->obs_buf[1], b->priv_obs_buf[0], b->priv_obs_buf[1], b->commands,
                        b->leg_phase};
    uintptr_t bits = 0;
    for (size_t i = 0; i < sizeof(al) / sizeof(al[0]); ++i) bits |= (uintptr_t)al[i];
    if (bits & 15)
      return fail(URLG_EINVAL, "urlg_env_bind: root_states, dof_state, contact_forces, actions, last_actions, last_dof_vel, "
                  "torques, obs / privileged obs slots, commands and leg_phase must be 16-byte aligned");
  }
  env->bufs = *b;
  env->bound = 1;
  return URLG_OK;
}

extern "C" int urlg_env_bind_frame_pool(urlg_env* env, const urlg_frame_pool* pool) {
  if (!env) return fail(URLG_EINVAL, "urlg_env_bind_frame_pool: null env");
  if (!pool) { env->pool.enabled = 0; return URLG_OK; }
  if (!pool->root || !pool->dof || !pool->contact || !pool->calls_base || pool->num_frames <= 0 || pool->substeps <= 0 ||
      pool->num_frames > 0x7fffffff || pool->substeps > 0x7fffffff)
    return fail(URLG_EINVAL, "urlg_env_bind_frame_pool: bad pool");
  if (env->desc.family != URLG_FAMILY_BASE && !pool->rigid_body)
    return fail(URLG_EINVAL, "urlg_env_bind_frame_pool: this env family reads rigid_body_states");
  const uintptr_t bits = (uintptr_t)pool->root | (uintptr_t)pool->dof | (uintptr_t)pool->contact;
  if (pool->frame_stride < 0 || (pool->frame_stride & 3))
    return fail(URLG_EINVAL, "urlg_env_bind_frame_pool: frame_stride must be 0 or a positive multiple of 4 floats");
  if ((bits & 15) || (env->desc.num_envs & 3))
    return fail(URLG_EINVAL, "urlg_env_bind_frame_pool: pool tensors must be 16-byte aligned and num_envs a multiple of 4");
  env->pool.pool = *pool;
  env->pool.enabled = 1;
  return URLG_OK;
}

extern "C" void urlg_env_destroy(urlg_env* env) {
  if (!env) return;
  if (env->scratch) cudaFree(env->scratch);
  delete env;
}

extern "C" int urlg_env_num_rng_slots(const urlg_env* env) {
  if (!env) return 0;
  return 14 + env->desc.num_dof + env->desc.num_obs + (env->desc.family == URLG_FAMILY_HANDSTAND ? 4 : 0);
}

static int pd_launch(urlg_env* env, const float* raw_actions, int64_t calls_off, void* stream) {
  if (!env || !env->bound) return fail(URLG_EINVAL, "urlg_pd_torques: env not bound");
  const int64_t total = (int64_t)env->desc.num_envs * env->desc.num_dof;
  if (total >= 0x7fffffffll) return fail(URLG_EINVAL, "urlg_pd_torques: num_envs * num_dof must be < 2^31");
  PoolArgs pa = env->pool;
  pa.calls_off = calls_off;
  if (((uintptr_t)raw_actions & 15) == 0 && (env->desc.num_dof & 3) == 0) {
    const uint32_t total4 = (uint32_t)(total / 4);
    pd_torques_vec4_kernel<<<(total4 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(env->desc, env->bufs, pa, raw_actions, total4);
  } else {
    pd_torques_kernel<<<grid_for(total, 256, 64), 256, 0, (cudaStream_t)stream>>>(env->desc, env->bufs, pa, raw_actions);
  }
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_pd_torques(urlg_env* env, const float* raw_actions, void* stream) {
  if (env && env->pool.enabled) return fail(URLG_EINVAL, "urlg_pd_torques: a frame pool is bound; call urlg_pd_torques_at");
  return pd_launch(env, raw_actions, 0, stream);
}
extern "C" int urlg_pd_torques_at(urlg_env* env, const float* raw_actions, int64_t calls_offset, void* stream) {
  return pd_launch(env, raw_actions, calls_offset, stream);
}

static int post_launch(urlg_env* env, const int64_t* step_base, int64_t step_offset, int obs_slot, int64_t calls_off,
                       void* stream) {
  if (!env || !env->bound) return fail(URLG_EINVAL, "urlg_post_physics: env not bound");
  if (obs_slot != 0 && obs_slot != 1) return fail(URLG_EINVAL, "urlg_post_physics: obs_slot must be 0 or 1");
  const int N = env->desc.num_envs;
  const int slots = urlg_env_num_rng_slots(env);
  cudaStream_t st = (cudaStream_t)stream;
  PoolArgs pa = env->pool;
  pa.calls_off = calls_off;
  K2Fn fn = k2_for(env->desc.num_dof, env->desc.family, env->tt.mask);
  fn<<<(N + 31) / 32, 32, env->smem_bytes, st>>>(env->desc, env->bufs, env->tt, pa, env->scratch, step_base, step_offset,
                                                 obs_slot, slots);
  URLG_LAUNCH_CHECK();
  URLG_CUDA(launch_pdl(finalize_step_kernel, dim3((N + 127) / 128), dim3(128), 0, st, env->desc, env->bufs, pa,
                       env->scratch, step_base, step_offset, slots));
  ++urlg::g_launches;
  return URLG_OK;
}

extern "C" int urlg_post_physics(urlg_env* env, const int64_t* step_base, int64_t step_offset, int obs_slot,
                                 void* stream) {
  if (env && env->pool.enabled) return fail(URLG_EINVAL, "urlg_post_physics: a frame pool is bound; call urlg_post_physics_at");
  return post_launch(env, step_base, step_offset, obs_slot, 0, stream);
}
extern "C" int urlg_post_physics_at(urlg_env* env, const int64_t* step_base, int64_t step_offset, int obs_slot,
                                    int64_t calls_offset, void* stream) {
  return post_launch(env, step_base, step_offset, obs_slot, calls_offset, stream);
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
  finalize_reset_kernel<<<1, 256, 0, st>>>(env->desc, env->bufs, env->scratch, n);
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

extern "C" int urlg_scatter_rows(float* dst, const float* src, const int32_t* ids, const int32_t* count, int row_floats,
                                 int max_rows, void* stream) {
  if (!dst || !src || !ids || !count || row_floats <= 0 || max_rows <= 0)
    return fail(URLG_EINVAL, "urlg_scatter_rows: bad argument");
  scatter_rows_kernel<<<grid_for((int64_t)max_rows * row_floats, 256, 2), 256, 0, (cudaStream_t)stream>>>(
      dst, src, ids, count, row_floats, max_rows);
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
