// ActorCritic rollout side ([UPSTREAM rsl_rl v1.0.2] modules/actor_critic.py, algorithms/ppo.py act;
// SURVEY.md §8(a).6 R5/R7, kernels K3/K4): workspace carving, bf16 weight images, MLP forward on the
// tcgen05 GEMM (bias+ELU fused in its epilogue), Gaussian sampling with counter-based Philox, and the
// transition store fused into the prep / sample kernels (no separate add_transitions pass).
#include "policy.cuh"
#include "philox.cuh"
#include <new>

namespace urlg {
namespace tc {
int launch_gemm(const urlg_gemm_desc* d, cudaStream_t st);
int launch_gemm_group(const urlg_gemm_desc* const* ds, int n, cudaStream_t st);
}

// ---------------------------------------------------------------- workspace carving
// bf16 slabs come from their own region so that, in split precision, the whole region can be duplicated
// (hi half, then lo half `off16` bytes later); everything else comes from the fp32 region behind it.
struct Carver {
  uint8_t *base16, *base32; int64_t off16, off32;
  template <typename T> T* take(int64_t n) {
    constexpr bool is16 = sizeof(T) == 2;
    int64_t& off = is16 ? off16 : off32;
    uint8_t* base = is16 ? base16 : base32;
    off = (off + 1023) & ~int64_t(1023);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * (int64_t)sizeof(T);
    return p;
  }
};

static int carve_net(Carver& c, NetWs& n, const urlg_mlp_desc& d, int R) {
  n.d = d;
  const int L = d.num_layers;
  if (L < 1 || L > URLG_MAX_LAYERS) return fail(URLG_EINVAL, "mlp: num_layers=%d out of range", L);
  for (int l = 0; l < L; ++l) {
    if (d.dims[l] <= 0 || d.dims[l + 1] <= 0) return fail(URLG_EINVAL, "mlp: bad dims");
    if (l > 0 && (d.dims[l] & 7)) return fail(URLG_EINVAL, "mlp: hidden width %d must be a multiple of 8", d.dims[l]);
    n.kpad[l] = round_up(d.dims[l], 8);
    n.w[l] = c.take<bf16>((int64_t)d.dims[l + 1] * n.kpad[l]);
  }
  n.x = c.take<bf16>((int64_t)R * n.kpad[0]);
  for (int l = 1; l < L; ++l) {
    n.h[l] = c.take<bf16>((int64_t)R * d.dims[l]);
    n.dz[l] = c.take<bf16>((int64_t)R * d.dims[l]);
  }
  n.out = c.take<float>((int64_t)R * d.dims[L]);
  n.opad = round_up(d.dims[L], 8);
  n.dout = c.take<bf16>((int64_t)R * n.opad);
  return URLG_OK;
}

static void carve_lstm(Carver& c, LstmWs& w, int in, int H, const int64_t* off, int R) {
  w.in = in; w.in_pad = round_up(in, 8); w.H = H; w.kcat = w.in_pad + H;
  for (int i = 0; i < 4; ++i) w.off[i] = off[i];
  w.w = c.take<bf16>((int64_t)4 * H * w.kcat);
  w.bias = c.take<float>(4 * H);
  w.xcat = c.take<bf16>((int64_t)R * w.kcat);
  w.gates = c.take<float>((int64_t)R * 4 * H);
  w.c = c.take<float>((int64_t)2 * R * H);
  w.h0 = c.take<float>((int64_t)R * H);
  w.dgates = c.take<bf16>((int64_t)R * 4 * H);
  w.dh = c.take<float>((int64_t)R * H);
  w.dc = c.take<float>((int64_t)R * H);
  w.dhrec = c.take<float>((int64_t)R * H);
  w.dx_head = c.take<bf16>((int64_t)R * H);
}

static int carve_pass(urlg_policy* p, Carver& c);

static int carve(urlg_policy* p, uint8_t* base, int64_t* total) {
  Carver c{nullptr, nullptr, 0, 0};
  if (int rc = carve_pass(p, c)) return rc;
  const int64_t s16 = (c.off16 + 1023) & ~int64_t(1023), s32 = (c.off32 + 1023) & ~int64_t(1023);
  const int copies = p->desc.precision ? 2 : 1;
  p->lo_off = p->desc.precision ? s16 / 2 : 0;      // in bf16 elements
  *total = copies * s16 + s32;
  if (base) {
    Carver d{base, base + copies * s16, 0, 0};
    return carve_pass(p, d);
  }
  return URLG_OK;
}

static int carve_pass(urlg_policy* p, Carver& c) {
  const urlg_policy_desc& d = p->desc;
  if (d.max_rows <= 0) return fail(URLG_EINVAL, "policy: max_rows must be positive");
  if (d.num_actions <= 0 || d.num_actions > 32) return fail(URLG_EINVAL, "policy: num_actions=%d unsupported (1..32)", d.num_actions);
  p->R = d.max_rows;
  p->T = d.max_seq > 0 ? d.max_seq : 1;
  if (int rc = carve_net(c, p->net[0], d.actor, p->R)) return rc;
  if (int rc = carve_net(c, p->net[1], d.critic, p->R)) return rc;
  if (d.actor.dims[d.actor.num_layers] != d.num_actions || d.critic.dims[d.critic.num_layers] != 1)
    return fail(URLG_EINVAL, "policy: actor must end in num_actions, critic in 1");
  if (d.recurrent) {
    if (d.rnn_hidden <= 0 || (d.rnn_hidden & 15)) return fail(URLG_EINVAL, "policy: rnn_hidden must be a multiple of 16");
    if (d.actor.dims[0] != d.rnn_hidden || d.critic.dims[0] != d.rnn_hidden)
      return fail(URLG_EINVAL, "policy: recurrent heads take the LSTM output");
    carve_lstm(c, p->lstm[0], d.num_actor_obs, d.rnn_hidden, d.lstm_a_off, p->R);
    carve_lstm(c, p->lstm[1], d.num_critic_obs, d.rnn_hidden, d.lstm_c_off, p->R);
  } else if (d.actor.dims[0] != d.num_actor_obs || d.critic.dims[0] != d.num_critic_obs) {
    return fail(URLG_EINVAL, "policy: MLP input widths must equal the observation widths");
  }
  p->red = c.take<float>(256 + 8);   // 256 block partials of the gradient norm + the last-block ticket (red[256], as uint)
  return URLG_OK;
}

// ---------------------------------------------------------------- bf16 weight images
__global__ void __launch_bounds__(256)
weight_image_kernel(bf16* __restrict__ dst, int64_t lo_off, const float* __restrict__ src, int rows, int cols, int kpad,
                    int dst_col0, int ld) {
  const int64_t total = (int64_t)rows * kpad;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / kpad), c = (int)(i % kpad);
    store_split(dst, lo_off, (size_t)r * ld + dst_col0 + c, c < cols ? src[(size_t)r * cols + c] : 0.0f);
  }
}
__global__ void bias_sum_kernel(float* __restrict__ dst, const float* __restrict__ a, const float* __restrict__ b, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = a[i] + b[i];
}

// All weight images of a policy in ONE launch (it runs after every Adam step): a table of up to 16 segments, each a
// [rows, kpad] image of a [rows, cols] fp32 matrix, or (rows == 0) a bias sum b_ih + b_hh of `cols` elements.
struct ImgSeg {
  bf16* dst; float* fdst; const float* src; const float* src2;
  int rows, cols, kpad, dst_col0, ld;
  int64_t start;                       // first flat element of this segment
};
struct ImgTable { int n; int64_t total; ImgSeg seg[16]; };

__global__ void __launch_bounds__(256) weight_images_kernel(const __grid_constant__ ImgTable tb, int64_t lo_off) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tb.total; i += (int64_t)gridDim.x * blockDim.x) {
    int k = 0;
    while (k + 1 < tb.n && i >= tb.seg[k + 1].start) ++k;
    const ImgSeg& sg = tb.seg[k];
    const int64_t j = i - sg.start;
    if (sg.rows == 0) {
      sg.fdst[j] = sg.src[j] + sg.src2[j];
    } else {
      const int r = (int)(j / sg.kpad), c = (int)(j % sg.kpad);
      store_split(sg.dst, lo_off, (size_t)r * sg.ld + sg.dst_col0 + c, c < sg.cols ? sg.src[(size_t)r * sg.cols + c] : 0.0f);
    }
  }
}

int sync_weights(urlg_policy* p, cudaStream_t st) {
  ImgTable tb;
  tb.n = 0; tb.total = 0;
  auto add = [&](bf16* dst, float* fdst, const float* src, const float* src2, int rows, int cols, int kpad, int col0, int ld) {
    ImgSeg& sg = tb.seg[tb.n++];
    sg.dst = dst; sg.fdst = fdst; sg.src = src; sg.src2 = src2;
    sg.rows = rows; sg.cols = cols; sg.kpad = kpad; sg.dst_col0 = col0; sg.ld = ld;
    sg.start = tb.total;
    tb.total += rows ? (int64_t)rows * kpad : cols;
  };
  for (int k = 0; k < 2; ++k) {
    NetWs& n = p->net[k];
    if (n.d.num_layers + (p->desc.recurrent ? 3 : 0) + tb.n > 16) return fail(URLG_EINVAL, "sync_weights: too many layers");
    for (int l = 0; l < n.d.num_layers; ++l)
      add(n.w[l], nullptr, p->params + n.d.w_off[l], nullptr, n.d.dims[l + 1], n.d.dims[l], n.kpad[l], 0, n.kpad[l]);
    if (p->desc.recurrent) {
      LstmWs& w = p->lstm[k];
      const int G = 4 * w.H;
      add(w.w, nullptr, p->params + w.off[0], nullptr, G, w.in, w.in_pad, 0, w.kcat);
      add(w.w, nullptr, p->params + w.off[1], nullptr, G, w.H, w.H, w.in_pad, w.kcat);
      add(nullptr, w.bias, p->params + w.off[2], p->params + w.off[3], 0, G, 0, 0, 0);
    }
  }
  int blocks = (int)((tb.total + 255) / 256);
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  weight_images_kernel<<<blocks, 256, 0, st>>>(tb, p->lo_off);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

// ---------------------------------------------------------------- MLP forward (x already staged in n.x)
// keep_hidden: activations stay in n.h[] (needed by the backward pass); the rollout reuses the same slabs.
static urlg_gemm_desc forward_desc(urlg_policy* p, NetWs& n, int l, int64_t rows, int max_ctas) {
  const int L = n.d.num_layers;
  urlg_gemm_desc g = {};
  g.A = l == 0 ? (const void*)n.x : (const void*)n.h[l];
  g.lda = l == 0 ? n.kpad[0] : n.d.dims[l];
  g.B = n.w[l]; g.ldb = n.kpad[l];
  g.M = (int)rows; g.N = n.d.dims[l + 1]; g.K = n.kpad[l];
  g.bias = p->params + n.d.b_off[l];
  if (l + 1 < L) { g.out_bf16 = n.h[l + 1]; g.ld_bf16 = n.d.dims[l + 1]; g.flags = URLG_GEMM_ELU; }
  else { g.out_f32 = n.out; g.ld_f32 = n.d.dims[L]; }
  g.max_ctas = max_ctas;
  with_lo(p, g);
  return g;
}

int mlp_forward(urlg_policy* p, NetWs& n, int64_t rows, cudaStream_t st, int max_ctas) {
  for (int l = 0; l < n.d.num_layers; ++l) {
    const urlg_gemm_desc g = forward_desc(p, n, l, rows, max_ctas);
    if (int rc = tc::launch_gemm(&g, st)) return rc;
  }
  return URLG_OK;
}

// Actor and critic side by side: layer l of both networks is ONE launch (two problems in the GEMM's tile space), so
// a forward of both is max(La, Lc) launches instead of La + Lc; the ~4 us fixed cost of a launch is paid once per layer.
int mlp_forward_pair(urlg_policy* p, NetWs& a, NetWs& c, int64_t rows, cudaStream_t st) {
  const int La = a.d.num_layers, Lc = c.d.num_layers;
  for (int l = 0; l < (La > Lc ? La : Lc); ++l) {
    urlg_gemm_desc g[2];
    const urlg_gemm_desc* ptr[2];
    int n = 0;
    if (l < La) { g[n] = forward_desc(p, a, l, rows, 0); ptr[n] = &g[n]; ++n; }
    if (l < Lc) { g[n] = forward_desc(p, c, l, rows, 0); ptr[n] = &g[n]; ++n; }
    if (int rc = tc::launch_gemm_group(ptr, n, st)) return rc;
  }
  return URLG_OK;
}

// ---------------------------------------------------------------- prep: obs -> storage slab + bf16 staging
__global__ void __launch_bounds__(256)
stage_obs_kernel(const float* __restrict__ src, int64_t rows, int cols, float* __restrict__ store, bf16* __restrict__ dst,
                 int64_t lo_off, int kpad, int ld) {
  const int64_t total = rows * kpad;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / kpad;
    const int c = (int)(i % kpad);
    float v = 0.0f;
    if (c < cols) {
      v = src[r * cols + c];
      if (store) store[r * cols + c] = v;
    }
    store_split(dst, lo_off, (size_t)(r * ld + c), v);
  }
}

// Actor and critic observations in one launch (rollout: saves a launch per env step); same element mapping as above.
__global__ void __launch_bounds__(256)
stage_obs_pair_kernel(const float* __restrict__ src_a, int cols_a, float* __restrict__ store_a, bf16* __restrict__ dst_a, int kpad_a,
                      const float* __restrict__ src_c, int cols_c, float* __restrict__ store_c, bf16* __restrict__ dst_c, int kpad_c,
                      int64_t rows, int64_t lo_off) {
  const int64_t total_a = rows * kpad_a, total = total_a + rows * kpad_c;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const bool crit = i >= total_a;
    const int64_t k = crit ? i - total_a : i;
    const int kpad = crit ? kpad_c : kpad_a, cols = crit ? cols_c : cols_a;
    const float* src = crit ? src_c : src_a;
    float* store = crit ? store_c : store_a;
    const int64_t r = k / kpad;
    const int c = (int)(k % kpad);
    float v = 0.0f;
    if (c < cols) {
      v = src[r * cols + c];
      if (store) store[r * cols + c] = v;
    }
    store_split(crit ? dst_c : dst_a, lo_off, (size_t)(r * kpad + c), v);
  }
}

int stage_obs(const float* src, int64_t rows, int cols, float* store, bf16* dst, int64_t lo_off, int kpad, int ld,
              cudaStream_t st) {
  const int64_t total = rows * kpad;
  int blocks = (int)((total + 255) / 256);
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  stage_obs_kernel<<<blocks, 256, 0, st>>>(src, rows, cols, store, dst, lo_off, kpad, ld);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

// ---------------------------------------------------------------- sample + log-prob + transition store
// Normal(mu, std).sample(): eps from Philox (Box-Muller on slots 2j, 2j+1 of stream kStreamPolicy, counter
// (global env, step)) or injected.  log_prob follows torch.distributions.Normal.log_prob term by term.
// One thread per (env, action): 32 envs x A threads per block, all loads / stores contiguous; the per-action log-prob
// terms meet in shared memory and are summed in action order by the env's first thread (the order a per-env loop uses).
constexpr int kSampleEnvs = 32;
__global__ void __launch_bounds__(1024)
sample_store_kernel(const float* __restrict__ mu_in, const float* __restrict__ val_in, const float* __restrict__ std_param,
                    const float* __restrict__ noise, urlg_storage s, int t, float* __restrict__ actions_out, int A,
                    uint32_t seed_lo, uint32_t seed_hi, int64_t env_offset, const int64_t* step_base, int64_t step_off) {
  extern __shared__ float terms[];                       // [kSampleEnvs * A]
  const int e = threadIdx.x / A, j = threadIdx.x - e * A;
  const int i = blockIdx.x * kSampleEnvs + e;
  if (i < s.N) {
    const size_t row = (size_t)t * s.N + i;
    const float mu = mu_in[(size_t)i * A + j];
    const float sd = mu * 0.0f + std_param[j];
    float eps;
    if (noise) {
      eps = noise[(size_t)i * A + j];
    } else {
      SlotRng rng;
      rng.init(seed_lo, seed_hi, kStreamPolicy, env_offset + i, eff_step(step_base, step_off), nullptr);
      const float u1 = 1.0f - rng.uniform(2 * j), u2 = rng.uniform(2 * j + 1);   // u1 in (0,1]
      eps = sqrtf(-2.0f * logf(u1)) * cosf(6.28318530717958647692f * u2);
    }
    const float a = mu + sd * eps;
    const float d = a - mu;
    terms[threadIdx.x] = -(d * d) / (2.0f * (sd * sd)) - logf(sd) - 0.918938533204672741780f;
    s.actions[row * A + j] = a;
    s.mu[row * A + j] = mu;
    s.sigma[row * A + j] = sd;
    actions_out[(size_t)i * A + j] = a;
  }
  __syncthreads();
  if (i < s.N && j == 0) {
    const size_t row = (size_t)t * s.N + i;
    float lp = 0.0f;
    for (int k = 0; k < A; ++k) lp += terms[e * A + k];
    s.actions_log_prob[row] = lp;
    s.values[row] = val_in[i];
  }
}

}  // namespace urlg

using namespace urlg;

extern "C" int64_t urlg_policy_workspace_bytes(const urlg_policy_desc* desc) {
  if (!desc) return fail(URLG_EINVAL, "urlg_policy_workspace_bytes: null desc");
  urlg_policy tmp = {};
  tmp.desc = *desc;
  int64_t total = 0;
  if (int rc = carve(&tmp, nullptr, &total)) return rc;
  return total;
}

extern "C" int urlg_policy_create(const urlg_policy_desc* desc, float* params, float* grads, float* exp_avg,
                                  float* exp_avg_sq, float* stats, void* workspace, int64_t workspace_bytes,
                                  urlg_policy** out) {
  if (!desc || !params || !grads || !exp_avg || !exp_avg_sq || !stats || !workspace || !out)
    return fail(URLG_EINVAL, "urlg_policy_create: null argument");
  if ((uintptr_t)workspace & 1023) return fail(URLG_EINVAL, "urlg_policy_create: workspace must be 1024-byte aligned");
  urlg_policy* p = new (std::nothrow) urlg_policy();
  if (!p) return fail(URLG_ENOMEM, "urlg_policy_create: out of host memory");
  p->desc = *desc;
  p->params = params; p->grads = grads; p->m = exp_avg; p->v = exp_avg_sq; p->stats = stats;
  int64_t total = 0;
  int rc = carve(p, static_cast<uint8_t*>(workspace), &total);
  if (!rc && total > workspace_bytes) rc = fail(URLG_EINVAL, "urlg_policy_create: workspace too small (%lld < %lld)",
                                                (long long)workspace_bytes, (long long)total);
  if (rc) { delete p; return rc; }
  if (cudaMemset(p->red + 256, 0, 8 * sizeof(float)) != cudaSuccess) {       // ticket of grad_norm_scalars_kernel starts at 0
    delete p;
    return fail(URLG_ECUDA, "urlg_policy_create: workspace memset: %s", cudaGetErrorString(cudaGetLastError()));
  }
  p->side = nullptr; p->ev_fork = nullptr; p->ev_join = nullptr; p->fz = nullptr; p->gacc = nullptr; p->wpart = nullptr; p->peer_bound = 0; p->norm_partials = nullptr; p->norm_partials_cap = 0; p->norm_done = 0;
  if (cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    urlg_policy_destroy(p);
    return fail(URLG_ECUDA, "urlg_policy_create: side stream / events: %s", cudaGetErrorString(cudaGetLastError()));
  }
  if (cudaMalloc(&p->gacc, ((size_t)desc->num_params + 4) * sizeof(long long)) != cudaSuccess ||
      cudaMemset(p->gacc, 0, ((size_t)desc->num_params + 4) * sizeof(long long)) != cudaSuccess) {
    urlg_policy_destroy(p);
    return fail(URLG_ECUDA, "urlg_policy_create: gradient accumulators: %s", cudaGetErrorString(cudaGetLastError()));
  }
  {
    // slabs for the split-K weight-gradient partials: a wgrad is split so that tiles x splits fills one wave of the
    // persistent grid (ppo_kernels.cu wgrad_desc), i.e. at most kNumSMs / tiles splits of [out, in] floats per layer
    int64_t total_f = 0;
    for (int q = 0; q < 2; ++q) {
      const urlg_mlp_desc& m = p->net[q].d;
      for (int l = 0; l < URLG_MAX_LAYERS; ++l) { p->wpart_off[q][l] = 0; p->wpart_cap[q][l] = 0; p->wpart_splits[q][l] = 0; }
      for (int l = 0; l < m.num_layers; ++l) {
        const int nout = m.dims[l + 1], nin = m.dims[l];
        const int bn = nin > 128 ? 128 : round_up(nin, 64);
        const int tiles = ((nout + 127) / 128) * ((nin + bn - 1) / bn);
        const int cap = tiles >= kNumSMs ? 1 : kNumSMs / tiles;
        p->wpart_off[q][l] = total_f;
        p->wpart_cap[q][l] = cap;
        total_f += ((int64_t)cap * nout * nin + 3) / 4 * 4;
      }
    }
    if (total_f > 0 && cudaMalloc(&p->wpart, (size_t)total_f * sizeof(float)) != cudaSuccess) {
      urlg_policy_destroy(p);
      return fail(URLG_ECUDA, "urlg_policy_create: weight-gradient partial slabs (%lld MB): %s",
                  (long long)(total_f * 4 >> 20), cudaGetErrorString(cudaGetLastError()));
    }
  }
  p->norm_partials_cap = (int)((desc->num_params + 4 + 255) / 256 + (desc->num_params / 4 + 255) / 256 + 2 * URLG_MAX_LAYERS + 8);
  if (cudaMalloc(&p->norm_partials, (size_t)p->norm_partials_cap * sizeof(float)) != cudaSuccess) {
    urlg_policy_destroy(p);
    return fail(URLG_ECUDA, "urlg_policy_create: norm partials: %s", cudaGetErrorString(cudaGetLastError()));
  }
  if (int rc2 = fused_build(p)) { urlg_policy_destroy(p); return rc2; }
  *out = p;
  return URLG_OK;
}

extern "C" void urlg_policy_destroy(urlg_policy* p) {
  if (!p) return;
  fused_free(p);
  if (p->gacc) cudaFree(p->gacc);
  if (p->wpart) cudaFree(p->wpart);
  if (p->norm_partials) cudaFree(p->norm_partials);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  if (p->side) cudaStreamDestroy(p->side);
  delete p;
}

extern "C" int urlg_policy_sync_weights(urlg_policy* p, void* stream) {
  if (!p) return fail(URLG_EINVAL, "urlg_policy_sync_weights: null policy");
  return sync_weights(p, (cudaStream_t)stream);
}

namespace urlg {
// lstm_kernels.cu: one rollout-mode LSTM step for Memory k (0 actor / 1 critic): saves the pre-step state
// to (save_h, save_c) if given, h/c [rows,H] fp32 updated in place, bf16 h' staged as the head's input.
int lstm_rollout_step(urlg_policy* p, int k, const float* x, int64_t rows, float* store_x, float* h, float* c,
                      float* save_h, float* save_c, cudaStream_t st);
}

extern "C" int urlg_policy_act(urlg_policy* p, const float* obs, const float* critic_obs, const urlg_storage* s, int t,
                               float* hidden, const float* noise, float* actions_out, const int64_t* step_base,
                               int64_t step_offset, void* stream) {
  if (!p || !obs || !s || !actions_out) return fail(URLG_EINVAL, "urlg_policy_act: null argument");
  if (t < 0 || t >= s->T) return fail(URLG_EINVAL, "Rollout buffer overflow");
  const urlg_policy_desc& d = p->desc;
  const int N = s->N;
  if (N > p->R) return fail(URLG_EINVAL, "urlg_policy_act: num_envs %d exceeds max_rows %d", N, p->R);
  if (s->num_obs != d.num_actor_obs || s->num_actions != d.num_actions ||
      (s->privileged_observations && s->num_priv != d.num_critic_obs))
    return fail(URLG_EINVAL, "urlg_policy_act: storage shape does not match the policy");
  cudaStream_t st = (cudaStream_t)stream;
  const float* cobs = critic_obs ? critic_obs : obs;
  float* store_o = s->observations + (size_t)t * N * s->num_obs;
  float* store_p = s->privileged_observations ? s->privileged_observations + (size_t)t * N * s->num_priv : nullptr;
  if (d.recurrent) {
    if (!hidden) return fail(URLG_EINVAL, "urlg_policy_act: recurrent policy needs hidden state");
    if (!s->hid_a_h || !s->hid_a_c || !s->hid_c_h || !s->hid_c_c)
      return fail(URLG_EINVAL, "urlg_policy_act: recurrent storage lacks saved hidden states");
  }
  if (!d.recurrent && p->fz) {
    // ONE launch: observation staging + slab copy, every layer of both networks, sampling / log-prob / transition store
    if (d.num_actions > 16) return fail(URLG_EINVAL, "urlg_policy_act: more than 16 actions");
    return fused_act(p, obs, cobs, s, t, noise, actions_out, step_base, step_offset, st);
  }
  if (!d.recurrent) {
    // both observation stagings, then one paired GEMM launch per layer on the caller's stream
    {
      const int ka = p->net[0].kpad[0], kc = p->net[1].kpad[0];
      const int64_t total = (int64_t)N * (ka + kc);
      int blocks = (int)((total + 255) / 256);
      if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
      stage_obs_pair_kernel<<<blocks, 256, 0, st>>>(obs, d.num_actor_obs, store_o, p->net[0].x, ka, cobs, d.num_critic_obs,
                                                    store_p, p->net[1].x, kc, N, p->lo_off);
      URLG_LAUNCH_CHECK();
    }
    if (int rc = mlp_forward_pair(p, p->net[0], p->net[1], N, st)) return rc;
  } else {
    // fork: the critic's LSTM step + head on the side stream, the actor's on the caller's.  Each chain is a handful of
    // small launches (staging, gate GEMM, cell, 2 head GEMMs), so two concurrent chains beat pairing the heads after
    // a join (measured: 2.34 vs 2.52 ms per 24-step rollout); each GEMM is capped to half the SMs so both fit.
    cudaStream_t sc = side_stream(p, st);
    URLG_CUDA(cudaEventRecord(p->ev_fork, st));
    URLG_CUDA(cudaStreamWaitEvent(sc, p->ev_fork, 0));
    const int cap = N <= 8192 ? kNumSMs / 2 : 0;
    const size_t HS = (size_t)N * d.rnn_hidden, slab = (size_t)t * HS;
    int rc_a = lstm_rollout_step(p, 0, obs, N, store_o, hidden, hidden + HS, s->hid_a_h + slab, s->hid_a_c + slab, st);
    int rc_c = lstm_rollout_step(p, 1, cobs, N, store_p, hidden + 2 * HS, hidden + 3 * HS, s->hid_c_h + slab, s->hid_c_c + slab, sc);
    if (!rc_a) rc_a = mlp_forward(p, p->net[0], N, st, cap);
    if (!rc_c) rc_c = mlp_forward(p, p->net[1], N, sc, cap);
    // join (always, so a capture in progress is never left with an unjoined stream)
    URLG_CUDA(cudaEventRecord(p->ev_join, sc));
    URLG_CUDA(cudaStreamWaitEvent(st, p->ev_join, 0));
    if (rc_a) return rc_a;
    if (rc_c) return rc_c;
  }
  if (d.num_actions > 32) return fail(URLG_EINVAL, "urlg_policy_act: more than 32 actions");
  sample_store_kernel<<<(N + kSampleEnvs - 1) / kSampleEnvs, kSampleEnvs * d.num_actions,
                        kSampleEnvs * d.num_actions * sizeof(float), st>>>(p->net[0].out, p->net[1].out, p->params + d.std_off, noise, *s, t,
                                                       actions_out, d.num_actions, d.seed_lo, d.seed_hi, d.env_offset,
                                                       step_base, step_offset);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_policy_forward(urlg_policy* p, int which, const float* x, int64_t rows, float* hidden_h,
                                   float* hidden_c, float* out, void* stream) {
  if (!p || !x || !out || (which != 0 && which != 1)) return fail(URLG_EINVAL, "urlg_policy_forward: bad argument");
  if (rows <= 0 || rows > p->R) return fail(URLG_EINVAL, "urlg_policy_forward: rows %lld outside 1..%d", (long long)rows, p->R);
  cudaStream_t st = (cudaStream_t)stream;
  const urlg_policy_desc& d = p->desc;
  NetWs& n = p->net[which];
  const int in = which == 0 ? d.num_actor_obs : d.num_critic_obs;
  if (!d.recurrent && p->fz) return fused_forward_one(p, which, x, rows, out, st);
  if (!d.recurrent) {
    if (int rc = stage_obs(x, rows, in, nullptr, n.x, p->lo_off, n.kpad[0], n.kpad[0], st)) return rc;
  } else {
    if (!hidden_h || !hidden_c) return fail(URLG_EINVAL, "urlg_policy_forward: recurrent policy needs hidden state");
    if (int rc = lstm_rollout_step(p, which, x, rows, nullptr, hidden_h, hidden_c, nullptr, nullptr, st)) return rc;
  }
  if (int rc = mlp_forward(p, n, rows, st)) return rc;
  URLG_CUDA(cudaMemcpyAsync(out, n.out, (size_t)rows * n.d.dims[n.d.num_layers] * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return URLG_OK;
}
