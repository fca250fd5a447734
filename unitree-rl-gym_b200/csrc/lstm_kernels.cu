// ActorCriticRecurrent's Memory = nn.LSTM(in, H, 1 layer) ([UPSTREAM rsl_rl v1.0.2]
// modules/actor_critic_recurrent.py; gate order i,f,g,o and weight layout pinned by
// deploy/pre_train/{g1,h1,h1_2}/motion.pt; kernels K3 (rollout) and K8 (update)).
//   gates = [x | h_prev] * [W_ih | W_hh]^T + (b_ih + b_hh)   one tcgen05 GEMM per step, K = in_pad + H
//   cell  : i,f,o = sigmoid, g = tanh, c' = f c + i g, h' = o tanh(c')
// Update mode replaces split_and_pad_trajectories/unpad by an equivalent masked time loop over the env
// slice: at a trajectory start (t == 0 or done at t-1) the state is re-seeded from the saved hidden states
// and no gradient flows across that boundary — no padding, no host sync.
#include "policy.cuh"

namespace urlg {
namespace tc { int launch_gemm(const urlg_gemm_desc* d, cudaStream_t st); }

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// ---------------------------------------------------------------- rollout: stage [x | h] (+ save pre-step state)
__global__ void __launch_bounds__(256)
lstm_stage_rollout_kernel(const float* __restrict__ x, int64_t rows, int in, int in_pad, int H, float* __restrict__ store_x,
                          const float* __restrict__ h, const float* __restrict__ c, float* __restrict__ save_h,
                          float* __restrict__ save_c, bf16* __restrict__ xcat, int64_t lo_off) {
  const int kcat = in_pad + H;
  const int64_t total = rows * kcat;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / kcat;
    const int col = (int)(i % kcat);
    float v = 0.0f;
    if (col < in) {
      v = x[r * in + col];
      if (store_x) store_x[r * in + col] = v;
    } else if (col >= in_pad) {
      const int j = col - in_pad;
      v = h[r * H + j];
      if (save_h) { save_h[r * H + j] = v; save_c[r * H + j] = c[r * H + j]; }
    }
    store_split(xcat, lo_off, (size_t)i, v);
  }
}

// ---------------------------------------------------------------- update: stage step t of the env slice
__global__ void __launch_bounds__(256)
lstm_stage_update_kernel(const float* __restrict__ obs /* [T,N,in] */, int in, int in_pad, int H, int t, int N, int env_begin,
                         int env_count, const uint8_t* __restrict__ dones, const float* __restrict__ saved_h,
                         const float* __restrict__ saved_c, bf16* __restrict__ xcat /* rows of step t */,
                         const bf16* __restrict__ headx_prev /* h' of step t-1 [env_count,H] */,
                         float* __restrict__ c_prev /* rows of step t */, const float* __restrict__ c_new_prev /* step t-1 */,
                         int64_t lo_off) {
  const int kcat = in_pad + H;
  const int64_t total = (int64_t)env_count * kcat;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i / kcat), col = (int)(i % kcat);
    const int env = env_begin + e;
    float v = 0.0f;
    if (col < in) {
      v = obs[((size_t)t * N + env) * in + col];
    } else if (col >= in_pad) {
      const int j = col - in_pad;
      const bool reseed = t == 0 || dones[(size_t)(t - 1) * N + env] != 0;
      const size_t sv = ((size_t)t * N + env) * H + j;
      if (reseed) {
        v = saved_h[sv];
        c_prev[(size_t)e * H + j] = saved_c[sv];
      } else {
        v = load_split(headx_prev, lo_off, (size_t)e * H + j);
        c_prev[(size_t)e * H + j] = c_new_prev[(size_t)e * H + j];
      }
    }
    store_split(xcat, lo_off, (size_t)i, v);
  }
}

// ---------------------------------------------------------------- cell forward
// gates: pre-activations in, activated (i,f,g,o) out (kept for the backward pass).
__global__ void __launch_bounds__(256)
lstm_cell_kernel(float* __restrict__ gates, int64_t rows, int H, const float* c_prev, float* c_new,
                 float* __restrict__ h_out, bf16* __restrict__ headx, int ld_headx, int64_t lo_off) {
  const int64_t total = rows * H;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = idx / H;
    const int j = (int)(idx % H);
    float* g = gates + r * 4 * H;
    const float i = sigmoidf_(g[j]), f = sigmoidf_(g[H + j]), gg = tanhf(g[2 * H + j]), o = sigmoidf_(g[3 * H + j]);
    const float c = f * c_prev[idx] + i * gg;
    const float h = o * tanhf(c);
    g[j] = i; g[H + j] = f; g[2 * H + j] = gg; g[3 * H + j] = o;
    c_new[idx] = c;
    if (h_out) h_out[idx] = h;
    store_split(headx, lo_off, (size_t)(r * ld_headx + j), h);
  }
}

// ---------------------------------------------------------------- cell backward (step t)
__global__ void __launch_bounds__(256)
lstm_cell_backward_kernel(const float* __restrict__ gates, int rows, int H, const float* __restrict__ c_prev,
                          const float* __restrict__ c_new, const bf16* __restrict__ dh_head, const float* __restrict__ dh_rec,
                          float* __restrict__ dc /* in: dL/dc_new from t+1, out: dL/dc_prev */,
                          const uint8_t* __restrict__ cut /* dones[t] of the slice: 1 = no gradient from t+1 */, int has_next,
                          bf16* __restrict__ dgates, int64_t lo_off, float* __restrict__ g_b_ih, float* __restrict__ g_b_hh) {
  extern __shared__ float acc[];   // [4H]
  for (int k = threadIdx.x; k < 4 * H; k += blockDim.x) acc[k] = 0.0f;
  __syncthreads();
  const int total = rows * H;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < total) {
    const int r = idx / H, j = idx % H;
    const bool carry = has_next && !cut[r];
    const float dh = load_split(dh_head, lo_off, (size_t)idx) + (carry ? dh_rec[idx] : 0.0f);
    const float dcn_in = carry ? dc[idx] : 0.0f;
    const float* g = gates + (size_t)r * 4 * H;
    const float i = g[j], f = g[H + j], gg = g[2 * H + j], o = g[3 * H + j];
    const float tc = tanhf(c_new[idx]);
    const float d_o = dh * tc;
    const float dcn = dcn_in + dh * o * (1.0f - tc * tc);
    const float di = dcn * gg, dg = dcn * i, df = dcn * c_prev[idx];
    dc[idx] = dcn * f;
    const float pi = di * i * (1.0f - i), pf = df * f * (1.0f - f), pg = dg * (1.0f - gg * gg), po = d_o * o * (1.0f - o);
    bf16* dgr = dgates + (size_t)r * 4 * H;
    store_split(dgr, lo_off, j, pi); store_split(dgr, lo_off, H + j, pf);
    store_split(dgr, lo_off, 2 * H + j, pg); store_split(dgr, lo_off, 3 * H + j, po);
    atomicAdd(acc + j, pi); atomicAdd(acc + H + j, pf); atomicAdd(acc + 2 * H + j, pg); atomicAdd(acc + 3 * H + j, po);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < 4 * H; k += blockDim.x) {
    const float a = acc[k];
    if (a != 0.0f) { atomicAdd(g_b_ih + k, a); atomicAdd(g_b_hh + k, a); }
  }
}

// ================================================================ update mode, persistent form (K8)
// The 24-step recurrence of a minibatch is ONE kernel per direction: env rows are independent, so a CTA owns a few
// rows for the whole sequence, keeps W_hh (fp32, 4H x H = 64 KB at H = 64) in shared memory and loops over t with the
// state (h, c / dh, dc) in registers — no per-step launches, no grid sync.  The input projection x_t W_ih^T + b for
// all T*E rows is one tensor-core GEMM before it, the weight gradients two split-K GEMMs after it.  The recurrent
// product itself runs in fp32 FMAs (exact; K = H is far too small for the tensor pipe to matter).

// x part of [x | h_prev] for every row of the slice at once (row = t*E + e); the h_prev part is written by the recurrence.
__global__ void __launch_bounds__(256)
lstm_stage_x_kernel(const float* __restrict__ obs /* [T,N,in] */, int in, int in_pad, int kcat, int T, int N, int env_begin,
                    int E, bf16* __restrict__ xcat, int64_t lo_off) {
  const int64_t total = (int64_t)T * E * in_pad;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / in_pad;
    const int col = (int)(i % in_pad);
    const int t = (int)(row / E), e = (int)(row % E);
    const float v = col < in ? obs[((size_t)t * N + env_begin + e) * in + col] : 0.0f;
    store_split(xcat, lo_off, (size_t)row * kcat + col, v);
  }
}

// Forward recurrence.  Thread (r, j): row r of the CTA, hidden unit j.  gates: in = x W_ih^T + b (pre-activation),
// out = activated (i, f, g, o) kept for the backward pass.
__global__ void __launch_bounds__(256)
lstm_recur_fwd_kernel(float* __restrict__ gates, int T, int E, int H, int N, int env_begin,
                      const uint8_t* __restrict__ dones, const float* __restrict__ saved_h, const float* __restrict__ saved_c,
                      const float* __restrict__ W_hh, float* __restrict__ c_prev, float* __restrict__ c_new,
                      bf16* __restrict__ xcat, int kcat, int in_pad, bf16* __restrict__ headx, int ld_headx, int64_t lo_off,
                      int rows_per_cta) {
  extern __shared__ __align__(16) float lsm[];
  float4* Wt = reinterpret_cast<float4*>(lsm);          // [k][j] -> (W[j][k], W[H+j][k], W[2H+j][k], W[3H+j][k])
  float* hs = lsm + 4 * H * H;                          // [rows_per_cta][H]
  for (int idx = threadIdx.x; idx < H * H; idx += blockDim.x) {
    const int k = idx / H, jj = idx % H;
    Wt[idx] = make_float4(W_hh[(size_t)jj * H + k], W_hh[(size_t)(H + jj) * H + k], W_hh[(size_t)(2 * H + jj) * H + k],
                          W_hh[(size_t)(3 * H + jj) * H + k]);
  }
  const int r = threadIdx.x / H, j = threadIdx.x % H;
  const int e = blockIdx.x * rows_per_cta + r;
  const bool valid = r < rows_per_cta && e < E;
  const int env = env_begin + (valid ? e : 0);
  float h = 0.0f, c = 0.0f;
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    const size_t row = (size_t)t * E + e;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (valid) {
      const float* g = gates + row * 4 * H;               // issue the projection loads first: latency hides under the dot
      acc = make_float4(g[j], g[H + j], g[2 * H + j], g[3 * H + j]);
      if (t == 0 || dones[(size_t)(t - 1) * N + env] != 0) {   // trajectory start: re-seed from the saved states
        const size_t sv = ((size_t)t * N + env) * H + j;
        h = saved_h[sv];
        c = saved_c[sv];
      }
      hs[r * H + j] = h;
      store_split(xcat, lo_off, row * kcat + in_pad + j, h);    // h_prev operand image for dW_hh
      c_prev[row * H + j] = c;
    }
    __syncthreads();
    if (valid) {
      const float* hr = hs + r * H;
#pragma unroll 8
      for (int k = 0; k < H; ++k) {
        const float hv = hr[k];
        const float4 w = Wt[k * H + j];
        acc.x += hv * w.x; acc.y += hv * w.y; acc.z += hv * w.z; acc.w += hv * w.w;
      }
      const float i = sigmoidf_(acc.x), f = sigmoidf_(acc.y), gg = tanhf(acc.z), o = sigmoidf_(acc.w);
      c = f * c + i * gg;
      h = o * tanhf(c);
      float* g = gates + row * 4 * H;
      g[j] = i; g[H + j] = f; g[2 * H + j] = gg; g[3 * H + j] = o;
      c_new[row * H + j] = c;
      store_split(headx, lo_off, row * ld_headx + j, h);
    }
    __syncthreads();
  }
}

// Backward recurrence (BPTT).  Thread (r, j) carries dL/dc and the recurrent dL/dh of its unit in registers.
__global__ void __launch_bounds__(256)
lstm_recur_bwd_kernel(const float* __restrict__ gates, int T, int E, int H, int N, int env_begin,
                      const uint8_t* __restrict__ dones, const float* __restrict__ c_prev, const float* __restrict__ c_new,
                      const bf16* __restrict__ dh_head, bf16* __restrict__ dgates, int64_t lo_off,
                      const float* __restrict__ W_hh, float* __restrict__ g_b_ih, float* __restrict__ g_b_hh,
                      int rows_per_cta) {
  extern __shared__ __align__(16) float lsm[];
  float* Wb = lsm;                                      // [4H][H] as stored by nn.LSTM
  float* dg = lsm + 4 * H * H;                          // [rows_per_cta][4H]
  for (int idx = threadIdx.x; idx < 4 * H * H; idx += blockDim.x) Wb[idx] = W_hh[idx];
  const int r = threadIdx.x / H, j = threadIdx.x % H;
  const int e = blockIdx.x * rows_per_cta + r;
  const bool valid = r < rows_per_cta && e < E;
  const int env = env_begin + (valid ? e : 0);
  float dh_rec = 0.0f, dc = 0.0f;
  float bi = 0.0f, bf = 0.0f, bg = 0.0f, bo = 0.0f;
  __syncthreads();
  for (int t = T - 1; t >= 0; --t) {
    const size_t row = (size_t)t * E + e;
    if (valid) {
      const size_t idx = row * H + j;
      const bool carry = t < T - 1 && dones[(size_t)t * N + env] == 0;   // a done at t cuts the gradient from t+1
      const float dh = load_split(dh_head, lo_off, idx) + (carry ? dh_rec : 0.0f);
      const float dcn_in = carry ? dc : 0.0f;
      const float* g = gates + row * 4 * H;
      const float i = g[j], f = g[H + j], gg = g[2 * H + j], o = g[3 * H + j];
      const float tc = tanhf(c_new[idx]);
      const float d_o = dh * tc;
      const float dcn = dcn_in + dh * o * (1.0f - tc * tc);
      const float di = dcn * gg, dgg = dcn * i, df = dcn * c_prev[idx];
      dc = dcn * f;
      const float pi = di * i * (1.0f - i), pf = df * f * (1.0f - f), pg = dgg * (1.0f - gg * gg), po = d_o * o * (1.0f - o);
      bf16* dgr = dgates + row * 4 * H;
      store_split(dgr, lo_off, j, pi); store_split(dgr, lo_off, H + j, pf);
      store_split(dgr, lo_off, 2 * H + j, pg); store_split(dgr, lo_off, 3 * H + j, po);
      float* d4 = dg + r * 4 * H;
      d4[j] = pi; d4[H + j] = pf; d4[2 * H + j] = pg; d4[3 * H + j] = po;
      bi += pi; bf += pf; bg += pg; bo += po;
    }
    __syncthreads();
    if (valid && t > 0) {                                 // dL/dh_{t-1} (recurrent part) = dgates_t W_hh
      const float* d4 = dg + r * 4 * H;
      float a = 0.0f;
#pragma unroll 8
      for (int q = 0; q < 4 * H; ++q) a += d4[q] * Wb[q * H + j];
      dh_rec = a;
    }
    __syncthreads();
  }
  // bias gradients: sum over the CTA's rows, then one atomic per gate unit (b_ih and b_hh get the same gradient)
  if (valid) {
    float* d4 = dg + r * 4 * H;
    d4[j] = bi; d4[H + j] = bf; d4[2 * H + j] = bg; d4[3 * H + j] = bo;
  }
  __syncthreads();
  const int live_rows = min(rows_per_cta, E - blockIdx.x * rows_per_cta);
  for (int q = threadIdx.x; q < 4 * H; q += blockDim.x) {
    float a = 0.0f;
    for (int rr = 0; rr < live_rows; ++rr) a += dg[rr * 4 * H + q];
    if (a != 0.0f) { atomicAdd(g_b_ih + q, a); atomicAdd(g_b_hh + q, a); }
  }
}

static inline int blocks_for(int64_t total, int cap = kNumSMs * 16) {
  int b = (int)((total + 255) / 256);
  return b > cap ? cap : (b < 1 ? 1 : b);
}

// Rows per CTA of the persistent recurrence (threads = rows * H <= 256), or 0 when W_hh (4H*H fp32) does not fit.
static int persistent_rows(int H) {
  if (H < 8 || H > 104 || (H & 3)) return 0;
  const int r = 256 / H;
  return r < 1 ? 0 : (r > 8 ? 8 : r);
}
static int recur_smem_attr(int smem) {
  static int set_to = 0;
  if (smem > set_to) {
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    set_to = smem;
  }
  return URLG_OK;
}

static int gates_gemm(urlg_policy* p, LstmWs& w, const bf16* xcat, float* gates, int64_t rows, cudaStream_t st) {
  urlg_gemm_desc g = {};
  g.A = xcat; g.lda = w.kcat;
  g.B = w.w; g.ldb = w.kcat;
  g.M = (int)rows; g.N = 4 * w.H; g.K = w.kcat;
  g.bias = w.bias;
  g.out_f32 = gates; g.ld_f32 = 4 * w.H;
  with_lo(p, g);
  return tc::launch_gemm(&g, st);
}

int lstm_rollout_step(urlg_policy* p, int k, const float* x, int64_t rows, float* store_x, float* h, float* c,
                      float* save_h, float* save_c, cudaStream_t st) {
  LstmWs& w = p->lstm[k];
  NetWs& n = p->net[k];
  lstm_stage_rollout_kernel<<<blocks_for(rows * w.kcat), 256, 0, st>>>(x, rows, w.in, w.in_pad, w.H, store_x, h, c, save_h,
                                                                       save_c, w.xcat, p->lo_off);
  URLG_LAUNCH_CHECK();
  if (int rc = gates_gemm(p, w, w.xcat, w.gates, rows, st)) return rc;
  lstm_cell_kernel<<<blocks_for(rows * w.H), 256, 0, st>>>(w.gates, rows, w.H, c, c, h, n.x, n.kpad[0], p->lo_off);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

static void slice_ptrs(const urlg_storage* s, int k, const float*& obs, int& in, const float*& sh, const float*& sc) {
  if (k == 0 || !s->privileged_observations) { obs = s->observations; in = s->num_obs; }
  else { obs = s->privileged_observations; in = s->num_priv; }
  sh = k == 0 ? s->hid_a_h : s->hid_c_h;
  sc = k == 0 ? s->hid_a_c : s->hid_c_c;
}

int lstm_update_forward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int env_count, cudaStream_t st) {
  LstmWs& w = p->lstm[k];
  NetWs& n = p->net[k];
  const int T = s->T, H = w.H, E = env_count;
  const float *obs, *sh, *sc; int in;
  slice_ptrs(s, k, obs, in, sh, sc);
  if (in != w.in) return fail(URLG_EINVAL, "lstm update: observation width %d != %d", in, w.in);
  if (!sh || !sc) return fail(URLG_EINVAL, "lstm update: storage lacks saved hidden states");
  if ((int64_t)T * E > p->R) return fail(URLG_EINVAL, "lstm update: T*envs exceeds max_rows");
  float* c_prev = w.c;                       // [T*E, H]
  float* c_new = w.c + (size_t)p->R * H;     // [T*E, H]
  const int rpc = persistent_rows(H);
  if (rpc > 0) {
    // persistent form: stage x for all rows, ONE projection GEMM (x W_ih^T + b), ONE recurrence kernel
    lstm_stage_x_kernel<<<blocks_for((int64_t)T * E * w.in_pad), 256, 0, st>>>(obs, in, w.in_pad, w.kcat, T, s->N, env_begin, E,
                                                                             w.xcat, p->lo_off);
    URLG_LAUNCH_CHECK();
    urlg_gemm_desc g = {};
    g.A = w.xcat; g.lda = w.kcat;
    g.B = w.w; g.ldb = w.kcat;
    g.M = T * E; g.N = 4 * H; g.K = w.in_pad;
    g.bias = w.bias;
    g.out_f32 = w.gates; g.ld_f32 = 4 * H;
    with_lo(p, g);
    if (int rc = tc::launch_gemm(&g, st)) return rc;
    const int smem = (4 * H * H + rpc * H) * (int)sizeof(float);
    if (int rc = recur_smem_attr(smem)) return rc;
    lstm_recur_fwd_kernel<<<(E + rpc - 1) / rpc, rpc * H, smem, st>>>(
        w.gates, T, E, H, s->N, env_begin, s->dones, sh, sc, p->params + w.off[1], c_prev, c_new, w.xcat, w.kcat, w.in_pad,
        n.x, n.kpad[0], p->lo_off, rpc);
    URLG_LAUNCH_CHECK();
    return URLG_OK;
  }
  for (int t = 0; t < T; ++t) {
    const size_t ro = (size_t)t * E;
    lstm_stage_update_kernel<<<blocks_for((int64_t)E * w.kcat), 256, 0, st>>>(
        obs, in, w.in_pad, H, t, s->N, env_begin, E, s->dones, sh, sc, w.xcat + ro * w.kcat,
        t ? n.x + (ro - E) * n.kpad[0] : nullptr, c_prev + ro * H, t ? c_new + (ro - E) * H : nullptr, p->lo_off);
    URLG_LAUNCH_CHECK();
    if (int rc = gates_gemm(p, w, w.xcat + ro * w.kcat, w.gates + ro * 4 * H, E, st)) return rc;
    lstm_cell_kernel<<<blocks_for((int64_t)E * H), 256, 0, st>>>(w.gates + ro * 4 * H, E, H, c_prev + ro * H, c_new + ro * H,
                                                                  nullptr, n.x + ro * n.kpad[0], n.kpad[0], p->lo_off);
    URLG_LAUNCH_CHECK();
  }
  return URLG_OK;
}

int lstm_update_backward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int env_count, cudaStream_t st) {
  LstmWs& w = p->lstm[k];
  const int T = s->T, H = w.H, E = env_count;
  float* c_prev = w.c;
  float* c_new = w.c + (size_t)p->R * H;
  const int total = E * H;
  const int rpc = persistent_rows(H);
  if (rpc > 0) {
    const int smem = (4 * H * H + rpc * 4 * H) * (int)sizeof(float);
    if (int rc = recur_smem_attr(smem)) return rc;
    lstm_recur_bwd_kernel<<<(E + rpc - 1) / rpc, rpc * H, smem, st>>>(
        w.gates, T, E, H, s->N, env_begin, s->dones, c_prev, c_new, w.dx_head, w.dgates, p->lo_off, p->params + w.off[1],
        p->grads + w.off[2], p->grads + w.off[3], rpc);
    URLG_LAUNCH_CHECK();
  }
  for (int t = rpc > 0 ? -1 : T - 1; t >= 0; --t) {
    const size_t ro = (size_t)t * E;
    lstm_cell_backward_kernel<<<(total + 255) / 256, 256, 4 * H * sizeof(float), st>>>(
        w.gates + ro * 4 * H, E, H, c_prev + ro * H, c_new + ro * H, w.dx_head + ro * H, w.dhrec, w.dc,
        s->dones + (size_t)t * s->N + env_begin, t < T - 1 ? 1 : 0, w.dgates + ro * 4 * H, p->lo_off, p->grads + w.off[2],
        p->grads + w.off[3]);
    URLG_LAUNCH_CHECK();
    if (t > 0) {   // dL/dh_{t-1} (recurrent part) = dgates_t * W_hh
      urlg_gemm_desc g = {};
      g.A = w.dgates + ro * 4 * H; g.lda = 4 * H;
      g.B = w.w + w.in_pad; g.ldb = w.kcat; g.b_mn = 1;
      g.M = E; g.N = H; g.K = 4 * H;
      g.block_n = round_up(H, 64);
      g.out_f32 = w.dhrec; g.ld_f32 = H;
      with_lo(p, g);
      if (int rc = tc::launch_gemm(&g, st)) return rc;
    }
  }
  // weight gradients over the whole sequence: dW_ih = dgates^T x, dW_hh = dgates^T h_prev (K = T*E)
  for (int part = 0; part < 2; ++part) {
    urlg_gemm_desc g = {};
    g.A = w.dgates; g.lda = 4 * H; g.a_mn = 1;
    g.B = part == 0 ? w.xcat : w.xcat + w.in_pad; g.ldb = w.kcat; g.b_mn = 1;
    g.M = 4 * H; g.N = part == 0 ? w.in : H; g.K = T * E;
    g.block_n = round_up(g.N, 64);
    g.out_f32 = p->grads + w.off[part]; g.ld_f32 = g.N;
    g.flags = URLG_GEMM_ATOMIC;
    const int tiles = (4 * H + 127) / 128;
    int splits = (kNumSMs + tiles - 1) / tiles;
    const int kblocks = (T * E + 63) / 64;
    if (splits > kblocks / 4) splits = kblocks / 4;
    g.k_splits = splits < 1 ? 1 : splits;
    with_lo(p, g);
    if (int rc = tc::launch_gemm(&g, st)) return rc;
  }
  return URLG_OK;
}

}  // namespace urlg
