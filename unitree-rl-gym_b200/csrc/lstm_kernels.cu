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
namespace tc {
int launch_gemm(const urlg_gemm_desc* d, cudaStream_t st);
int launch_gemm_group(const urlg_gemm_desc* const* ds, int n, cudaStream_t st);
}
// lstm_tc.cu: the recurrence on tcgen05 (cluster size, or 0 when the Memory is outside its limits / URLG_LSTM_TC=0)
int lstm_tc_clusters(const urlg_policy* p, const LstmWs& w);
int lstm_tc_forward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int E, const float* saved_h,
                    const float* saved_c, cudaStream_t st);
int lstm_tc_backward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int E, cudaStream_t st);

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
                          bf16* __restrict__ dgates, int64_t lo_off, long long* __restrict__ g_b_ih, long long* __restrict__ g_b_hh) {
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
    if (a != 0.0f) { tc::fx_add(g_b_ih + k, a, tc::kFxScale); tc::fx_add(g_b_hh + k, a, tc::kFxScale); }
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

// W_hh [4H, H] (nn.LSTM layout) -> [k][j][gate] fp32 image for the forward recurrence.
__global__ void __launch_bounds__(256)
lstm_wt_image_kernel(const float* __restrict__ W_hh, int H, float* __restrict__ Wt) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 4 * H * H) return;
  const int q = idx / H, k = idx % H, gate = q / H, jj = q % H;
  Wt[((size_t)k * H + jj) * 4 + gate] = W_hh[idx];
}

// Forward recurrence.  Thread (rg, j): hidden unit j of RPT consecutive rows (row group rg of the CTA); the W_hh
// column it needs is read from shared memory once per step and reused for its RPT rows.  gates: in = x W_ih^T + b
// (pre-activation), out = activated (i, f, g, o) kept for the backward pass.  The next step's inputs are prefetched
// into registers while the current step computes.
template <int RPT>
__global__ void __launch_bounds__(256)
lstm_recur_fwd_kernel(float* __restrict__ gates, int T, int E, int H, int N, int env_begin,
                      const uint8_t* __restrict__ dones, const float* __restrict__ saved_h, const float* __restrict__ saved_c,
                      const float* __restrict__ W_hh, float* __restrict__ c_prev, float* __restrict__ c_new,
                      bf16* __restrict__ xcat, int kcat, int in_pad, bf16* __restrict__ headx, int ld_headx, int64_t lo_off) {
  extern __shared__ __align__(16) float lsm[];
  float4* Wt = reinterpret_cast<float4*>(lsm);          // [k][j] -> (W[j][k], W[H+j][k], W[2H+j][k], W[3H+j][k])
  float* hs = lsm + 4 * H * H;                          // [k][rows_per_cta]: a thread's RPT rows are contiguous
  // W_hh arrives already in the [k][j][gate] layout (lstm_wt_image_kernel): coalesced 16-byte loads
  for (int idx = threadIdx.x; idx < H * H; idx += blockDim.x) Wt[idx] = reinterpret_cast<const float4*>(W_hh)[idx];
  const int RG = blockDim.x / H, rows_cta = RG * RPT;
  const int rg = threadIdx.x / H, j = threadIdx.x % H;
  const int e0 = blockIdx.x * rows_cta + rg * RPT;
  float h[RPT], c[RPT];
  float4 pre[RPT];                                      // prefetched projection of the current step
  bool seed[RPT];
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    h[i] = 0.0f; c[i] = 0.0f; seed[i] = true;
    pre[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (e0 + i < E) {
      const float* g = gates + ((size_t)e0 + i) * 4 * H;
      pre[i] = make_float4(g[j], g[H + j], g[2 * H + j], g[3 * H + j]);
    }
  }
  __syncthreads();
  for (int t = 0; t < T; ++t) {
    float4 acc[RPT];
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      const int e = e0 + i;
      acc[i] = pre[i];
      if (e < E) {
        const size_t row = (size_t)t * E + e;
        if (seed[i]) {                                    // trajectory start: re-seed from the saved states
          const size_t sv = ((size_t)t * N + env_begin + e) * H + j;
          h[i] = saved_h[sv];
          c[i] = saved_c[sv];
        }
        store_split(xcat, lo_off, row * kcat + in_pad + j, h[i]);    // h_prev operand image for dW_hh
        c_prev[row * H + j] = c[i];
      }
      hs[j * rows_cta + rg * RPT + i] = h[i];
    }
    __syncthreads();
    // prefetch step t+1 (projection + boundary flag) while this step's dot products run
    if (t + 1 < T) {
#pragma unroll
      for (int i = 0; i < RPT; ++i) {
        const int e = e0 + i;
        if (e < E) {
          const float* g = gates + ((size_t)(t + 1) * E + e) * 4 * H;
          pre[i] = make_float4(g[j], g[H + j], g[2 * H + j], g[3 * H + j]);
          seed[i] = dones[(size_t)t * N + env_begin + e] != 0;
        }
      }
    }
    const float* hk = hs + rg * RPT;
#pragma unroll 4
    for (int k = 0; k < H; ++k) {
      const float4 w = Wt[k * H + j];
      float hv[RPT];
      if (RPT == 4) {
        const float4 q = *reinterpret_cast<const float4*>(hk + k * rows_cta);
        hv[0] = q.x; hv[1 % RPT] = q.y; hv[2 % RPT] = q.z; hv[3 % RPT] = q.w;
      } else if (RPT == 2) {
        const float2 q = *reinterpret_cast<const float2*>(hk + k * rows_cta);
        hv[0] = q.x; hv[1 % RPT] = q.y;
      } else {
        hv[0] = hk[k * rows_cta];
      }
#pragma unroll
      for (int i = 0; i < RPT; ++i) {
        acc[i].x += hv[i] * w.x; acc[i].y += hv[i] * w.y; acc[i].z += hv[i] * w.z; acc[i].w += hv[i] * w.w;
      }
    }
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      const int e = e0 + i;
      if (e < E) {
        const size_t row = (size_t)t * E + e;
        const float ig = sigmoidf_(acc[i].x), fg = sigmoidf_(acc[i].y), gg = tanhf(acc[i].z), og = sigmoidf_(acc[i].w);
        c[i] = fg * c[i] + ig * gg;
        h[i] = og * tanhf(c[i]);
        float* g = gates + row * 4 * H;
        g[j] = ig; g[H + j] = fg; g[2 * H + j] = gg; g[3 * H + j] = og;
        c_new[row * H + j] = c[i];
        store_split(headx, lo_off, row * ld_headx + j, h[i]);
      }
    }
    __syncthreads();
  }
}

// Backward recurrence (BPTT).  Thread (rg, j) carries dL/dc and the recurrent dL/dh of unit j for its RPT rows.
template <int RPT>
__global__ void __launch_bounds__(256)
lstm_recur_bwd_kernel(const float* __restrict__ gates, int T, int E, int H, int N, int env_begin,
                      const uint8_t* __restrict__ dones, const float* __restrict__ c_prev, const float* __restrict__ c_new,
                      const bf16* __restrict__ dh_head, bf16* __restrict__ dgates, int64_t lo_off,
                      const float* __restrict__ W_hh, long long* __restrict__ g_b_ih, long long* __restrict__ g_b_hh) {
  extern __shared__ __align__(16) float lsm[];
  float* Wb = lsm;                                      // [4H][H] as stored by nn.LSTM
  float* dg = lsm + 4 * H * H;                          // [4H][rows_per_cta]
  for (int idx = threadIdx.x; idx < 4 * H * H; idx += blockDim.x) Wb[idx] = W_hh[idx];
  const int RG = blockDim.x / H, rows_cta = RG * RPT;
  const int rg = threadIdx.x / H, j = threadIdx.x % H;
  const int e0 = blockIdx.x * rows_cta + rg * RPT;
  float dh_rec[RPT], dc[RPT], bsum[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
  for (int i = 0; i < RPT; ++i) { dh_rec[i] = 0.0f; dc[i] = 0.0f; }
  __syncthreads();
  for (int t = T - 1; t >= 0; --t) {
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      const int e = e0 + i;
      float pi = 0.0f, pf = 0.0f, pg = 0.0f, po = 0.0f;
      if (e < E) {
        const size_t row = (size_t)t * E + e, idx = row * H + j;
        const bool carry = t < T - 1 && dones[(size_t)t * N + env_begin + e] == 0;   // a done at t cuts the gradient from t+1
        const float dh = load_split(dh_head, lo_off, idx) + (carry ? dh_rec[i] : 0.0f);
        const float dcn_in = carry ? dc[i] : 0.0f;
        const float* g = gates + row * 4 * H;
        const float ig = g[j], fg = g[H + j], gg = g[2 * H + j], og = g[3 * H + j];
        const float tc = tanhf(c_new[idx]);
        const float d_o = dh * tc;
        const float dcn = dcn_in + dh * og * (1.0f - tc * tc);
        const float di = dcn * gg, dgg = dcn * ig, df = dcn * c_prev[idx];
        dc[i] = dcn * fg;
        pi = di * ig * (1.0f - ig); pf = df * fg * (1.0f - fg); pg = dgg * (1.0f - gg * gg); po = d_o * og * (1.0f - og);
        bf16* dgr = dgates + row * 4 * H;
        store_split(dgr, lo_off, j, pi); store_split(dgr, lo_off, H + j, pf);
        store_split(dgr, lo_off, 2 * H + j, pg); store_split(dgr, lo_off, 3 * H + j, po);
        bsum[0] += pi; bsum[1] += pf; bsum[2] += pg; bsum[3] += po;
      }
      const int col = rg * RPT + i;
      dg[j * rows_cta + col] = pi; dg[(H + j) * rows_cta + col] = pf;
      dg[(2 * H + j) * rows_cta + col] = pg; dg[(3 * H + j) * rows_cta + col] = po;
    }
    __syncthreads();
    if (t > 0) {                                          // dL/dh_{t-1} (recurrent part) = dgates_t W_hh
      float a[RPT];
#pragma unroll
      for (int i = 0; i < RPT; ++i) a[i] = 0.0f;
      const float* dq = dg + rg * RPT;
#pragma unroll 4
      for (int q = 0; q < 4 * H; ++q) {
        const float w = Wb[q * H + j];
        if (RPT == 4) {
          const float4 d = *reinterpret_cast<const float4*>(dq + q * rows_cta);
          a[0] += d.x * w; a[1 % RPT] += d.y * w; a[2 % RPT] += d.z * w; a[3 % RPT] += d.w * w;
        } else if (RPT == 2) {
          const float2 d = *reinterpret_cast<const float2*>(dq + q * rows_cta);
          a[0] += d.x * w; a[1 % RPT] += d.y * w;
        } else {
          a[0] += dq[q * rows_cta] * w;
        }
      }
#pragma unroll
      for (int i = 0; i < RPT; ++i) dh_rec[i] = a[i];
    }
    __syncthreads();
  }
  // bias gradients: per-thread sums over (t, own rows) -> per-CTA sums over row groups -> one atomic per gate unit
  dg[j * RG + rg] = bsum[0]; dg[(H + j) * RG + rg] = bsum[1]; dg[(2 * H + j) * RG + rg] = bsum[2]; dg[(3 * H + j) * RG + rg] = bsum[3];
  __syncthreads();
  for (int q = threadIdx.x; q < 4 * H; q += blockDim.x) {
    float a = 0.0f;
    for (int rr = 0; rr < RG; ++rr) a += dg[q * RG + rr];
    if (a != 0.0f) { tc::fx_add(g_b_ih + q, a, tc::kFxScale); tc::fx_add(g_b_hh + q, a, tc::kFxScale); }
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
// Rows per thread: W_hh traffic from shared memory is per thread and per step, so more rows per thread amortise it;
// keep at least ~one CTA per SM.
static int rows_per_thread(int E, int rpc) {
  if (E >= 4 * rpc * kNumSMs) return 4;
  if (E >= 2 * rpc * kNumSMs) return 2;
  return 1;
}
static int recur_smem_attr(int smem) {
  static int set_to = 0;
  if (smem > set_to) {
    const int cap = 200 * 1024;
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_fwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap));
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_fwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap));
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_fwd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap));
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap));
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_bwd_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap));
    URLG_CUDA(cudaFuncSetAttribute(lstm_recur_bwd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap));
    set_to = cap;
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
    // tensor-core recurrence: input projection, recurrent product and cell update of all T steps in one cluster kernel
    if (lstm_tc_clusters(p, w) > 0) return lstm_tc_forward(p, k, s, env_begin, E, sh, sc, st);
    urlg_gemm_desc g = {};
    g.A = w.xcat; g.lda = w.kcat;
    g.B = w.w; g.ldb = w.kcat;
    g.M = T * E; g.N = 4 * H; g.K = w.in_pad;
    g.bias = w.bias;
    g.out_f32 = w.gates; g.ld_f32 = 4 * H;
    with_lo(p, g);
    if (int rc = tc::launch_gemm(&g, st)) return rc;
    const int rpt = rows_per_thread(E, rpc);
    const int rows_cta = rpc * rpt;
    const int smem = (4 * H * H + rows_cta * H) * (int)sizeof(float);
    if (int rc = recur_smem_attr(smem)) return rc;
    const int grid = (E + rows_cta - 1) / rows_cta;
    float* wt_img = w.dhrec;                   // [R, H] scratch of the per-step path, unused here (R*H >= 4H*H)
    if ((int64_t)p->R * H < 4 * H * H) return fail(URLG_EINVAL, "lstm update: max_rows too small for the W_hh image");
    lstm_wt_image_kernel<<<(4 * H * H + 255) / 256, 256, 0, st>>>(p->params + w.off[1], H, wt_img);
    URLG_LAUNCH_CHECK();
#define URLG_FWD(RPT)                                                                                                  \
    lstm_recur_fwd_kernel<RPT><<<grid, rpc * H, smem, st>>>(w.gates, T, E, H, s->N, env_begin, s->dones, sh, sc,        \
                                                           wt_img, c_prev, c_new, w.xcat, w.kcat,                       \
                                                           w.in_pad, n.x, n.kpad[0], p->lo_off)
    if (rpt == 4) URLG_FWD(4); else if (rpt == 2) URLG_FWD(2); else URLG_FWD(1);
#undef URLG_FWD
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
  const bool on_tc = rpc > 0 && lstm_tc_clusters(p, w) > 0;
  if (on_tc) {
    if (int rc = lstm_tc_backward(p, k, s, env_begin, E, st)) return rc;
  } else if (rpc > 0) {
    const int rpt = rows_per_thread(E, rpc);
    const int rows_cta = rpc * rpt;
    const int smem = (4 * H * H + rows_cta * 4 * H) * (int)sizeof(float);
    if (int rc = recur_smem_attr(smem)) return rc;
    const int grid = (E + rows_cta - 1) / rows_cta;
#define URLG_BWD(RPT)                                                                                                  \
    lstm_recur_bwd_kernel<RPT><<<grid, rpc * H, smem, st>>>(w.gates, T, E, H, s->N, env_begin, s->dones, c_prev, c_new, \
                                                           w.dx_head, w.dgates, p->lo_off, p->params + w.off[1],        \
                                                           p->gacc + w.off[2], p->gacc + w.off[3])
    if (rpt == 4) URLG_BWD(4); else if (rpt == 2) URLG_BWD(2); else URLG_BWD(1);
#undef URLG_BWD
    URLG_LAUNCH_CHECK();
  }
  for (int t = rpc > 0 ? -1 : T - 1; t >= 0; --t) {
    const size_t ro = (size_t)t * E;
    lstm_cell_backward_kernel<<<(total + 255) / 256, 256, 4 * H * sizeof(float), st>>>(
        w.gates + ro * 4 * H, E, H, c_prev + ro * H, c_new + ro * H, w.dx_head + ro * H, w.dhrec, w.dc,
        s->dones + (size_t)t * s->N + env_begin, t < T - 1 ? 1 : 0, w.dgates + ro * 4 * H, p->lo_off, p->gacc + w.off[2],
        p->gacc + w.off[3]);
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
  // (both parts in ONE launch: they read the same dgates and write disjoint gradient slices)
  urlg_gemm_desc gs[2];
  const urlg_gemm_desc* ptr[2] = {&gs[0], &gs[1]};
  for (int part = 0; part < 2; ++part) {
    urlg_gemm_desc& g = gs[part];
    g = urlg_gemm_desc{};
    g.A = w.dgates; g.lda = 4 * H; g.a_mn = 1;
    g.B = part == 0 ? w.xcat : w.xcat + w.in_pad; g.ldb = w.kcat; g.b_mn = 1;
    g.M = 4 * H; g.N = part == 0 ? w.in : H; g.K = T * E;
    g.block_n = round_up(g.N, 64);
    g.out_f32 = p->grads + w.off[part]; g.ld_f32 = g.N;
    g.out_fx = reinterpret_cast<int64_t*>(p->gacc + w.off[part]);
    g.flags = URLG_GEMM_ATOMIC;
    const int tiles = 2 * ((4 * H + 127) / 128);          // both parts share the wave
    int splits = kNumSMs / tiles;
    const int kblocks = (T * E + 63) / 64;
    if (splits > kblocks / 4) splits = kblocks / 4;
    g.k_splits = splits < 1 ? 1 : splits;
    with_lo(p, g);
  }
  if (int rc = tc::launch_gemm_group(ptr, 2, st)) return rc;
  return URLG_OK;
}

}  // namespace urlg
