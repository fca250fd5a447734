// PPO.update, one minibatch ([UPSTREAM rsl_rl v1.0.2] algorithms/ppo.py update, storage mini_batch_generator;
// SURVEY.md §3.4, §8(a).6 R8/R10, kernels K6/K7):
//   gather_obs_kernel   minibatch rows -> bf16 operand slabs (the other 7 tensors are read through the index
//                       inside the loss head, never materialised)
//   ppo_head_kernel     Gaussian log-prob / entropy / KL / clipped surrogate / clipped value loss and their
//                       gradients wrt mu, value and std; last-layer bias gradients; loss sums
//   backward            dgrad + wgrad on the tcgen05 GEMM (MN-major operands, ELU' and bias-gradient
//                       column sums fused in the dgrad epilogue, split-K partial slabs for wgrad summed in split
//                       order by grads_finalize_kernel: bit-reproducible without atomics)
//   apply               grad-norm, device-side adaptive-KL learning rate, clip, Adam + bf16 weight images
#include "policy.cuh"
#include "peer.cuh"

namespace urlg {
namespace tc {
int launch_gemm(const urlg_gemm_desc* d, cudaStream_t st);
int launch_gemm_group(const urlg_gemm_desc* const* ds, int n, cudaStream_t st);
}

int lstm_update_forward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int env_count, cudaStream_t st);
int lstm_update_backward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int env_count, cudaStream_t st);

__device__ __forceinline__ int64_t row_index(const int64_t* __restrict__ indices, int64_t r, int N, int env_begin, int env_count) {
  if (indices) return indices[r];
  return (r / env_count) * (int64_t)N + env_begin + (r % env_count);
}

// ---------------------------------------------------------------- K6: gather
__global__ void __launch_bounds__(256)
gather_obs_kernel(const float* __restrict__ src, int cols, const int64_t* __restrict__ indices, int64_t rows,
                  bf16* __restrict__ dst, bf16* __restrict__ dst2 /* optional second consumer of the same rows */,
                  int64_t lo_off, int kpad) {
  if (((cols | kpad) & 3) == 0 && (lo_off & 3) == 0 && ((uintptr_t)src & 15) == 0 &&
      (((uintptr_t)dst | (uintptr_t)dst2) & 7) == 0 && rows * (kpad >> 2) < 0x7fffffff) {
    // 4 columns per thread: one float4 load, packed bf16x4 stores of the high and the low parts, 32-bit index math
    const int k4 = kpad >> 2, c4n = cols >> 2;
    const int total = (int)rows * k4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
      const int r = i / k4, c = i - r * k4;
      float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      if (c < c4n) v = __ldg(reinterpret_cast<const float4*>(src + indices[r] * cols) + c);
      const __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
      const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
      const __nv_bfloat162 l0 = __floats2bfloat162_rn(v.x - f0.x, v.y - f0.y), l1 = __floats2bfloat162_rn(v.z - f1.x, v.w - f1.y);
      uint2 hi, lo;
      hi.x = *reinterpret_cast<const uint32_t*>(&h0); hi.y = *reinterpret_cast<const uint32_t*>(&h1);
      lo.x = *reinterpret_cast<const uint32_t*>(&l0); lo.y = *reinterpret_cast<const uint32_t*>(&l1);
      const size_t o = (size_t)i * 4;
      *reinterpret_cast<uint2*>(dst + o) = hi;
      if (lo_off) *reinterpret_cast<uint2*>(dst + o + lo_off) = lo;
      if (dst2) {
        *reinterpret_cast<uint2*>(dst2 + o) = hi;
        if (lo_off) *reinterpret_cast<uint2*>(dst2 + o + lo_off) = lo;
      }
    }
    return;
  }
  const int64_t total = rows * kpad;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / kpad;
    const int c = (int)(i % kpad);
    const float v = c < cols ? src[indices[r] * cols + c] : 0.0f;
    store_split(dst, lo_off, (size_t)i, v);
    if (dst2) store_split(dst2, lo_off, (size_t)i, v);
  }
}

// ---------------------------------------------------------------- loss head
// One thread per (row, action), 32 rows per block: ~60 warps per SM and a few hundred SASS instructions.  (The first
// version ran one thread per row with the action loops unrolled 32-wide: 7.8k SASS instructions executed once per warp
// at 5 warps per SM -- ncu: 8.6 % issue-active, `no_instruction` the top stall, 25 us per launch.)
//   phase 1  per (row, j): log-prob and KL terms -> shared memory
//   leader   (j == 0): sums them in action order (bit-identical to a per-row loop), ratio, surrogate, value loss,
//            d loss / d log-prob and d loss / d value of the row
//   phase 2  per (row, j): d loss / d mu_j, d loss / d sigma_j; column sums over the block's rows -> atomics
constexpr int kHeadRows = 32;
constexpr int kMaxA = 32;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

__global__ void __launch_bounds__(kHeadRows * kMaxA)
ppo_head_kernel(const float* __restrict__ mu_new, const float* __restrict__ val_new, const float* __restrict__ std_param,
                const urlg_storage s, const int64_t* __restrict__ indices, int64_t rows, int env_begin, int env_count,
                const urlg_ppo_hparams hp, bf16* __restrict__ dmu, int ld_dmu, bf16* __restrict__ dval, int ld_dval,
                int64_t lo_off,
                long long* __restrict__ g_std, long long* __restrict__ g_bias_mu, long long* __restrict__ g_bias_val,
                long long* __restrict__ sums /* [4] fixed point: KL, surrogate, value loss, entropy */) {
  extern __shared__ float sh[];
  const int A = s.num_actions;
  float* t_a = sh;                         // [rows][A]  log-prob terms, then d loss / d mu
  float* t_b = t_a + kHeadRows * A;        // [rows][A]  KL terms, then d loss / d sigma
  float* row_dlp = t_b + kHeadRows * A;    // [rows]
  float* row_s = row_dlp + kHeadRows;      // [rows][5]  kl, surrogate, value loss, entropy, d loss / d value
  float* c_log = row_s + 5 * kHeadRows;    // [A]        log sigma_j
  const int tid = threadIdx.x;
  const int e = tid / A, j = tid - e * A;
  const int64_t r = blockIdx.x * (int64_t)kHeadRows + e;
  const bool ok = r < rows;
  const float inv_rows = 1.0f / (float)rows;
  if (tid < A) c_log[tid] = logf(std_param[tid]);
  __syncthreads();
  int64_t q = 0;
  float d = 0.0f, var = 1.0f, sd = 1.0f;
  if (ok) {
    q = row_index(indices, r, s.N, env_begin, env_count);
    const float mu = mu_new[r * A + j];
    sd = std_param[j];
    const float a = s.actions[q * A + j], mo = s.mu[q * A + j], so = s.sigma[q * A + j];
    d = a - mu;
    var = sd * sd;
    t_a[tid] = -(d * d) / (2.0f * var) - c_log[j] - 0.918938533204672741780f;
    t_b[tid] = logf(sd / so + 1.e-5f) + (so * so + (mo - mu) * (mo - mu)) / (2.0f * var) - 0.5f;
  }
  __syncthreads();
  if (j == 0) {
    float kl = 0.0f, surr = 0.0f, vloss = 0.0f, ent = 0.0f, dv = 0.0f, dlp = 0.0f;
    if (ok) {
      float lp = 0.0f;
      for (int k = 0; k < A; ++k) {
        lp += t_a[e * A + k];
        kl += t_b[e * A + k];
        ent += 0.5f + 0.918938533204672741780f + c_log[k];
      }
      const float adv = s.advantages[q], old_lp = s.actions_log_prob[q];
      const float ratio = expf(lp - old_lp);
      const float lo = 1.0f - hp.clip_param, hi = 1.0f + hp.clip_param;
      const float rc = fminf(fmaxf(ratio, lo), hi);
      const float s1 = -adv * ratio, s2 = -adv * rc;
      surr = fmaxf(s1, s2);
      // d max(s1,s2)/d ratio (torch.max splits ties evenly; clamp passes gradient on [lo,hi] inclusive)
      const float in_rng = (ratio >= lo && ratio <= hi) ? 1.0f : 0.0f;
      float dr;
      if (s1 > s2) dr = -adv;
      else if (s1 < s2) dr = -adv * in_rng;
      else dr = 0.5f * (-adv) + 0.5f * (-adv) * in_rng;
      dlp = dr * ratio * inv_rows;
      const float v = val_new[r], ret = s.returns[q], vo = s.values[q];
      if (hp.use_clipped_value_loss) {
        const float dvv = v - vo;
        const float vc = vo + fminf(fmaxf(dvv, -hp.clip_param), hp.clip_param);
        const float l1 = (v - ret) * (v - ret), l2 = (vc - ret) * (vc - ret);
        vloss = fmaxf(l1, l2);
        const float in_v = (dvv >= -hp.clip_param && dvv <= hp.clip_param) ? 1.0f : 0.0f;
        const float g1 = 2.0f * (v - ret), g2 = 2.0f * (vc - ret) * in_v;
        dv = l1 > l2 ? g1 : (l1 < l2 ? g2 : 0.5f * g1 + 0.5f * g2);
      } else {
        vloss = (ret - v) * (ret - v);
        dv = 2.0f * (v - ret);
      }
      dv *= hp.value_loss_coef * inv_rows;
      if (ld_dval == 8 && (lo_off & 7) == 0 && ((uintptr_t)dval & 15) == 0) {
        const bf16 h = __float2bfloat16(dv);
        const bf16 l = __float2bfloat16(dv - __bfloat162float(h));
        bf16* o = dval + (size_t)r * 8;
        *reinterpret_cast<uint4*>(o) = make_uint4((uint32_t)__bfloat16_as_ushort(h), 0u, 0u, 0u);
        if (lo_off) *reinterpret_cast<uint4*>(o + lo_off) = make_uint4((uint32_t)__bfloat16_as_ushort(l), 0u, 0u, 0u);
      } else {
        store_split(dval, lo_off, (size_t)(r * ld_dval), dv);
        for (int k = 1; k < ld_dval; ++k) store_split(dval, lo_off, (size_t)(r * ld_dval + k), 0.0f);
      }
      for (int k = A; k < ld_dmu; ++k) store_split(dmu, lo_off, (size_t)(r * ld_dmu + k), 0.0f);   // pad columns
    }
    row_dlp[e] = dlp;
    row_s[e * 5 + 0] = kl; row_s[e * 5 + 1] = surr; row_s[e * 5 + 2] = vloss; row_s[e * 5 + 3] = ent; row_s[e * 5 + 4] = dv;
  }
  __syncthreads();
  float dmv = 0.0f, dsv = 0.0f;
  if (ok) {
    const float dlp = row_dlp[e];
    dmv = (d / var) * dlp;                                   // d logp / d mu    * d loss / d logp
    dsv = (d * d / (var * sd) - 1.0f / sd) * dlp;            // d logp / d sigma * d loss / d logp
    store_split(dmu, lo_off, (size_t)(r * ld_dmu + j), dmv);
  }
  if (e < kHeadRows) { t_a[tid] = dmv; t_b[tid] = dsv; }
  __syncthreads();
  // column sums over the block's rows (fixed order), then one atomic per column and block
  if (tid < A) {
    float sm = 0.0f, ss = 0.0f;
    for (int k = 0; k < kHeadRows; ++k) { sm += t_a[k * A + tid]; ss += t_b[k * A + tid]; }
    tc::fx_add(g_bias_mu + tid, sm, tc::kFxScale);
    if (blockIdx.x == 0) ss += -hp.entropy_coef / std_param[tid];   // d(-c_ent * mean entropy)/d sigma_j
    tc::fx_add(g_std + tid, ss, tc::kFxScale);
  } else if (tid >= 32 && tid < 37) {
    const int k5 = tid - 32;
    float a = 0.0f;
    for (int k = 0; k < kHeadRows; ++k) a += row_s[k * 5 + k5];
    static_assert(URLG_STAT_KL_SUM == 0 && URLG_STAT_SURR_SUM == 1 && URLG_STAT_VLOSS_SUM == 2 && URLG_STAT_ENT_SUM == 3, "sum slots");
    if (k5 < 4) tc::fx_add(sums + k5, a, tc::kFxStatScale);
    else tc::fx_add(g_bias_val, a, tc::kFxScale);
  }
}

}  // namespace urlg

namespace urlg {

// ---------------------------------------------------------------- MLP backward (dout already holds dL/d out)
// Layer l: a_l = (l == 0 ? x : h[l]), G_l = dL/d(pre-activation of its output) = (l == L-1 ? dout : dz[l+1]).
//   wgrad  dW_l[out,in] += G_l^T a_l          (A = G_l MN-major, B = a_l MN-major, split-K atomics)
//   dgrad  dz[l] = (G_l W_l) * ELU'(h[l])     (B = W_l image MN-major), column sums -> db_{l-1}
// dx_out (optional, bf16 [rows, dims[0]]): gradient wrt the network input (recurrent heads), no activation.
static urlg_gemm_desc wgrad_desc(urlg_policy* p, NetWs& n, int l, int64_t rows) {
  const int L = n.d.num_layers;
  const bf16* G = l == L - 1 ? n.dout : n.dz[l + 1];
  const int ldG = l == L - 1 ? n.opad : n.d.dims[l + 1];
  const int nout = n.d.dims[l + 1], nin = n.d.dims[l];
  urlg_gemm_desc g = {};
  g.A = G; g.lda = ldG; g.a_mn = 1;
  g.B = l == 0 ? (const void*)n.x : (const void*)n.h[l];
  g.ldb = l == 0 ? n.kpad[0] : nin; g.b_mn = 1;
  g.M = nout; g.N = nin; g.K = (int)rows;
  g.out_f32 = p->grads + n.d.w_off[l]; g.ld_f32 = nin;
  g.out_fx = reinterpret_cast<int64_t*>(p->gacc + n.d.w_off[l]);
  g.flags = URLG_GEMM_ATOMIC;
  const int bn = nin > 256 ? 256 : round_up(nin, 64);
  g.block_n = bn;
  // split K so that tiles x splits fills ONE wave of the persistent grid (split-bf16 tiles are at most 128 wide):
  // measured on B200, one wave beats two (wgrad1 23.0 vs 25.8 us, wgrad0 12.1 vs 13.6): half the partial tiles
  // and twice the k-blocks per tile to pipeline
  const int bn_eff = bn > 128 ? 128 : bn;
  const int tiles = ((nout + 127) / 128) * ((nin + bn_eff - 1) / bn_eff);
  // the wgrads of BOTH networks of a level share the wave (mlp_backward_pair): tiles x splits x 2 ~ one wave.  Measured
  // (Go2 minibatch): same launch times as per-problem one-wave splits, half the partial-slab volume for the finalize
  static const int share = []() { const char* e = getenv("URLG_WGRAD_SHARE"); return e ? atoi(e) : 2; }();
  int splits = kNumSMs / (tiles * (share > 0 ? share : 1));
  const int kblocks = (int)((rows + 63) / 64);
  if (splits > kblocks / 4) splits = kblocks / 4;
  if (splits < 1) splits = 1;
  g.k_splits = splits;
  // partial slabs instead of atomics (policy.cuh wpart): grads_finalize_kernel sums them in split order
  const int q = &n == &p->net[0] ? 0 : &n == &p->net[1] ? 1 : -1;
  if (q >= 0 && p->wpart) {
    const int eff = urlg_tc_gemm_splits((int)rows, splits);
    if (eff <= p->wpart_cap[q][l]) {
      g.out_part = p->wpart + p->wpart_off[q][l];
      g.out_fx = nullptr;
      p->wpart_splits[q][l] = eff;
    }
  }
  with_lo(p, g);
  return g;
}

// false when layer l has no data gradient to produce (l == 0 without dx_out)
static bool dgrad_desc(urlg_policy* p, NetWs& n, int l, int64_t rows, bf16* dx_out, urlg_gemm_desc& g) {
  if (!(l >= 1 || dx_out)) return false;
  const int L = n.d.num_layers;
  const bf16* G = l == L - 1 ? n.dout : n.dz[l + 1];
  const int ldG = l == L - 1 ? n.opad : n.d.dims[l + 1];
  const int nout = n.d.dims[l + 1], nin = n.d.dims[l];
  g = urlg_gemm_desc{};
  g.A = G; g.lda = ldG;
  g.B = n.w[l]; g.ldb = n.kpad[l]; g.b_mn = 1;
  g.M = (int)rows; g.N = nin; g.K = nout;
  g.block_n = nin > 256 ? 256 : round_up(nin, 64);
  if (l >= 1) {
    g.out_bf16 = n.dz[l]; g.ld_bf16 = nin;
    g.dact_ref = n.h[l]; g.ld_ref = nin;
    g.colsum = p->grads + n.d.b_off[l - 1];
    g.colsum_fx = reinterpret_cast<int64_t*>(p->gacc + n.d.b_off[l - 1]);
  } else {
    g.out_bf16 = dx_out; g.ld_bf16 = nin;
  }
  with_lo(p, g);
  return true;
}

int mlp_backward(urlg_policy* p, NetWs& n, int64_t rows, bf16* dx_out, cudaStream_t st) {
  for (int l = n.d.num_layers - 1; l >= 0; --l) {
    urlg_gemm_desc g[2];
    const urlg_gemm_desc* ptr[2];
    int k = 0;
    g[k] = wgrad_desc(p, n, l, rows); ptr[k] = &g[k]; ++k;
    if (dgrad_desc(p, n, l, rows, dx_out, g[k])) { ptr[k] = &g[k]; ++k; }
    if (int rc = tc::launch_gemm_group(ptr, k, st)) return rc;
  }
  return URLG_OK;
}

// Both networks, top layer down: level k = layer La-1-k of the actor and Lc-1-k of the critic.  A level's weight
// gradients and data gradients of both networks only READ that level's incoming gradient (written by the previous
// launch) and forward activations, and write disjoint outputs, so they are one launch of up to 4 problems: a
// backward of both networks is max(La, Lc) launches instead of 2 (La + Lc) - 2.  Long tiles (split-K weight
// gradients) first.
int mlp_backward_pair(urlg_policy* p, NetWs& a, NetWs& c, int64_t rows, bf16* dxa, bf16* dxc, cudaStream_t st) {
  const int La = a.d.num_layers, Lc = c.d.num_layers;
  for (int k = 0; k < (La > Lc ? La : Lc); ++k) {
    const int la = La - 1 - k, lc = Lc - 1 - k;
    urlg_gemm_desc g[4];
    const urlg_gemm_desc* ptr[4];
    int n = 0;
    if (la >= 0) { g[n] = wgrad_desc(p, a, la, rows); ptr[n] = &g[n]; ++n; }
    if (lc >= 0) { g[n] = wgrad_desc(p, c, lc, rows); ptr[n] = &g[n]; ++n; }
    if (la >= 0 && dgrad_desc(p, a, la, rows, dxa, g[n])) { ptr[n] = &g[n]; ++n; }
    if (lc >= 0 && dgrad_desc(p, c, lc, rows, dxc, g[n])) { ptr[n] = &g[n]; ++n; }
    if (int rc = tc::launch_gemm_group(ptr, n, st)) return rc;
  }
  return URLG_OK;
}

// ---------------------------------------------------------------- apply: norm, LR rule, clip, Adam
// Gradient norm + the host-side scalar logic of PPO.update in ONE launch: every block writes its partial sum of squares,
// the block that draws the last ticket sums the (<= 256) partials in block order (fixed-shape tree: the norm, hence the
// clip coefficient and the whole Adam step, is bit-reproducible, so data-parallel ranks stay bit-identical) and then
// runs the LR rule / Adam bias corrections / loss accumulators as one thread -- no .item(), no KL branch sync.
// Tail shared by the two gradient-norm kernels: block partial -> ticket -> the last block sums the partials in block order
// and runs the scalar logic as one thread.  `epoch` (peer exchange) is advanced by that block when given.
__device__ __forceinline__ void norm_tail(float acc, float* __restrict__ partials, unsigned int* __restrict__ ticket,
                                          float* __restrict__ stats, const urlg_ppo_hparams& hp, float rows_total,
                                          uint32_t* epoch) {
  acc = warp_sum(acc);
  __shared__ float sh[256];
  __shared__ bool last;
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.0f;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) a += sh[k];
    partials[blockIdx.x] = a;
    __threadfence();                                              // partial visible before the ticket is drawn
    last = atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1;     // wraps to 0 on the last draw: ready for the next launch
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  {
    float a = 0.0f;                                 // thread t adds partials t, t + 256, ... in that order (<= 256 blocks: one term)
    for (unsigned k = threadIdx.x; k < gridDim.x; k += 256) a += __ldcg(partials + k);
    sh[threadIdx.x] = a;
  }
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x != 0) return;
  if (epoch) *epoch = *epoch + 1u;          // every block read the old epoch before it drew its ticket
  stats[URLG_STAT_GNORM_SQ] = sh[0];
  float lr = stats[URLG_STAT_LR];
  const float kl_mean = stats[URLG_STAT_KL_SUM] / rows_total;
  stats[URLG_STAT_KL_MEAN] = kl_mean;
  if (hp.adaptive_lr) {
    if (kl_mean > hp.desired_kl * 2.0f) lr = fmaxf(hp.lr_min, lr / 1.5f);
    else if (kl_mean < hp.desired_kl / 2.0f && kl_mean > 0.0f) lr = fminf(hp.lr_max, lr * 1.5f);
  }
  stats[URLG_STAT_LR] = lr;
  const float step = stats[URLG_STAT_STEP] + 1.0f;
  stats[URLG_STAT_STEP] = step;
  stats[URLG_STAT_BC1] = 1.0f - powf(hp.beta1, step);
  stats[URLG_STAT_BC2] = 1.0f - powf(hp.beta2, step);
  const float total_norm = sqrtf(sh[0]);
  stats[URLG_STAT_CLIP_COEF] = hp.max_grad_norm > 0.0f ? fminf(hp.max_grad_norm / (total_norm + 1e-6f), 1.0f) : 1.0f;
  stats[URLG_STAT_ACC_VLOSS] += stats[URLG_STAT_VLOSS_SUM] / rows_total;
  stats[URLG_STAT_ACC_SURR] += stats[URLG_STAT_SURR_SUM] / rows_total;
  stats[URLG_STAT_ACC_N] += 1.0f;
}

__global__ void __launch_bounds__(256)
grad_norm_scalars_kernel(const float* __restrict__ g, int64_t n, float scale, float* __restrict__ partials,
                         unsigned int* __restrict__ ticket, float* __restrict__ stats, const urlg_ppo_hparams hp,
                         float rows_total) {
  float acc = 0.0f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float x = g[i] * scale;
    acc += x * x;
  }
  norm_tail(acc, partials, ticket, stats, hp, rows_total, nullptr);
}

// Data-parallel variant (a peer exchange is bound, include/urlg_b200.h urlg_policy_bind_peer_exchange): the all-reduce IS
// this kernel's load phase.  Every rank's grads_finalize_kernel has published [gradient | 4 loss sums] into its exchange
// half and flagged the peers; here every block waits for the flags, sums the `world` halves in rank order over NVLink
// (bit-identical on every rank) into the local [grads | sums] block and accumulates the norm of the scaled sum on the fly.
__global__ void __launch_bounds__(256)
peer_reduce_norm_scalars_kernel(const urlg_peer_exchange x, float* __restrict__ g, int64_t n, int64_t n_total, float scale,
                                float* __restrict__ partials, unsigned int* __restrict__ ticket, float* __restrict__ stats,
                                const urlg_ppo_hparams hp, float rows_total) {
  const uint32_t epoch = *x.epoch;
  peer_wait_flags(x, epoch + 1u);
  const size_t half = (size_t)(epoch & 1u) * x.count;
  float acc = 0.0f;
  const int64_t n4 = (n_total + 3) >> 2;          // the halves are padded to a multiple of 4 floats (zero beyond n_total)
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll 1
    for (int r0 = 0; r0 < x.world; r0 += 4) {
      float4 v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[j] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (r0 + j < x.world) v[j] = ld_relaxed_sys_v4(x.bufs[r0 + j] + half + 4 * i);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (r0 + j < x.world) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
    }
    const float e[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t k = 4 * i + j;
      if (k < n_total) g[k] = e[j];              // [grads (n) | KL, surrogate, value, entropy sums]: stats follow the gradients
      if (k < n) { const float y = e[j] * scale; acc += y * y; }
    }
  }
  norm_tail(acc, partials, ticket, stats, hp, rows_total, x.epoch);
}

// Two-shot form of the same kernel (urlg_peer_exchange.mode 1; 4 and 8 ranks): rank r sums only slice r of the published
// halves (world - 1 remote reads of n / world floats) and stores the reduced slice into every rank's result area (world - 1
// remote writes); the block that finishes the slice last flags the peers' second flag row; then every block waits until
// all `world` slices have arrived in its OWN result area and reads it locally.  Per rank 2 (world - 1) / world x n floats
// cross NVLink instead of (world - 1) x n.  The grid (<= 256 blocks) is co-resident, so the blocks that spin in the
// second wait cannot starve the ones still reducing.  Each element is reduced by exactly one rank in rank order and
// broadcast: bit-identical on every rank.
__global__ void __launch_bounds__(256)
peer_reduce2_norm_scalars_kernel(const urlg_peer_exchange x, float* __restrict__ g, int64_t n, int64_t n_total, float scale,
                                 float* __restrict__ partials, unsigned int* __restrict__ ticket, float* __restrict__ stats,
                                 const urlg_ppo_hparams hp, float rows_total) {
  const uint32_t epoch = *x.epoch;
  peer_wait_flags(x, epoch + 1u, 0);
  const size_t half = (size_t)(epoch & 1u) * x.count, res = 2 * (size_t)x.count;
  const int64_t n4 = (n_total + 3) >> 2;
  const int64_t s4 = (n4 + x.world - 1) / x.world, lo = (int64_t)x.rank * s4, hi = lo + s4 < n4 ? lo + s4 : n4;
  for (int64_t i = lo + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < hi; i += (int64_t)gridDim.x * blockDim.x) {
    float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll 1
    for (int r0 = 0; r0 < x.world; r0 += 8) {
      float4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[j] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (r0 + j < x.world) v[j] = ld_relaxed_sys_v4(x.bufs[r0 + j] + half + 4 * i);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (r0 + j < x.world) { a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w; }
    }
#pragma unroll 1
    for (int q = 0; q < x.world; ++q) reinterpret_cast<float4*>(x.bufs[q] + res)[i] = a;
  }
  __shared__ bool slice_last;
  __threadfence_system();                          // this block's slice stores are ordered at system scope
  __syncthreads();
  if (threadIdx.x == 0) slice_last = atomicInc(x.ticket + 1, gridDim.x - 1) == gridDim.x - 1;
  __syncthreads();
  if (slice_last) peer_signal(x, epoch + 1u, 1);
  peer_wait_flags(x, epoch + 1u, 1);
  const float* mine = x.bufs[x.rank] + res;
  float acc = 0.0f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 a = ld_relaxed_sys_v4(mine + 4 * i);
    const float e[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t k = 4 * i + j;
      if (k < n_total) g[k] = e[j];
      if (k < n) { const float y = e[j] * scale; acc += y * y; }
    }
  }
  norm_tail(acc, partials, ticket, stats, hp, rows_total, x.epoch);
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
            const float* __restrict__ stats, const urlg_ppo_hparams hp) {
  const float lr = stats[URLG_STAT_LR];
  const float gs = hp.grad_scale * stats[URLG_STAT_CLIP_COEF];
  const float step_size = lr / stats[URLG_STAT_BC1];
  const float bc2_sqrt = sqrtf(stats[URLG_STAT_BC2]);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * gs;
    const float mi = m[i] + (gi - m[i]) * (1.0f - hp.beta1);          // exp_avg.lerp_(grad, 1-beta1)
    const float vi = v[i] * hp.beta2 + (1.0f - hp.beta2) * gi * gi;   // mul_(beta2).addcmul_(grad, grad, 1-beta2)
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + hp.eps;
    p[i] = p[i] - step_size * (mi / denom);
  }
}

// Adam + the bf16 (hi, lo) operand images of the weights it has just updated, in one pass (feed-forward policies): the
// thread that owns parameter i also owns its image element (the K-padding columns of an image are zero from the full
// sync at create / load and never change), so the separate weight_images launch after every optimiser step is gone.
struct WImgSeg { long long off; int n, cols, ld; bf16* dst; };
struct WImgTable { WImgSeg s[2 * URLG_MAX_LAYERS]; int num; };
__global__ void __launch_bounds__(256)
adam_images_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
                   const float* __restrict__ stats, const urlg_ppo_hparams hp, const __grid_constant__ WImgTable wt,
                   int64_t lo_off) {
  const float lr = stats[URLG_STAT_LR];
  const float gs = hp.grad_scale * stats[URLG_STAT_CLIP_COEF];
  const float step_size = lr / stats[URLG_STAT_BC1];
  const float bc2_sqrt = sqrtf(stats[URLG_STAT_BC2]);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i] * gs;
    const float mi = m[i] + (gi - m[i]) * (1.0f - hp.beta1);
    const float vi = v[i] * hp.beta2 + (1.0f - hp.beta2) * gi * gi;
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + hp.eps;
    const float pn = p[i] - step_size * (mi / denom);
    p[i] = pn;
#pragma unroll 1
    for (int k = 0; k < wt.num; ++k) {
      const long long j = i - wt.s[k].off;
      if (j >= 0 && j < wt.s[k].n) {
        const int r = (int)j / wt.s[k].cols, c = (int)j - r * wt.s[k].cols;
        store_split(wt.s[k].dst, lo_off, (size_t)r * wt.s[k].ld + c, pn);
        break;
      }
    }
  }
}

// gacc (fixed point) -> flat fp32 gradient + the four per-minibatch sums; re-zeroes the accumulators for the next minibatch
// Weight matrices whose split-K partials were stored as slabs (PartTable) are summed from them in split order instead:
// blocks >= blocks_b walk the flat parameter index (fixed-point accumulators; slab-backed elements are skipped),
// blocks [0, blocks_b) sum the slabs: four consecutive elements of a slab-backed matrix per thread.
struct PartSeg { long long off; int n, splits; const float* ws; int item_end; };
struct PartTable { PartSeg s[2 * URLG_MAX_LAYERS]; int num; };
__device__ __forceinline__ float grads_finalize_body(long long* __restrict__ gacc, float* __restrict__ grads,
                                                    float* __restrict__ stats, int64_t n, const PartTable& pt, int blocks_b,
                                                    float* __restrict__ xb) {
  // the slab sums (blocks [0, blocks_b)) are the long blocks: they are dispatched first
  if ((int)blockIdx.x >= blocks_b) {
    const int64_t i = (int64_t)((int)blockIdx.x - blocks_b) * blockDim.x + threadIdx.x;
    if (i < n) {
#pragma unroll 1
      for (int s = 0; s < pt.num; ++s)
        if (i >= pt.s[s].off && i < pt.s[s].off + pt.s[s].n) return 0.0f;
      const float gv = (float)((double)gacc[i] * (1.0 / 1099511627776.0));
      grads[i] = gv;
      if (xb) xb[i] = gv;
      gacc[i] = 0;
      return gv * gv;
    } else if (i < n + 4) {
      const float sv = (float)((double)gacc[i] * (1.0 / 1073741824.0));      // URLG_STAT_KL_SUM .. URLG_STAT_ENT_SUM
      stats[i - n] = sv;
      if (xb) xb[i] = sv;
      gacc[i] = 0;
    }
    return 0.0f;
  }
  // one thread per four consecutive elements; the slabs are read sixteen at a time (independent, predicated 16-byte loads:
  // the kernel is a handful of L2 round trips) and added in split order in double precision
  const int item = (int)blockIdx.x * (int)blockDim.x + (int)threadIdx.x;
  int k = 0;
  while (k < pt.num && item >= pt.s[k].item_end) ++k;
  if (k >= pt.num) return 0.0f;
  const PartSeg& sg = pt.s[k];
  const int e0 = 4 * (item - (k ? pt.s[k - 1].item_end : 0));
  const int sn = sg.n, splits = sg.splits;
  const float* w = sg.ws + e0;
  float* out = grads + sg.off + e0;
  if ((sn & 3) == 0) {
    // fp32 adds in split order (a fixed order: still bit-reproducible; <= 148 terms, relative error ~1e-6); the
    // float -> double conversions of a double-precision sum ran on the 16-lane XU pipe and dominated this kernel
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
    for (int z = 0; z < splits; z += 16) {
      float4 tv[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        tv[j] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (z + j < splits) tv[j] = __ldcg(reinterpret_cast<const float4*>(w + (size_t)(z + j) * sn));
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) { a0 += tv[j].x; a1 += tv[j].y; a2 += tv[j].z; a3 += tv[j].w; }
    }
    out[0] = a0; out[1] = a1; out[2] = a2; out[3] = a3;
    if (xb) { float* xo = xb + sg.off + e0; xo[0] = a0; xo[1] = a1; xo[2] = a2; xo[3] = a3; }
    return a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3;
  } else {
    float sq = 0.0f;
    for (int j = 0; j < 4 && e0 + j < sn; ++j) {
      float a = 0.0f;
      for (int z = 0; z < splits; ++z) a += __ldcg(w + j + (size_t)z * sn);
      out[j] = a;
      if (xb) xb[sg.off + e0 + j] = a;
      sq += a * a;
    }
    return sq;
  }
}

// `pub` (data-parallel, a peer exchange is bound): every value is also written into this rank's exchange half and the
// block that draws the last ticket flags the peers -- the publish step of the gradient all-reduce costs no launch.
// `nrm` (single rank: nothing happens to the gradient between this kernel and the optimiser step): the squared gradient
// norm is accumulated from the values as they are written and the last block runs the scalar logic of urlg_ppo_apply
// (norm, LR rule, clip coefficient, Adam bias corrections), so apply is the Adam launch alone.
struct NormArgs { int enabled; float* partials; unsigned int* ticket; urlg_ppo_hparams hp; float rows_total; };
__global__ void __launch_bounds__(256) grads_finalize_kernel(long long* __restrict__ gacc, float* __restrict__ grads,
                                                             float* __restrict__ stats, int64_t n,
                                                             const __grid_constant__ PartTable pt, int blocks_b,
                                                             const urlg_peer_exchange pub, const __grid_constant__ NormArgs nrm) {
  uint32_t epoch = 0;
  float* xb = nullptr;
  if (pub.world > 0) {
    epoch = *pub.epoch;
    xb = pub.bufs[pub.rank] + (size_t)(epoch & 1u) * pub.count;
  }
  const float sq = grads_finalize_body(gacc, grads, stats, n, pt, blocks_b, xb);
  if (nrm.enabled) {
    norm_tail(sq, nrm.partials, nrm.ticket, stats, nrm.hp, nrm.rows_total, nullptr);
    return;
  }
  if (pub.world == 0) return;
  __shared__ bool last;
  __threadfence_system();          // the exchange half is read by other GPUs: every block orders its own stores at system scope
  __syncthreads();
  if (threadIdx.x == 0) last = atomicInc(pub.ticket, gridDim.x - 1) == gridDim.x - 1;
  __syncthreads();
  if (last) peer_signal(pub, epoch + 1u);
}

__global__ void zero_minibatch_stats_kernel(float* stats) {
  const int i = threadIdx.x;
  if (i == URLG_STAT_KL_SUM || i == URLG_STAT_SURR_SUM || i == URLG_STAT_VLOSS_SUM || i == URLG_STAT_ENT_SUM ||
      i == URLG_STAT_GNORM_SQ)
    stats[i] = 0.0f;
}

}  // namespace urlg

using namespace urlg;

extern "C" int urlg_ppo_minibatch_grads(urlg_policy* p, const urlg_storage* s, const int64_t* indices, int64_t rows,
                                        int32_t env_begin, int32_t env_count, const urlg_ppo_hparams* h, void* stream) {
  if (!p || !s || !h) return fail(URLG_EINVAL, "urlg_ppo_minibatch_grads: null argument");
  const urlg_policy_desc& d = p->desc;
  cudaStream_t st = (cudaStream_t)stream;
  if (d.recurrent) {
    if (env_count <= 0 || env_begin < 0 || env_begin + env_count > s->N)
      return fail(URLG_EINVAL, "urlg_ppo_minibatch_grads: bad env slice [%d,+%d) of %d", env_begin, env_count, s->N);
    rows = (int64_t)s->T * env_count;
    indices = nullptr;
  } else if (!indices || rows <= 0) {
    return fail(URLG_EINVAL, "urlg_ppo_minibatch_grads: indices/rows are required for a feed-forward policy");
  }
  if (rows > p->R) return fail(URLG_EINVAL, "urlg_ppo_minibatch_grads: rows %lld exceed max_rows %d", (long long)rows, p->R);
  // when the caller laid the stats block directly behind the flat gradients (the data-parallel layout, one all-reduce over
  // [grads | KL, surrogate, value, entropy sums]) the same memset clears the four sums; GNORM_SQ is overwritten by
  // urlg_ppo_apply before anything reads it
  static_assert(URLG_STAT_KL_SUM == 0 && URLG_STAT_ENT_SUM == 3, "the four per-minibatch sums lead the stats block");
  // every contribution of this minibatch goes into p->gacc (zero on entry: create / the previous finalize left it so);
  // grads_finalize_kernel at the end writes the whole flat gradient and the four sums, so nothing is cleared here
  NetWs &na = p->net[0], &nc = p->net[1];
  for (int q = 0; q < 2; ++q)
    for (int l = 0; l < URLG_MAX_LAYERS; ++l) p->wpart_splits[q][l] = 0;
  const bool fused_fwd = !d.recurrent && p->fz != nullptr;
  if (fused_fwd) {
    // gather + every forward layer of both networks in one launch; x and the hidden activations are saved for the backward
    if (int rc = fused_update_forward(p, s, indices, rows, st)) return rc;
  } else if (!d.recurrent) {
    const float* cobs = s->privileged_observations ? s->privileged_observations : s->observations;
    int blocks = (int)((rows * na.kpad[0] + 255) / 256);
    const bool shared = cobs == s->observations && na.kpad[0] == nc.kpad[0];   // no privileged obs: one gather feeds both nets
    gather_obs_kernel<<<blocks, 256, 0, st>>>(s->observations, d.num_actor_obs, indices, rows, na.x, shared ? nc.x : nullptr,
                                              p->lo_off, na.kpad[0]);
    URLG_LAUNCH_CHECK();
    if (!shared) {
      blocks = (int)((rows * nc.kpad[0] + 255) / 256);
      gather_obs_kernel<<<blocks, 256, 0, st>>>(cobs, d.num_critic_obs, indices, rows, nc.x, nullptr, p->lo_off, nc.kpad[0]);
      URLG_LAUNCH_CHECK();
    }
  } else {
    // the two memories are independent chains of latency-bound kernels (staging, projection GEMM, one persistent
    // recurrence of ~100 us on a fraction of the SMs): the critic's runs on the side stream beside the actor's
    URLG_CUDA(cudaEventRecord(p->ev_fork, st));
    URLG_CUDA(cudaStreamWaitEvent(side_stream(p, st), p->ev_fork, 0));
    const int rc_a = lstm_update_forward(p, 0, s, env_begin, env_count, st);
    const int rc_c = lstm_update_forward(p, 1, s, env_begin, env_count, side_stream(p, st));
    URLG_CUDA(cudaEventRecord(p->ev_join, side_stream(p, st)));       // join always: a capture must not end with a forked stream
    URLG_CUDA(cudaStreamWaitEvent(st, p->ev_join, 0));
    if (rc_a) return rc_a;
    if (rc_c) return rc_c;
  }
  if (!fused_fwd)
    if (int rc = mlp_forward_pair(p, na, nc, rows, st)) return rc;
  if (d.num_actions < 2 || d.num_actions > kMaxA) return fail(URLG_EINVAL, "urlg_ppo_minibatch_grads: 2..%d actions", kMaxA);
  const int head_smem = (2 * kHeadRows * d.num_actions + 6 * kHeadRows + d.num_actions) * (int)sizeof(float);
  ppo_head_kernel<<<(int)((rows + kHeadRows - 1) / kHeadRows), kHeadRows * d.num_actions, head_smem, st>>>(
      na.out, nc.out, p->params + d.std_off, *s, indices, rows, env_begin, env_count, *h, na.dout, na.opad, nc.dout,
      nc.opad, p->lo_off, p->gacc + d.std_off, p->gacc + na.d.b_off[na.d.num_layers - 1],
      p->gacc + nc.d.b_off[nc.d.num_layers - 1], p->gacc + d.num_params);
  URLG_LAUNCH_CHECK();
  if (int rc = mlp_backward_pair(p, na, nc, rows, d.recurrent ? p->lstm[0].dx_head : nullptr,
                                 d.recurrent ? p->lstm[1].dx_head : nullptr, st)) return rc;
  if (d.recurrent) {
    URLG_CUDA(cudaEventRecord(p->ev_fork, st));
    URLG_CUDA(cudaStreamWaitEvent(side_stream(p, st), p->ev_fork, 0));
    const int rc_a = lstm_update_backward(p, 0, s, env_begin, env_count, st);
    const int rc_c = lstm_update_backward(p, 1, s, env_begin, env_count, side_stream(p, st));
    URLG_CUDA(cudaEventRecord(p->ev_join, side_stream(p, st)));
    URLG_CUDA(cudaStreamWaitEvent(st, p->ev_join, 0));
    if (rc_a) return rc_a;
    if (rc_c) return rc_c;
  }
  {
    const int64_t n = d.num_params + 4;
    PartTable pt;
    pt.num = 0;
    int items = 0;
    for (int q = 0; q < 2; ++q) {
      const urlg_mlp_desc& m = p->net[q].d;
      for (int l = 0; l < m.num_layers; ++l)
        if (p->wpart_splits[q][l] > 0) {
          PartSeg& sg = pt.s[pt.num++];
          sg.off = m.w_off[l]; sg.n = m.dims[l + 1] * m.dims[l]; sg.splits = p->wpart_splits[q][l];
          sg.ws = p->wpart + p->wpart_off[q][l];
          items += (sg.n + 3) / 4;
          sg.item_end = items;
        }
    }
    const int blocks_a = (int)((n + 255) / 256), blocks_b = (items + 255) / 256;
    urlg_peer_exchange pub = {};
    if (p->peer_bound) pub = p->peer;
    // single rank: the gradient norm and the scalar logic of urlg_ppo_apply ride on this kernel (see NormArgs)
    static const bool fuse_norm = []() { const char* e = getenv("URLG_FUSED_NORM"); return !(e && e[0] == '0'); }();
    NormArgs nrm = {};
    p->norm_done = 0;
    if (fuse_norm && !p->peer_bound && h->grad_scale == 1.0f && h->kl_count_scale == 1.0f && p->norm_partials &&
        blocks_a + blocks_b <= p->norm_partials_cap) {
      nrm.enabled = 1; nrm.partials = p->norm_partials; nrm.ticket = reinterpret_cast<unsigned int*>(p->red + 257);
      nrm.hp = *h; nrm.rows_total = (float)rows;
      p->norm_done = 1;
    }
    grads_finalize_kernel<<<blocks_a + blocks_b, 256, 0, st>>>(p->gacc, p->grads, p->stats, d.num_params, pt, blocks_b, pub, nrm);
    URLG_LAUNCH_CHECK();
  }
  p->last_rows = rows;
  return URLG_OK;
}

extern "C" int urlg_ppo_apply(urlg_policy* p, const urlg_ppo_hparams* h, void* stream) {
  if (!p || !h) return fail(URLG_EINVAL, "urlg_ppo_apply: null argument");
  if (p->last_rows <= 0) return fail(URLG_EINVAL, "urlg_ppo_apply: no minibatch gradients pending");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = p->desc.num_params;
  int blocks = (int)((n + 255) / 256);
  if (blocks > kNumSMs * 4) blocks = kNumSMs * 4;
  const int nparts = blocks < 256 ? blocks : 256;      // p->red holds 256 floats
  if (p->norm_done) {
    // grads_finalize_kernel has already produced the norm and the scalars for exactly these hyper-parameters
    if (p->peer_bound || h->grad_scale != 1.0f || h->kl_count_scale != 1.0f)
      return fail(URLG_EINVAL, "urlg_ppo_apply: grad_scale / kl_count_scale / peer binding changed since urlg_ppo_minibatch_grads");
    p->norm_done = 0;
  } else {
    if (p->peer_bound && p->peer.mode == 1)
      peer_reduce2_norm_scalars_kernel<<<nparts, 256, 0, st>>>(p->peer, p->grads, n, n + 4, h->grad_scale, p->red,
                                                               reinterpret_cast<unsigned int*>(p->red + 256), p->stats, *h,
                                                               (float)p->last_rows * h->kl_count_scale);
    else if (p->peer_bound)
      peer_reduce_norm_scalars_kernel<<<nparts, 256, 0, st>>>(p->peer, p->grads, n, n + 4, h->grad_scale, p->red,
                                                              reinterpret_cast<unsigned int*>(p->red + 256), p->stats, *h,
                                                              (float)p->last_rows * h->kl_count_scale);
    else
      grad_norm_scalars_kernel<<<nparts, 256, 0, st>>>(p->grads, n, h->grad_scale, p->red,
                                                       reinterpret_cast<unsigned int*>(p->red + 256), p->stats, *h,
                                                       (float)p->last_rows * h->kl_count_scale);
    URLG_LAUNCH_CHECK();
  }
  static const bool fuse_img = []() { const char* e = getenv("URLG_ADAM_IMAGES"); return !(e && e[0] == '0'); }();
  if (fuse_img && !p->desc.recurrent) {
    WImgTable wt;
    wt.num = 0;
    for (int q = 0; q < 2; ++q) {
      const NetWs& nw = p->net[q];
      for (int l = 0; l < nw.d.num_layers; ++l) {
        WImgSeg& sg = wt.s[wt.num++];
        sg.off = nw.d.w_off[l]; sg.n = nw.d.dims[l + 1] * nw.d.dims[l]; sg.cols = nw.d.dims[l]; sg.ld = nw.kpad[l];
        sg.dst = nw.w[l];
      }
    }
    adam_images_kernel<<<blocks, 256, 0, st>>>(p->params, p->grads, p->m, p->v, n, p->stats, *h, wt, p->lo_off);
    URLG_LAUNCH_CHECK();
    return URLG_OK;
  }
  adam_kernel<<<blocks, 256, 0, st>>>(p->params, p->grads, p->m, p->v, n, p->stats, *h);
  URLG_LAUNCH_CHECK();
  return sync_weights(p, st);
}

extern "C" int urlg_policy_bind_peer_exchange(urlg_policy* p, const urlg_peer_exchange* x) {
  if (!p) return fail(URLG_EINVAL, "urlg_policy_bind_peer_exchange: null policy");
  if (!x) { p->peer_bound = 0; return URLG_OK; }
  if (!x->bufs || !x->flags || !x->epoch || !x->ticket || x->world < 1 || x->world > URLG_PEER_MAX_WORLD || x->rank < 0 ||
      x->rank >= x->world)
    return fail(URLG_EINVAL, "urlg_policy_bind_peer_exchange: bad exchange (rank %d of %d, at most %d ranks)", x->rank, x->world,
                URLG_PEER_MAX_WORLD);
  if (x->mode != 0 && x->mode != 1) return fail(URLG_EINVAL, "urlg_policy_bind_peer_exchange: mode %d (0 one-shot, 1 two-shot)", x->mode);
  if (x->count < p->desc.num_params + 4 || (x->count & 3))
    return fail(URLG_EINVAL, "urlg_policy_bind_peer_exchange: exchange halves of %lld floats, need >= %lld (multiple of 4)",
                (long long)x->count, (long long)p->desc.num_params + 4);
  if (p->stats != p->grads + p->desc.num_params)
    return fail(URLG_EINVAL, "urlg_policy_bind_peer_exchange: the stats block must follow the flat gradients directly "
                             "([grads | KL, surrogate, value, entropy sums ...] is what the ranks exchange)");
  p->peer = *x;
  p->peer_bound = 1;
  return URLG_OK;
}
