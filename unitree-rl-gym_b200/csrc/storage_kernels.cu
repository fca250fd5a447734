// RolloutStorage side of the PPO half ([UPSTREAM rsl_rl v1.0.2] storage/rollout_storage.py, algorithms/ppo.py
// process_env_step; SURVEY.md §8(a).6 R7/R8, kernels K4/K5).
//   process_env_step_kernel : timeout bootstrap + reward/done store + recurrent hidden-state reset
//   gae_scan_kernel         : one thread per env, reverse scan over T in registers, coalesced over envs at
//                             each t; block-reduced double moments of the raw advantages
//   gae_normalize_kernel    : (adv - mean) / (unbiased std + 1e-8)
// All tensors are time-major [T,N,1]; 25 algorithmic bytes per transition.
#include "common.cuh"
#include <mutex>

namespace urlg {

__global__ void __launch_bounds__(256)
process_env_step_kernel(const urlg_storage s, int t, const float* __restrict__ rew, const uint8_t* __restrict__ dones,
                        const uint8_t* __restrict__ time_outs, float gamma, float* __restrict__ hidden) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.N) return;
  const size_t o = (size_t)t * s.N + i;
  float r = rew[i];
  if (time_outs) r += gamma * (s.values[o] * (float)time_outs[i]);
  s.rewards[o] = r;
  const uint8_t d = dones[i];
  s.dones[o] = d;
  if (hidden && d) {
    const int H = s.rnn_hidden;
    for (int k = 0; k < 4; ++k) {
      float* h = hidden + ((size_t)k * s.N + i) * H;
      for (int j = 0; j < H; ++j) h[j] = 0.0f;
    }
  }
}

__global__ void __launch_bounds__(256)
gae_scan_kernel(const urlg_storage s, const float* __restrict__ last_values, float gamma, float lam,
                double* __restrict__ moments) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double sum = 0.0, sumsq = 0.0;
  if (i < s.N) {
    float adv = 0.0f;
    float next_v = last_values[i];
    for (int t = s.T - 1; t >= 0; --t) {
      const size_t o = (size_t)t * s.N + i;
      const float v = s.values[o];
      const float nt = 1.0f - (float)s.dones[o];
      const float delta = s.rewards[o] + nt * gamma * next_v - v;
      adv = delta + nt * gamma * lam * adv;
      const float ret = adv + v;
      s.returns[o] = ret;
      const float a = ret - v;          // advantages = returns - values (as the reference recomputes it)
      s.advantages[o] = a;
      sum += (double)a;
      sumsq += (double)a * (double)a;
      next_v = v;
    }
  }
  __shared__ double sh[2][8];
  for (int off = 16; off > 0; off >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, off);
    sumsq += __shfl_xor_sync(0xffffffffu, sumsq, off);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh[0][w] = sum; sh[1][w] = sumsq; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { a += sh[0][k]; b += sh[1][k]; }
    atomicAdd(moments + 0, a);
    atomicAdd(moments + 1, b);
    if (blockIdx.x == 0) atomicAdd(moments + 2, (double)s.T * (double)s.N);
  }
}

__global__ void __launch_bounds__(256)
gae_normalize_kernel(float* __restrict__ adv, int64_t total, const double* __restrict__ moments) {
  const double n = moments[2];
  const double mean = moments[0] / n;
  double var = (moments[1] - moments[0] * mean) / (n - 1.0);   // unbiased (torch.std default)
  if (var < 0.0) var = 0.0;
  const float m = (float)mean, inv = 1.0f / ((float)sqrt(var) + 1e-8f);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
    adv[i] = (adv[i] - m) * inv;
}

// ---------------------------------------------------------------- allocation-free GAE (urlg_gae_noalloc)
// Same scan and normalisation as above, but the moments scratch is a file-scope device array instead of a caller
// buffer: per-block partial (sum, sumsq) pairs, folded in block order by one small kernel (no atomics), then the
// existing gae_normalize_kernel.  Every GPU that loads the library holds the array until exit (64.25 KB).
constexpr int kNoallocMaxBlocks = URLG_GAE_NOALLOC_MAX_ENVS / 256;          // one 256-thread block per 256 envs
// [0, 2*kNoallocMaxBlocks): partials (sum, sumsq) per scan block; [2*kNoallocMaxBlocks, +3): moments (sum, sumsq, count)
static __device__ __align__(256) double g_gae_noalloc_scratch[2 * kNoallocMaxBlocks + 32];

__global__ void __launch_bounds__(256)
gae_scan_partials_kernel(const urlg_storage s, const float* __restrict__ last_values, float gamma, float lam,
                         double* __restrict__ partials /* [gridDim.x][2] */) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double sum = 0.0, sumsq = 0.0;
  if (i < s.N) {
    float adv = 0.0f;
    float next_v = last_values[i];
    for (int t = s.T - 1; t >= 0; --t) {
      const size_t o = (size_t)t * s.N + i;
      const float v = s.values[o];
      const float nt = 1.0f - (float)s.dones[o];
      const float delta = s.rewards[o] + nt * gamma * next_v - v;
      adv = delta + nt * gamma * lam * adv;
      const float ret = adv + v;
      s.returns[o] = ret;
      const float a = ret - v;
      s.advantages[o] = a;
      sum += (double)a;
      sumsq += (double)a * (double)a;
      next_v = v;
    }
  }
  __shared__ double sh[2][8];
  for (int off = 16; off > 0; off >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, off);
    sumsq += __shfl_xor_sync(0xffffffffu, sumsq, off);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh[0][w] = sum; sh[1][w] = sumsq; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int k = 0; k < (int)(blockDim.x >> 5); ++k) { a += sh[0][k]; b += sh[1][k]; }
    partials[2 * blockIdx.x] = a;            // written, not accumulated: nothing in the scratch has to start at zero
    partials[2 * blockIdx.x + 1] = b;
  }
}

// One block: moments = (sum of partial sums, sum of partial sumsq, count), folded in a fixed order.
__global__ void __launch_bounds__(256)
gae_fold_partials_kernel(const double* __restrict__ partials, int nblocks, double count, double* __restrict__ moments) {
  double a = 0.0, b = 0.0;
  for (int k = threadIdx.x; k < nblocks; k += blockDim.x) { a += partials[2 * k]; b += partials[2 * k + 1]; }
  __shared__ double sh[2][256];
  sh[0][threadIdx.x] = a; sh[1][threadIdx.x] = b;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if ((int)threadIdx.x < off) { sh[0][threadIdx.x] += sh[0][threadIdx.x + off]; sh[1][threadIdx.x] += sh[1][threadIdx.x + off]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { moments[0] = sh[0][0]; moments[1] = sh[1][0]; moments[2] = count; }
}

}  // namespace urlg

using namespace urlg;

static int check_storage(const urlg_storage* s) {
  if (!s || s->T <= 0 || s->N <= 0 || !s->rewards || !s->values || !s->dones || !s->returns || !s->advantages)
    return fail(URLG_EINVAL, "urlg_storage: missing tensor or bad shape");
  return URLG_OK;
}

extern "C" int urlg_process_env_step(const urlg_storage* s, int t, const float* rew, const uint8_t* dones,
                                     const uint8_t* time_outs, float gamma, float* hidden, void* stream) {
  if (int rc = check_storage(s)) return rc;
  if (t < 0 || t >= s->T) return fail(URLG_EINVAL, "Rollout buffer overflow");   // same text as the reference's assert
  if (!rew || !dones) return fail(URLG_EINVAL, "urlg_process_env_step: rew/dones are required");
  process_env_step_kernel<<<(s->N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*s, t, rew, dones, time_outs, gamma, hidden);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

extern "C" int urlg_gae(const urlg_storage* s, const float* last_values, float gamma, float lam, double* moments,
                        int phase, void* stream) {
  if (int rc = check_storage(s)) return rc;
  if (!moments) return fail(URLG_EINVAL, "urlg_gae: moments scratch is required");
  cudaStream_t st = (cudaStream_t)stream;
  if (phase & 1) {
    if (!last_values) return fail(URLG_EINVAL, "urlg_gae: last_values is required");
    URLG_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
    gae_scan_kernel<<<(s->N + 255) / 256, 256, 0, st>>>(*s, last_values, gamma, lam, moments);
    URLG_LAUNCH_CHECK();
  }
  if (phase & 2) {
    const int64_t total = (int64_t)s->T * s->N;
    int blocks = (int)((total + 255) / 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    gae_normalize_kernel<<<blocks, 256, 0, st>>>(s->advantages, total, moments);
    URLG_LAUNCH_CHECK();
  }
  return URLG_OK;
}

// Device address of g_gae_noalloc_scratch on the current device.  In host code the array's name is only a host-side
// stand-in for the symbol, never a device pointer: the address is asked of the runtime once per device, with that
// device current, and cached.  cudaGetSymbolAddress allocates nothing (the array is part of the loaded module).
static int gae_noalloc_scratch_ptr(double** out) {
  constexpr int kMaxDevices = 64;
  static double* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  URLG_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(URLG_EINVAL, "urlg_gae_noalloc: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    double* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_gae_noalloc_scratch);
    if (e != cudaSuccess || !ptr)
      return fail(URLG_ECUDA, "urlg_gae_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    ptrs[dev] = ptr;
  }
  *out = ptrs[dev];
  return URLG_OK;
}

extern "C" int urlg_gae_noalloc(const urlg_storage* s, const float* last_values, float gamma, float lam, void* stream) {
  if (int rc = check_storage(s)) return rc;
  if (!last_values) return fail(URLG_EINVAL, "urlg_gae_noalloc: last_values is required");
  if (s->N > URLG_GAE_NOALLOC_MAX_ENVS)
    return fail(URLG_EINVAL, "urlg_gae_noalloc: num_envs %d exceeds the static scratch limit of %d envs "
                             "(use urlg_gae with a caller-provided moments buffer)", s->N, URLG_GAE_NOALLOC_MAX_ENVS);
  double* scratch = nullptr;
  if (int rc = gae_noalloc_scratch_ptr(&scratch)) return rc;
  double* moments = scratch + 2 * kNoallocMaxBlocks;
  cudaStream_t st = (cudaStream_t)stream;
  const int nblocks = (s->N + 255) / 256;
  gae_scan_partials_kernel<<<nblocks, 256, 0, st>>>(*s, last_values, gamma, lam, scratch);
  URLG_LAUNCH_CHECK();
  gae_fold_partials_kernel<<<1, 256, 0, st>>>(scratch, nblocks, (double)s->T * (double)s->N, moments);
  URLG_LAUNCH_CHECK();
  const int64_t total = (int64_t)s->T * s->N;
  int blocks = (int)((total + 255) / 256);
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  gae_normalize_kernel<<<blocks, 256, 0, st>>>(s->advantages, total, moments);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}
