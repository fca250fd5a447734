// Data-parallel gradient exchange over peer memory (C1/C2 of SURVEY.md 8(a): the one real exchange step of the path).
// The reference is single-GPU; [UPSTREAM rsl_rl v1.0.2] algorithms/ppo.py update is where the flat gradient of a
// minibatch has to become the mean over ranks before clip_grad_norm_ / Adam.
//
// Instead of a NCCL ring (45 us for 1.5 MB at 8 ranks, launch- and hop-latency bound) every rank keeps an exchange buffer
// in symmetric memory that all peers have mapped (NVLink 5 / NVSwitch peer loads):
//
//   publish   data -> own buffer half (epoch & 1); the block that draws the last ticket fences at system scope and
//             stores epoch + 1 into word [rank] of every peer's flag row
//   reduce    every block waits until all `world` words of its OWN flag row have reached epoch + 1, then
//             data[i] = sum over ranks r = 0 .. world-1 (in rank order: every rank computes bit-identical sums) of
//             peer r's half; the last block advances the device-resident epoch
//
// One-shot: each rank reads world x n floats once; no second hop, no staging through a ring.  Two halves are enough:
// a rank overwrites half h at exchange e + 2 only after its reduce of exchange e + 1 has seen every peer's publish of
// e + 1, which is stream-ordered behind that peer's reduce (= last read) of exchange e.  Epoch and flags live in device
// memory and only ever grow, so the pair of launches is capturable in a CUDA graph and replayable.
#include "common.cuh"
#include "peer.cuh"

namespace urlg {

__global__ void __launch_bounds__(256) peer_publish_kernel(const urlg_peer_exchange x, const float* __restrict__ data, int64_t n) {
  const uint32_t epoch = *x.epoch;
  float* mine = x.bufs[x.rank] + (size_t)(epoch & 1u) * x.count;
  const int64_t n4 = n >> 2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
    reinterpret_cast<float4*>(mine)[i] = reinterpret_cast<const float4*>(data)[i];
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) mine[(n4 << 2) + threadIdx.x] = data[(n4 << 2) + threadIdx.x];
  __shared__ bool last;
  __threadfence_system();          // the half is read by other GPUs: every block orders its own stores at system scope
  __syncthreads();
  if (threadIdx.x == 0) last = atomicInc(x.ticket, gridDim.x - 1) == gridDim.x - 1;      // wraps to 0 on the last draw
  __syncthreads();
  if (!last) return;
  peer_signal(x, epoch + 1u);                   // every block's stores are ordered before the flags
}

__global__ void __launch_bounds__(256) peer_reduce_kernel(const urlg_peer_exchange x, float* __restrict__ data, int64_t n) {
  const uint32_t epoch = *x.epoch;
  peer_wait_flags(x, epoch + 1u);
  const size_t half = (size_t)(epoch & 1u) * x.count;
  const int64_t n4 = n >> 2;
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
    reinterpret_cast<float4*>(data)[i] = a;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = (n4 << 2) + threadIdx.x;
    float a = 0.0f;
    for (int r = 0; r < x.world; ++r) a += ld_relaxed_sys(x.bufs[r] + half + i);
    data[i] = a;
  }
  __shared__ bool last;
  __syncthreads();
  if (threadIdx.x == 0) last = atomicInc(x.ticket + 1, gridDim.x - 1) == gridDim.x - 1;
  __syncthreads();
  if (last && threadIdx.x == 0) *x.epoch = epoch + 1u;      // every block has read the old value (its ticket came first)
}

}  // namespace urlg

using namespace urlg;

extern "C" int urlg_peer_allreduce(const urlg_peer_exchange* x, float* data, int64_t n, void* stream) {
  if (!x || !data || !x->bufs || !x->flags || !x->epoch || !x->ticket) return fail(URLG_EINVAL, "urlg_peer_allreduce: null argument");
  if (x->world < 1 || x->world > URLG_PEER_MAX_WORLD || x->rank < 0 || x->rank >= x->world)
    return fail(URLG_EINVAL, "urlg_peer_allreduce: rank %d of world %d (at most %d ranks)", x->rank, x->world, URLG_PEER_MAX_WORLD);
  if (n <= 0 || n > x->count) return fail(URLG_EINVAL, "urlg_peer_allreduce: %lld floats do not fit the exchange buffer (%lld)", (long long)n, (long long)x->count);
  if (((uintptr_t)data & 15) || (x->count & 3)) return fail(URLG_EINVAL, "urlg_peer_allreduce: data must be 16-byte aligned and count a multiple of 4");
  cudaStream_t st = (cudaStream_t)stream;
  int blocks = (int)(((n >> 2) + 255) / 256);
  if (blocks < 1) blocks = 1;
  if (blocks > kNumSMs * 4) blocks = kNumSMs * 4;
  peer_publish_kernel<<<blocks, 256, 0, st>>>(*x, data, n);
  URLG_LAUNCH_CHECK();
  peer_reduce_kernel<<<blocks, 256, 0, st>>>(*x, data, n);
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}
