// System-scope loads / stores of the peer-memory gradient exchange (peer_xchg.cu, ppo_kernels.cu): flags are published with
// release stores and polled with acquire loads at .sys scope; peer buffers are read with relaxed .sys loads after the acquire.
#pragma once
#include "common.cuh"

namespace urlg {

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_relaxed_sys_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Flag rows of a rank: row 0 = "rank r has published exchange e", row 1 (two-shot) = "rank r has stored its reduced slice".
// Spin until every word of this rank's flag row has reached `want` (threads 0 .. world-1 poll one word each); traps after
// ~10 s instead of hanging the GPU when a peer has died.  Call from all threads of the block.
__device__ __forceinline__ void peer_wait_flags(const urlg_peer_exchange& x, uint32_t want, int row = 0) {
  if ((int)threadIdx.x < x.world) {
    const uint32_t* f = x.flags[x.rank] + row * URLG_PEER_MAX_WORLD + threadIdx.x;
    const long long t0 = clock64();
    for (uint32_t spins = 1; (int32_t)(ld_acquire_sys(f) - want) < 0; ++spins)
      if ((spins & 0x3ffu) == 0 && clock64() - t0 > 20000000000LL) __trap();
  }
  __syncthreads();
}
// The last block of a publishing grid: order every block's stores before the flags, then flag all peers.
__device__ __forceinline__ void peer_signal(const urlg_peer_exchange& x, uint32_t value, int row = 0) {
  __threadfence_system();
  if ((int)threadIdx.x < x.world) st_release_sys(x.flags[threadIdx.x] + row * URLG_PEER_MAX_WORLD + x.rank, value);
}

}  // namespace urlg
