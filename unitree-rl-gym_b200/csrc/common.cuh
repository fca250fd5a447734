// Shared host/device plumbing for liburlg_b200.so (sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include "../../include/urlg_b200.h"

namespace urlg {

extern thread_local char g_err[512];
extern long long g_launches;   // kernels launched by this library since load (urlg_launch_count)

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define URLG_CUDA(expr)                                                                  \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess)                                                              \
      return urlg::fail(URLG_ECUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,         \
                        cudaGetErrorString(e__));                                        \
  } while (0)

// ---- programmatic dependent launch (PDL) -------------------------------------------------------------------------
// A kernel that starts with pdl_enter() and is launched through launch_pdl() lets the next such kernel in the stream
// onto SMs the current grid has already left (launch latency and set-up overlap the tail); griddepcontrol.wait then
// holds it until the previous grid has completed and flushed.  Used by the tcgen05 GEMM (8 back-to-back launches per
// env step, 22 per minibatch per network): measured on B200, rollout 2.55 -> 2.26 ms, PPO iteration 12.4 -> 11.6 ms.
// Extending it to the small elementwise kernels was measured too and is slower (12.4 ms), so they keep plain launches.
// URLG_PDL=0 launches everything with plain stream order.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
inline bool pdl_enabled() {
  static const bool on = []() { const char* e = getenv("URLG_PDL"); return !(e && e[0] == '0'); }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define URLG_LAUNCH_CHECK()          \
  do {                               \
    ++urlg::g_launches;              \
    URLG_CUDA(cudaGetLastError());   \
  } while (0)

__device__ __forceinline__ int64_t eff_step(const int64_t* base, int64_t off) {
  return (base ? *base : 0) + off;
}

constexpr int kNumSMs = 148;

namespace tc {
// value -> 64-bit fixed point and add: integer atomics commute, so a sum accumulated this way does not depend on the
// arrival order of its contributions (bit-reproducible gradients; include/urlg_b200.h URLG_FX_SHIFT)
__device__ __forceinline__ void fx_add(long long* p, float v, float scale) {
  atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)__float2ll_rn(v * scale));
}
constexpr float kFxScale = 1099511627776.0f;        // 2^URLG_FX_SHIFT
constexpr float kFxStatScale = 1073741824.0f;       // 2^URLG_FX_STAT_SHIFT
}  // namespace tc

}  // namespace urlg
