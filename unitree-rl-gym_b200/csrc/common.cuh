// Shared host/device plumbing for liburlg_b200.so (sm_100a only).
#pragma once
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

#define URLG_LAUNCH_CHECK()          \
  do {                               \
    ++urlg::g_launches;              \
    URLG_CUDA(cudaGetLastError());   \
  } while (0)

__device__ __forceinline__ int64_t eff_step(const int64_t* base, int64_t off) {
  return (base ? *base : 0) + off;
}

constexpr int kNumSMs = 148;

}  // namespace urlg
