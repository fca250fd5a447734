// Counter-based Philox4x32-10 (Salmon et al., Random123).  Stream layout is shared with
// oracle/philox.py: key=(seed_lo, seed_hi^stream), counter=(global_env, step_lo, step_hi, block),
// slot s -> block s/4 lane s%4, uniform = (x>>8)*2^-24 in [0,1).
#pragma once
#include <cstdint>

namespace urlg {

constexpr uint32_t kPhiloxM0 = 0xD2511F53u, kPhiloxM1 = 0xCD9E8D57u;
constexpr uint32_t kPhiloxW0 = 0x9E3779B9u, kPhiloxW1 = 0xBB67AE85u;
constexpr uint32_t kStreamEnv = 0, kStreamPolicy = 1;

// noinline: called from a dozen sites in the env kernels; one copy keeps their instruction footprint in the I-cache
static __device__ __noinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(kPhiloxM0, c.x), lo0 = kPhiloxM0 * c.x;
    uint32_t hi1 = __umulhi(kPhiloxM1, c.z), lo1 = kPhiloxM1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += kPhiloxW0;
    k1 += kPhiloxW1;
  }
  return c;
}

// inlined form for fully unrolled call sites (the key schedule is then shared between the calls)
__device__ __forceinline__ uint4 philox4x32_10_inl(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(kPhiloxM0, c.x), lo0 = kPhiloxM0 * c.x;
    uint32_t hi1 = __umulhi(kPhiloxM1, c.z), lo1 = kPhiloxM1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
    k0 += kPhiloxW0;
    k1 += kPhiloxW1;
  }
  return c;
}

// inlined form with the ten round keys precomputed (kernel-parameter / constant-bank operands)
__device__ __forceinline__ uint4 philox4x32_10_keys(uint4 c, const uint32_t (&k0)[10], const uint32_t (&k1)[10]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(kPhiloxM0, c.x), lo0 = kPhiloxM0 * c.x;
    uint32_t hi1 = __umulhi(kPhiloxM1, c.z), lo1 = kPhiloxM1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k0[r], lo1, hi0 ^ c.w ^ k1[r], lo0);
  }
  return c;
}

__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// Per-(env, step) slot reader with a one-block cache: slots are consumed mostly in order.
struct SlotRng {
  uint32_t env, step_lo, step_hi, k0, k1;
  int cur_blk;
  uint4 cur;
  const float* override_row;  // optional uniforms [num_slots] for this env (debug / parity injection)

  __device__ __forceinline__ void init(uint32_t seed_lo, uint32_t seed_hi, uint32_t stream, int64_t global_env,
                                       int64_t step, const float* ovr) {
    env = (uint32_t)global_env;
    step_lo = (uint32_t)(step & 0xffffffff);
    step_hi = (uint32_t)((uint64_t)step >> 32);
    k0 = seed_lo;
    k1 = seed_hi ^ stream;
    cur_blk = -1;
    override_row = ovr;
  }
  __device__ __forceinline__ uint32_t raw(int slot) {
    int blk = slot >> 2;
    if (blk != cur_blk) {
      cur = philox4x32_10(make_uint4(env, step_lo, step_hi, (uint32_t)blk), k0, k1);
      cur_blk = blk;
    }
    int l = slot & 3;
    return l == 0 ? cur.x : l == 1 ? cur.y : l == 2 ? cur.z : cur.w;
  }
  __device__ __forceinline__ float uniform(int slot) {
    if (override_row) return override_row[slot];
    return u01(raw(slot));
  }
};

}  // namespace urlg
