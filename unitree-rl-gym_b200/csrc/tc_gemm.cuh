// Warp-specialised tcgen05 GEMM for the PPO half: D[M,N] = sum_k A[m,k] * B[n,k], bf16 operands staged
// in shared memory by TMA (SWIZZLE_128B), fp32 accumulator in TMEM, fused epilogue.
//   warp 0 : TMA producer     warp 1 : TMEM alloc + single-thread MMA issue     warps 2-5 : epilogue
// Either operand may be K-major (row-major [rows, k]) or MN-major (row-major [k, rows]); the second form is
// what dgrad (B = W[N_out, K_in]) and wgrad (A = dY[batch, N_out], B = X[batch, K_in]) need, so no transposed
// copies of weights or activations are ever made.
#pragma once
#include <cuda.h>
#include "tc_ptx.cuh"

namespace urlg {
namespace tc {

constexpr int kBlockM = 128, kBlockK = 64, kMaxStages = 4;
constexpr int kGemmThreads = 192;
constexpr int kFlagElu = 1, kFlagAtomic = 2;

struct GemmArgs {
  int M, N, K;
  int a_mn, b_mn;        // 0 = K-major, 1 = MN-major
  int block_n;           // UMMA N (multiple of 16, <= 256; multiple of 64 when b_mn)
  int k_per_split;       // multiple of 64
  float* out_f32; int ld_f32;
  __nv_bfloat16* out_bf16; int ld_bf16;
  const float* bias;                              // [N] added before the activation
  const __nv_bfloat16* dact_ref; int ld_ref;      // value *= ELU'(y) with y = post-activation at the same (m,n)
  float* colsum;                                  // [N] += column sums of the stored values (bias gradients)
  int flags;
  int split;             // 1: operands are (hi, lo) bf16 pairs, D = Ahi*Bhi + Alo*Bhi + Ahi*Blo (~fp32 operand precision)
  int stages;            // smem pipeline depth (fits 227 KB)
  __nv_bfloat16* out_lo; // split: low part of the bf16 output (same ld as out_bf16)
  const __nv_bfloat16* dact_lo;   // split: low part of dact_ref
};

__host__ __device__ inline int stage_bytes_a() { return kBlockM * kBlockK * 2; }
__host__ __device__ inline int stage_bytes_b(int block_n) { return block_n * kBlockK * 2; }
inline int gemm_stage_bytes(int block_n, int split) { return (split ? 2 : 1) * (stage_bytes_a() + stage_bytes_b(block_n)); }
inline int gemm_stages_for(int block_n, int split) {
  int s = (227 * 1024 - 1024 - 256) / gemm_stage_bytes(block_n, split);
  return s > kMaxStages ? kMaxStages : s;
}
inline int gemm_smem_bytes(int block_n, int split, int stages) { return stages * gemm_stage_bytes(block_n, split) + 1024 + 256; }

__device__ __forceinline__ float elu1(float x) { return x > 0.0f ? x : expm1f(x); }

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmAlo, const __grid_constant__ CUtensorMap tmBlo, const GemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte alignment
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int bytes_a = stage_bytes_a(), bytes_b = stage_bytes_b(g.block_n);
  const int kStages = g.stages;
  // layout: A[stages] | B[stages] | (split) Alo[stages] | Blo[stages] | barriers
  const uint32_t sA = base, sB = base + kStages * bytes_a;
  const uint32_t sAlo = sB + kStages * bytes_b, sBlo = sAlo + (g.split ? kStages * bytes_a : 0);
  const uint32_t bars = sBlo + (g.split ? kStages * bytes_b : 0);   // full[kStages], empty[kStages], tmem_full, tmem_ptr
  const uint32_t full0 = bars, empty0 = bars + 8 * kMaxStages, tmem_full = bars + 16 * kMaxStages;
  const uint32_t tmem_slot = tmem_full + 8;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * kBlockM, n0 = blockIdx.y * g.block_n;
  const int k_begin = blockIdx.z * g.k_per_split;
  const int k_end = min(g.K, k_begin + g.k_per_split);
  const int num_kb = (k_end - k_begin + kBlockK - 1) / kBlockK;
  const uint32_t ncols = tmem_cols_for(g.block_n);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (g.split) { tma_prefetch_desc(&tmAlo); tma_prefetch_desc(&tmBlo); }
    for (int s = 0; s < kStages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot_ptr;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % kStages, ph = (kb / kStages) & 1;
        mbar_wait(empty0 + 8 * s, ph ^ 1);
        const uint32_t full = full0 + 8 * s;
        mbar_arrive_expect_tx(full, (g.split ? 2 : 1) * (bytes_a + bytes_b));
        const int k0 = k_begin + kb * kBlockK;
        for (int part = 0; part <= g.split; ++part) {
          const CUtensorMap* ta = part ? &tmAlo : &tmA;
          const CUtensorMap* tb = part ? &tmBlo : &tmB;
          const uint32_t a_dst = (part ? sAlo : sA) + s * bytes_a, b_dst = (part ? sBlo : sB) + s * bytes_b;
          if (!g.a_mn) {
            tma_load_2d(a_dst, ta, full, k0, m0);                        // box {64 k, 128 m}
          } else {
            tma_load_2d(a_dst, ta, full, m0, k0);                        // box {64 m, 64 k} x 2
            tma_load_2d(a_dst + 8192, ta, full, m0 + 64, k0);
          }
          if (!g.b_mn) {
            tma_load_2d(b_dst, tb, full, k0, n0);                        // box {64 k, block_n}
          } else {
            for (int j = 0; j < g.block_n / 64; ++j) tma_load_2d(b_dst + 8192 * j, tb, full, n0 + 64 * j, k0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(kBlockM, g.block_n, g.a_mn, g.b_mn);
      const uint32_t a_lbo = g.a_mn ? 8192u : 16u, b_lbo = g.b_mn ? 8192u : 16u;
      const uint32_t a_kstep = g.a_mn ? 2048u : 32u, b_kstep = g.b_mn ? 2048u : 32u;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % kStages, ph = (kb / kStages) & 1;
        mbar_wait(full0 + 8 * s, ph);
        tc_fence_after();
        const uint32_t a_src = sA + s * bytes_a, b_src = sB + s * bytes_b;
        const uint32_t al_src = sAlo + s * bytes_a, bl_src = sBlo + s * bytes_b;
#pragma unroll
        for (int k = 0; k < kBlockK / 16; ++k) {
          const uint64_t ad = make_sdesc_sw128(a_src + k * a_kstep, a_lbo, 1024);
          const uint64_t bd = make_sdesc_sw128(b_src + k * b_kstep, b_lbo, 1024);
          umma_bf16(tmem_d, ad, bd, idesc, (kb | k) ? 1u : 0u);
          if (g.split) {
            const uint64_t ald = make_sdesc_sw128(al_src + k * a_kstep, a_lbo, 1024);
            const uint64_t bld = make_sdesc_sw128(bl_src + k * b_kstep, b_lbo, 1024);
            umma_bf16(tmem_d, ald, bd, idesc, 1u);
            umma_bf16(tmem_d, ad, bld, idesc, 1u);
          }
        }
        umma_commit(empty0 + 8 * s);             // frees this smem stage when the MMAs retire
      }
      umma_commit(tmem_full);                     // accumulator complete
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global =====
    const int lg = warp & 3;                       // TMEM lane quarter this warp may access
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row = m0 + lg * 32 + lane;
    const bool row_ok = row < g.M;
    for (int c0 = 0; c0 < g.block_n; c0 += 16) {
      const int col0 = n0 + c0;
      if (col0 >= g.N) break;                      // uniform across the warp
      float v[16];
      if (num_kb > 0) {
        tmem_ld16(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.0f;
      }
      const int ncol = min(16, g.N - col0);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        if (i < ncol) {
          float x = v[i];
          if (g.bias) x += g.bias[col0 + i];
          if (g.flags & kFlagElu) x = elu1(x);
          if (g.dact_ref && row_ok) {
            float y = __bfloat162float(g.dact_ref[(size_t)row * g.ld_ref + col0 + i]);
            if (g.dact_lo) y += __bfloat162float(g.dact_lo[(size_t)row * g.ld_ref + col0 + i]);
            x *= (y > 0.0f ? 1.0f : y + 1.0f);
          }
          v[i] = row_ok ? x : 0.0f;
        } else {
          v[i] = 0.0f;
        }
      }
      if (row_ok) {
        if (g.out_f32) {
          float* o = g.out_f32 + (size_t)row * g.ld_f32 + col0;
          if (g.flags & kFlagAtomic) {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < ncol) atomicAdd(o + i, v[i]);
          } else if (ncol == 16 && ((g.ld_f32 & 3) == 0) && ((col0 & 3) == 0)) {
#pragma unroll
            for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(o)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < ncol) o[i] = v[i];
          }
        }
        if (g.out_bf16) {
          __nv_bfloat16* o = g.out_bf16 + (size_t)row * g.ld_bf16 + col0;
          __nv_bfloat16* ol = g.out_lo ? g.out_lo + (size_t)row * g.ld_bf16 + col0 : nullptr;
          if (ncol == 16 && ((g.ld_bf16 & 7) == 0) && ((col0 & 7) == 0)) {
            uint32_t p[8], q[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
              p[i] = *reinterpret_cast<uint32_t*>(&h);
              __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * i] - __low2float(h), v[2 * i + 1] - __high2float(h));
              q[i] = *reinterpret_cast<uint32_t*>(&l);
            }
            reinterpret_cast<uint4*>(o)[0] = make_uint4(p[0], p[1], p[2], p[3]);
            reinterpret_cast<uint4*>(o)[1] = make_uint4(p[4], p[5], p[6], p[7]);
            if (ol) {
              reinterpret_cast<uint4*>(ol)[0] = make_uint4(q[0], q[1], q[2], q[3]);
              reinterpret_cast<uint4*>(ol)[1] = make_uint4(q[4], q[5], q[6], q[7]);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < ncol) {
              const __nv_bfloat16 h = __float2bfloat16(v[i]);
              o[i] = h;
              if (ol) ol[i] = __float2bfloat16(v[i] - __bfloat162float(h));
            }
          }
        }
      }
      if (g.colsum) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float s = v[i];
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
          if (lane == 0 && i < ncol) atomicAdd(g.colsum + col0 + i, s);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_d, ncols);
}

}  // namespace tc
}  // namespace urlg
