// Warp-specialised tcgen05 GEMM for the PPO half: D[M,N] = sum_k A[m,k] * B[n,k], bf16 operands staged
// in shared memory by TMA (SWIZZLE_128B), fp32 accumulator in TMEM, fused epilogue.
//   warp 0 : TMA producer     warp 1 : TMEM alloc + single-thread MMA issue     warps 2-17 : epilogue
// Persistent over tiles with a double-buffered TMEM accumulator.
// Either operand may be K-major (row-major [rows, k]) or MN-major (row-major [k, rows]); the second form is
// what dgrad (B = W[N_out, K_in]) and wgrad (A = dY[batch, N_out], B = X[batch, K_in]) need, so no transposed
// copies of weights or activations are ever made.
#pragma once
#include <cuda.h>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace urlg {
namespace tc {

constexpr int kBlockM = 128, kBlockK = 64, kMaxStages = 4;
constexpr int kEpiWarps = 16;                     // 4 per TMEM lane quarter, each a quarter of the tile's column chunks
constexpr int kGemmThreads = (2 + kEpiWarps) * 32;
constexpr int kFlagElu = 1, kFlagAtomic = 2;

struct GemmArgs {
  int M, N, K;
  int a_mn, b_mn;        // 0 = K-major, 1 = MN-major
  int block_n;           // UMMA N (multiple of 16, <= 256; multiple of 64 when b_mn)
  int k_per_split;       // multiple of 64
  int tiles_m, tiles_n, splits;   // persistent tile space: tile id = (z * tiles_n + y) * tiles_m + x
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
  int tma_ref;           // dact_ref chunks arrive by TMA tensor loads, one chunk ahead (tmR / tmRlo valid; needs tma_out)
  int tma_out;           // bf16 outputs leave through per-warp shared staging + TMA tensor stores (tmO / tmOlo valid)
  long long* out_fx;     // fixed-point accumulation targets (see urlg_gemm_desc): replace the fp32 atomics when set
  long long* colsum_fx;
  float* out_part;       // split-K partial slabs [splits][M][ld_f32] (plain stores; replaces out_f32 / out_fx when set)
};


__host__ __device__ inline int stage_bytes_a() { return kBlockM * kBlockK * 2; }
__host__ __device__ inline int stage_bytes_b(int block_n) { return block_n * kBlockK * 2; }
inline int gemm_stage_bytes(int block_n, int split) { return (split ? 2 : 1) * (stage_bytes_a() + stage_bytes_b(block_n)); }
// Epilogue staging: per warp one 32-row x 16-column bf16 chunk for the high and one for the low part (1 KB each).
constexpr int kEpiStageBytes = 2048;
inline int gemm_epi_bytes(int tma_out, int tma_ref) { return tma_out ? kEpiWarps * kEpiStageBytes * (tma_ref ? 2 : 1) : 0; }
inline int gemm_stages_for(int block_n, int split, int tma_out, int tma_ref) {
  int s = (227 * 1024 - 1024 - 256 - gemm_epi_bytes(tma_out, tma_ref)) / gemm_stage_bytes(block_n, split);
  return s > kMaxStages ? kMaxStages : s;
}
inline int gemm_smem_bytes(int block_n, int split, int stages, int tma_out, int tma_ref) {
  return stages * gemm_stage_bytes(block_n, split) + gemm_epi_bytes(tma_out, tma_ref) + 1024 + 256;
}

// ELU(x) = x > 0 ? x : exp(x) - 1.  The epilogue is instruction-bound (ncu: IPC-limited, tensor pipe ~10% busy), and
// expm1f is ~20 of its ~35 instructions per element; ex2.approx (2 ulp) - 1 has an absolute error <= 2.5e-7, far inside
// the 1e-4 forward / 1e-3 update tolerances (the values feed bf16 hi+lo operands of ~2^-17 relative resolution).
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// one FMUL + one MUFU (no denormal range fix-up: exp(x) below 2^-126 flushes to 0, i.e. ELU = -1 exactly as fp32 rounds it)
__device__ __forceinline__ float elu1(float x) { return x > 0.0f ? x : ex2_ftz(x * 1.4426950408889634f) - 1.0f; }
__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

// A launch carries up to kMaxGroup independent problems (e.g. the same layer of the actor and of the critic, or a
// layer's dgrad and wgrad): their tiles form one persistent tile space, so small problems share a launch's fixed cost
// (~4 us: launch, TMEM allocation, descriptor fetch, pipeline fill and drain) and fill each other's wave tails.
constexpr int kMaxGroup = 4;
enum { kTmA = 0, kTmB, kTmAlo, kTmBlo, kTmO, kTmOlo, kTmR, kTmRlo, kTmCount };
// G = capacity (1, 2 or 4): the parameter block is 1.2 KB per problem and is fetched at every launch, so a single
// problem is launched through the G = 1 instantiation (measured: +0.7 us per launch with the 4-problem block).
template <int G>
struct GroupArgs {
  CUtensorMap tm[G][kTmCount];
  GemmArgs g[G];
  int tile_end[G];           // exclusive prefix sums of the problems' tile counts
  int num;
  int stages, split;         // common smem pipeline depth / operand mode
  int bytes_b_max;           // stage stride of the B ring (largest block_n of the group)
  int acc_cols;              // TMEM columns per accumulator buffer (largest block_n of the group)
  int any_tma_out, any_tma_ref;
};
struct TileRef { int p, tx, ty, tz; };
template <int G>
__device__ __forceinline__ TileRef locate(const GroupArgs<G>& P, int tile) {
  TileRef t;
  t.p = 0;
  if (G > 1) {
    while (t.p + 1 < P.num && tile >= P.tile_end[t.p]) ++t.p;
  }
  const int local = tile - (t.p ? P.tile_end[t.p - 1] : 0);
  const GemmArgs& g = P.g[t.p];
  t.tx = local % g.tiles_m;
  const int r = local / g.tiles_m;
  t.ty = r % g.tiles_n;
  t.tz = r / g.tiles_n;
  return t;
}

// Persistent kernel: each CTA walks tiles blockIdx.x, blockIdx.x + gridDim.x, ...  The accumulator is double
// buffered in TMEM (2 x tmem_cols_for(block_n) columns), so the MMAs of tile i+1 run under the epilogue of tile i;
// the epilogue (the expensive part for these skinny GEMMs: bias, ELU / ELU', hi+lo split, column sums) is spread
// over 16 warps = 4 per scheduler.
template <int G>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ GroupArgs<G> P) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte alignment
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int bytes_a = stage_bytes_a(), bytes_b = P.bytes_b_max;         // ring strides; a tile loads its own block_n rows
  const int kStages = P.stages;
  // layout: A[stages] | B[stages] | (split) Alo[stages] | Blo[stages] | (tma_out) epilogue staging | barriers
  const uint32_t sA = base, sB = base + kStages * bytes_a;
  const uint32_t sAlo = sB + kStages * bytes_b, sBlo = sAlo + (P.split ? kStages * bytes_a : 0);
  const uint32_t sEpi = sBlo + (P.split ? kStages * bytes_b : 0);           // 1024-byte aligned (stage sizes are 8 KB multiples)
  const uint32_t epi_stride = P.any_tma_ref ? 2 * kEpiStageBytes : kEpiStageBytes;
  const uint32_t bars = sEpi + (P.any_tma_out ? kEpiWarps * epi_stride : 0);
  // full[kMaxStages], empty[kMaxStages], tmem_full[2], tmem_empty[2], tmem_ptr
  const uint32_t full0 = bars, empty0 = bars + 8 * kMaxStages, tfull0 = bars + 16 * kMaxStages;
  const uint32_t tempty0 = tfull0 + 16, tmem_slot = tempty0 + 16;
  const uint32_t rbar0 = bars + 128;                     // ref_full[kEpiWarps]: one per epilogue warp
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t acc_stride = (uint32_t)P.acc_cols;
  const uint32_t ncols = 2 * acc_stride;
  const int num_tiles = P.tile_end[P.num - 1];

  if (warp == 0 && lane == 0) {
    for (int q = 0; q < P.num; ++q) {
      const GemmArgs& g = P.g[q];
      tma_prefetch_desc(&P.tm[q][kTmA]);
      tma_prefetch_desc(&P.tm[q][kTmB]);
      if (P.split) { tma_prefetch_desc(&P.tm[q][kTmAlo]); tma_prefetch_desc(&P.tm[q][kTmBlo]); }
      if (g.tma_out) { tma_prefetch_desc(&P.tm[q][kTmO]); if (g.out_lo) tma_prefetch_desc(&P.tm[q][kTmOlo]); }
      if (g.tma_ref) { tma_prefetch_desc(&P.tm[q][kTmR]); if (g.dact_lo) tma_prefetch_desc(&P.tm[q][kTmRlo]); }
    }
    for (int w = 0; w < kEpiWarps; ++w) mbar_init(rbar0 + 8 * w, 1);
    for (int s = 0; s < kStages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tfull0 + 8 * a, 1); mbar_init(tempty0 + 8 * a, kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, ncols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // Programmatic dependent launch: the next kernel in the stream may start its own set-up (barriers, TMEM, descriptor
  // prefetch) on SMs this grid has already left; nothing above touches global memory, everything below waits for
  // the previous grid to have completed and flushed.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int kbt = 0;                                                       // k-blocks issued so far (all tiles)
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const TileRef tw = locate(P, tile);
        const GemmArgs& g = P.g[tw.p];
        const int m0 = tw.tx * kBlockM, n0 = tw.ty * g.block_n;
        const int k_begin = tw.tz * g.k_per_split;
        const int k_end = min(g.K, k_begin + g.k_per_split);
        const int num_kb = (k_end - k_begin + kBlockK - 1) / kBlockK;
        for (int kb = 0; kb < num_kb; ++kb, ++kbt) {
          const int s = kbt % kStages, ph = (kbt / kStages) & 1;
          mbar_wait(empty0 + 8 * s, ph ^ 1);
          const uint32_t full = full0 + 8 * s;
          mbar_arrive_expect_tx(full, (P.split ? 2 : 1) * (bytes_a + stage_bytes_b(g.block_n)));
          const int k0 = k_begin + kb * kBlockK;
          for (int part = 0; part <= P.split; ++part) {
            const CUtensorMap* ta = &P.tm[tw.p][part ? kTmAlo : kTmA];
            const CUtensorMap* tb = &P.tm[tw.p][part ? kTmBlo : kTmB];
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
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      int kbt = 0, it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const TileRef tw = locate(P, tile);
        const GemmArgs& g = P.g[tw.p];
        const uint32_t idesc = make_idesc_bf16(kBlockM, g.block_n, g.a_mn, g.b_mn);
        const uint32_t a_lbo = g.a_mn ? 8192u : 16u, b_lbo = g.b_mn ? 8192u : 16u;
        const uint32_t a_kstep = g.a_mn ? 2048u : 32u, b_kstep = g.b_mn ? 2048u : 32u;
        const int k_begin = tw.tz * g.k_per_split;
        const int k_end = min(g.K, k_begin + g.k_per_split);
        const int num_kb = (k_end - k_begin + kBlockK - 1) / kBlockK;
        const int acc = it & 1;
        mbar_wait(tempty0 + 8 * acc, ((it >> 1) & 1) ^ 1);               // epilogue drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * acc_stride;
        for (int kb = 0; kb < num_kb; ++kb, ++kbt) {
          const int s = kbt % kStages, ph = (kbt / kStages) & 1;
          mbar_wait(full0 + 8 * s, ph);
          tc_fence_after();
          const uint32_t a_src = sA + s * bytes_a, b_src = sB + s * bytes_b;
          const uint32_t al_src = sAlo + s * bytes_a, bl_src = sBlo + s * bytes_b;
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            const uint64_t ad = make_sdesc_sw128(a_src + k * a_kstep, a_lbo, 1024);
            const uint64_t bd = make_sdesc_sw128(b_src + k * b_kstep, b_lbo, 1024);
            umma_bf16(tmem_d, ad, bd, idesc, (kb | k) ? 1u : 0u);
            if (P.split) {
              const uint64_t ald = make_sdesc_sw128(al_src + k * a_kstep, a_lbo, 1024);
              const uint64_t bld = make_sdesc_sw128(bl_src + k * b_kstep, b_lbo, 1024);
              umma_bf16(tmem_d, ald, bd, idesc, 1u);
              umma_bf16(tmem_d, ad, bld, idesc, 1u);
            }
          }
          umma_commit(empty0 + 8 * s);             // frees this smem stage when the MMAs retire
        }
        umma_commit(tfull0 + 8 * acc);              // accumulator complete
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global =====
    const int lg = warp & 3;                       // TMEM lane quarter this warp may access (warp id % 4)
    const int cg = (warp - 2) >> 2;                // column group 0..3: chunks cg, cg+4, ...
    // this warp's staging chunk: row r (= lane) at 32 r bytes, its two 16-byte halves swapped when r & 4 (SWIZZLE_32B)
    const uint32_t stage_hi = sEpi + (uint32_t)(warp - 2) * epi_stride, stage_lo = stage_hi + 1024;
    const uint32_t ref_hi = stage_hi + kEpiStageBytes, ref_lo = ref_hi + 1024, rbar = rbar0 + 8 * (warp - 2);
    const uint32_t st_row = (uint32_t)lane * 32u, st_sw = (uint32_t)((lane >> 2) & 1) << 4;
    int it = 0;
    // dact_ref prefetch cursor: walks this warp's valid chunks (of the problems that read a reference) in processing
    // order, one chunk ahead of the consumer
    int ptile = blockIdx.x, pc = -1;
    uint32_t ref_phase = 0;
    auto prefetch_ref = [&]() {
      for (;;) {
        if (ptile >= num_tiles) return;
        const TileRef pw = locate(P, ptile);
        const GemmArgs& q = P.g[pw.p];
        pc = pc < 0 ? 16 * cg : pc + 64;
        if (q.tma_ref && pc < q.block_n && pw.ty * q.block_n + pc < q.N) {
          if (lane == 0) {
            mbar_arrive_expect_tx(rbar, q.dact_lo ? 2048u : 1024u);
            tma_load_2d(ref_hi, &P.tm[pw.p][kTmR], rbar, pw.ty * q.block_n + pc, pw.tx * kBlockM + lg * 32);
            if (q.dact_lo) tma_load_2d(ref_lo, &P.tm[pw.p][kTmRlo], rbar, pw.ty * q.block_n + pc, pw.tx * kBlockM + lg * 32);
          }
          return;
        }
        ptile += gridDim.x;
        pc = -1;
      }
    };
    if (P.any_tma_ref) prefetch_ref();
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const TileRef tw = locate(P, tile);
      const GemmArgs& g = P.g[tw.p];
      const CUtensorMap* tmO = &P.tm[tw.p][kTmO];
      const CUtensorMap* tmOlo = &P.tm[tw.p][kTmOlo];
      const bool vec_bf16 = (g.ld_bf16 & 7) == 0, vec_f32 = (g.ld_f32 & 3) == 0, vec_ref = (g.ld_ref & 7) == 0;
      const int m0 = tw.tx * kBlockM, n0 = tw.ty * g.block_n;
      const int acc = it & 1;
      mbar_wait(tfull0 + 8 * acc, (it >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * acc_stride;
      const int row = m0 + lg * 32 + lane;
      const bool row_ok = row < g.M;
      for (int c0 = 16 * cg; c0 < g.block_n; c0 += 64) {
        const int col0 = n0 + c0;
        if (col0 >= g.N) break;                      // uniform across the warp
        uint4 rh0, rh1, rl0 = make_uint4(0, 0, 0, 0), rl1 = rl0;
        if (g.tma_ref) {
          mbar_wait(rbar, ref_phase);
          ref_phase ^= 1;
          rh0 = ld_shared_v4(ref_hi + st_row + st_sw);
          rh1 = ld_shared_v4(ref_hi + st_row + (16u ^ st_sw));
          if (g.dact_lo) {
            rl0 = ld_shared_v4(ref_lo + st_row + st_sw);
            rl1 = ld_shared_v4(ref_lo + st_row + (16u ^ st_sw));
          }
          __syncwarp();
          prefetch_ref();                          // the next chunk's reference lands while this one is processed
        }
        float v[16];
        tmem_ld16(tmem_d + ((uint32_t)(lg * 32) << 16) + (uint32_t)c0, v);
        const int ncol = min(16, g.N - col0);
        const bool full = ncol == 16;
        if (g.bias) {
          if (full && ((uintptr_t)(g.bias + col0) & 15) == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(g.bias + col0) + i);
              v[4 * i] += b4.x; v[4 * i + 1] += b4.y; v[4 * i + 2] += b4.z; v[4 * i + 3] += b4.w;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < ncol) v[i] += g.bias[col0 + i];
          }
        }
        if (g.flags & kFlagElu) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = elu1(v[i]);
        }
        if (g.tma_ref) {
          const uint32_t aw[8] = {rh0.x, rh0.y, rh0.z, rh0.w, rh1.x, rh1.y, rh1.z, rh1.w};
          const uint32_t bw[8] = {rl0.x, rl0.y, rl0.z, rl0.w, rl1.x, rl1.y, rl1.z, rl1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {            // clipped rows / columns read zeros: factor 1, zeroed below
            const float y0 = bf16lo(aw[i]) + bf16lo(bw[i]), y1 = bf16hi(aw[i]) + bf16hi(bw[i]);
            v[2 * i] *= (y0 > 0.0f ? 1.0f : y0 + 1.0f);
            v[2 * i + 1] *= (y1 > 0.0f ? 1.0f : y1 + 1.0f);
          }
        } else if (g.dact_ref && row_ok) {
          const __nv_bfloat16* rp = g.dact_ref + (size_t)row * g.ld_ref + col0;
          const __nv_bfloat16* rl = g.dact_lo ? g.dact_lo + (size_t)row * g.ld_ref + col0 : nullptr;
          if (full && vec_ref && (((uintptr_t)rp | (uintptr_t)rl) & 15) == 0) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const uint4 a = __ldg(reinterpret_cast<const uint4*>(rp) + h);
              uint4 b = make_uint4(0, 0, 0, 0);
              if (rl) b = __ldg(reinterpret_cast<const uint4*>(rl) + h);
              const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float y0 = bf16lo(aw[i]) + bf16lo(bw[i]), y1 = bf16hi(aw[i]) + bf16hi(bw[i]);
                v[8 * h + 2 * i] *= (y0 > 0.0f ? 1.0f : y0 + 1.0f);
                v[8 * h + 2 * i + 1] *= (y1 > 0.0f ? 1.0f : y1 + 1.0f);
              }
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < ncol) {
              float y = __bfloat162float(rp[i]);
              if (rl) y += __bfloat162float(rl[i]);
              v[i] *= (y > 0.0f ? 1.0f : y + 1.0f);
            }
          }
        }
        if (!full || __any_sync(0xffffffffu, !row_ok)) {       // warp-uniform: interior tiles skip the 16 selects
#pragma unroll
          for (int i = 0; i < 16; ++i) if (!row_ok || i >= ncol) v[i] = 0.0f;
        }
        if (g.out_bf16 && g.tma_out) {
          // pack (hi, lo), stage the 32 x 16 chunk in shared memory and let TMA write full 32-byte row segments
          // (a row-per-thread st.global touches 32 lines per instruction; rows >= M / columns >= N are clipped by TMA)
          uint32_t p[8], q[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            p[i] = *reinterpret_cast<uint32_t*>(&h);
            __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * i] - bf16lo(p[i]), v[2 * i + 1] - bf16hi(p[i]));
            q[i] = *reinterpret_cast<uint32_t*>(&l);
          }
          if (lane == 0) bulk_wait_read0();          // the previous chunk's stores have finished reading the staging
          __syncwarp();
          st_shared_v4(stage_hi + st_row + st_sw, p[0], p[1], p[2], p[3]);
          st_shared_v4(stage_hi + st_row + (16u ^ st_sw), p[4], p[5], p[6], p[7]);
          if (g.out_lo) {
            st_shared_v4(stage_lo + st_row + st_sw, q[0], q[1], q[2], q[3]);
            st_shared_v4(stage_lo + st_row + (16u ^ st_sw), q[4], q[5], q[6], q[7]);
          }
          fence_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(tmO, stage_hi, col0, m0 + lg * 32);
            if (g.out_lo) tma_store_2d(tmOlo, stage_lo, col0, m0 + lg * 32);
            bulk_commit();
          }
        }
        if (row_ok && g.out_part) {
          float* o = g.out_part + ((size_t)tw.tz * g.M + row) * g.ld_f32 + col0;
          if (full && vec_f32 && ((uintptr_t)o & 15) == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(o)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < ncol) o[i] = v[i];
          }
        } else if (row_ok && g.out_fx) {
          long long* o = g.out_fx + (size_t)row * g.ld_f32 + col0;
#pragma unroll
          for (int i = 0; i < 16; ++i) if (i < ncol) fx_add(o + i, v[i], kFxScale);
        } else if (row_ok) {
          if (g.out_f32) {
            float* o = g.out_f32 + (size_t)row * g.ld_f32 + col0;
            const bool v4 = full && vec_f32 && ((uintptr_t)o & 15) == 0;
            if (g.flags & kFlagAtomic) {
              if (v4) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                  atomicAdd(reinterpret_cast<float4*>(o) + i, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
              } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) if (i < ncol) atomicAdd(o + i, v[i]);
              }
            } else if (v4) {
#pragma unroll
              for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(o)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) if (i < ncol) o[i] = v[i];
            }
          }
          if (g.out_bf16 && !g.tma_out) {
            __nv_bfloat16* o = g.out_bf16 + (size_t)row * g.ld_bf16 + col0;
            __nv_bfloat16* ol = g.out_lo ? g.out_lo + (size_t)row * g.ld_bf16 + col0 : nullptr;
            if (full && vec_bf16 && (((uintptr_t)o | (uintptr_t)ol) & 15) == 0) {
              uint32_t p[8], q[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
                p[i] = *reinterpret_cast<uint32_t*>(&h);
                __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * i] - bf16lo(p[i]), v[2 * i + 1] - bf16hi(p[i]));
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
        if (g.colsum || g.colsum_fx) {
          // 16 columns x 32 rows -> 16 sums with a transposing butterfly (16 shuffles): after the steps below lane l
          // (l even) holds the sum of column 8*b4 + 4*b3 + 2*b2 + b1 (bK = bit K of l).
          float w8[8], w4[4], w2[2], w1;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const bool up = lane & 16;
            const float send = up ? v[i] : v[i + 8], keep = up ? v[i + 8] : v[i];
            w8[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const bool up = lane & 8;
            const float send = up ? w8[i] : w8[i + 4], keep = up ? w8[i + 4] : w8[i];
            w4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
          }
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const bool up = lane & 4;
            const float send = up ? w4[i] : w4[i + 2], keep = up ? w4[i + 2] : w4[i];
            w2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
          }
          {
            const bool up = lane & 2;
            const float send = up ? w2[0] : w2[1], keep = up ? w2[1] : w2[0];
            w1 = keep + __shfl_xor_sync(0xffffffffu, send, 2);
          }
          w1 += __shfl_xor_sync(0xffffffffu, w1, 1);
          const int c = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
          if ((lane & 1) == 0 && c < ncol) {
            if (g.colsum_fx) fx_add(g.colsum_fx + col0 + c, w1, kFxScale);
            else atomicAdd(g.colsum + col0 + c, w1);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8 * acc);     // this warp is done reading accumulator `acc`
    }
    if (P.any_tma_out && lane == 0) bulk_wait_read0();   // staging must outlive the last stores' reads
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, ncols);
}

}  // namespace tc
}  // namespace urlg
