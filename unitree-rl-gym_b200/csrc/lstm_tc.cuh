// LSTM recurrence of PPO.update on tcgen05 (K8): the 24-step forward and backward time loops of one Memory
// ([UPSTREAM rsl_rl v1.0.2] modules/actor_critic_recurrent.py Memory.forward in batch mode + autograd's BPTT;
// shapes legged_gym/envs/h1/h1_config.py:103-115: nn.LSTM(in, 64, 1)) as ONE kernel per direction.
//
// A cluster of NC CTAs owns a tile of 128 env rows for the whole sequence; CTA r of the cluster owns hidden units
// [r*HU, (r+1)*HU), HU = 64/NC, of every gate.  Per step t:
//
//   forward   gates[128, 4*HU] (TMEM, fp32) = x_t . W_ih[own rows]^T          issued one step ahead (no dependence on h)
//                                           + h_{t-1} . W_hh[own rows]^T      K = 64: the full hidden state
//             drain: + bias, sigmoid/tanh, c' = f c + i g, h' = o tanh c'  -> saved gates / cell states (fp32, global),
//                    the h' slice as bf16 (hi, lo) into the CTA's own operand tile, then copied into the operand tiles
//                    of the other NC-1 CTAs by cp.async.bulk over distributed shared memory (completes on their mbarrier)
//   backward  drain: dgates_t from (dL/dh_t from the head + recurrent part, dL/dc) and the saved gates -> dgates image
//                    (global, for the weight-gradient GEMMs) and the CTA's A tile [128, 4*HU] (bf16 hi, lo)
//             partial dL/dh_{t-1}[128, 64] (TMEM) = dgates_t[own columns] . W_hh[own rows]      K = 4*HU
//             reduce-scatter: each CTA sends the [128, HU] block of its partial that belongs to CTA d to d (bulk copy
//                    over DSMEM), and sums the NC blocks it receives
//
// Operand precision is split bf16 (hi + lo, 3 MMAs per product, fp32 accumulate) like tc_gemm.cuh.  Weights stay
// resident in shared memory for the whole sequence (TMA, SWIZZLE_128B); x tiles arrive by TMA from the staged bf16
// image; the h / dgates operand tiles are written by the drain warps.  The h tile uses the no-swizzle K-major core
// matrix layout so a CTA's unit slice is one contiguous block (one bulk copy per peer).
//
// All bf16 images leave / enter through TMA: the operand tiles are columns of core matrices ([8-column chunk][128 rows]
// [16 bytes]), which is exactly the shared-memory side of a {8 columns, 128 rows, 1 step} tensor box, so the h_prev part
// of the [x | h] image, the h_t image of the MLP head and the dgates image are stored straight out of the tiles the MMAs
// read, and dL/dh from the head is loaded the same way (3-D maps [step, env, column]: a ragged last tile is clipped per
// step).  The fp32 state kept for the backward pass (activated gates, c_{t-1}, tanh c_t) uses a private layout
// [4-column chunk][row][4] per (step, tile, CTA), so every warp access is 512 contiguous bytes.
//
// Warp roles (18 warps): 0 = TMA producer, 1 = MMA issuer + TMEM owner, 2..17 = drains (4 per TMEM lane quarter).
#pragma once
#include <cuda.h>
#include "tc_ptx.cuh"
#include "policy.cuh"

namespace urlg {
namespace lstmtc {

using namespace urlg::tc;

constexpr int kH = 64;
constexpr int kDrainWarps = 16;
constexpr int kDrainThreads = kDrainWarps * 32;
constexpr int kThreads = (2 + kDrainWarps) * 32;
constexpr int kTileBytes = 128 * 64 * 2;            // one 128-row x 64-column bf16 tile

template <int NC>
struct Cfg {
  static constexpr int HU = kH / NC;                // hidden units owned by a CTA
  static constexpr int NG = 4 * HU;                 // gate columns owned by a CTA (gate-major: g * HU + local unit)
  static constexpr int UT = HU / 4;                 // units per drain thread (4 drain warps per lane quarter)
  static constexpr int WB = NG * 128;               // bytes of one weight tile [NG rows, 64 k] (forward)
  static constexpr int kSliceBytes = HU * 256;      // a CTA's unit slice of a no-swizzle h tile (HU/8 chunks x 2 KB)
  // forward shared memory: Wx(hi,lo) Wh(hi,lo) | X(hi,lo) | H[2](hi,lo) | h_t slice [2](hi,lo) | bias | barriers
  static constexpr int kFwdSmem = 4 * WB + 2 * kTileBytes + 4 * kTileBytes + 4 * kSliceBytes + NG * 4 + 128 + 1024;
  // backward shared memory: W(hi,lo) [64 units, NG k] | A(hi,lo) dgates tile [128, NG] | staging [2][NC][128][HU] f32 |
  // reduce [2][NC][128][HU] f32 | dL/dh_t slice from the head [2](hi,lo) | barriers
  static constexpr int kBlkBytes = 128 * HU * 4;    // one [128, HU] fp32 block of the reduce-scatter
  static constexpr int KB = NG / 64;                // 64-wide K blocks of the backward product (K = NG)
  static constexpr int kBwdSmem = 2 * KB * 8192 + 2 * KB * kTileBytes + 4 * NC * kBlkBytes + 4 * kSliceBytes + 128 + 1024;
};

struct FwdArgs {
  CUtensorMap x_hi, x_lo;        // staged x image [T*E, in_pad] (row pitch kcat), box {64, 128}
  CUtensorMap wx_hi, wx_lo;      // W_ih image [256, in_pad] (row pitch kcat), box {64, HU or 256}
  CUtensorMap wh_hi, wh_lo;      // W_hh image [256, 64] (same rows, column offset in_pad), box {64, HU or 256}
  const float* bias;             // [256] b_ih + b_hh
  CUtensorMap xh_hi, xh_lo;      // h_prev part of the [x | h] image as [T, E, 64] (row pitch kcat), box {8, 128, 1}
  CUtensorMap hx_hi, hx_lo;      // h_t image = input of the MLP head, [T, E, 64] (row pitch ld_headx), box {8, 128, 1}
  float *gates, *c_prev, *tanh_c;// private layout (see top): activated gates, c_{t-1}, tanh c_t -- read by the backward kernel
  const uint8_t* dones;          // [T, N]
  const float *saved_h, *saved_c;// [T, N, 64] hidden states saved by the rollout
  int T, E, N, env_begin, kx_steps;
};

struct BwdArgs {
  CUtensorMap w_hi, w_lo;        // permuted W_hh^T image [64 units, 256 k'], k' = r*NG + g*HU + j; box {64, 64}
  CUtensorMap dh_hi, dh_lo;      // dL/dh_t from the MLP head as [T, E, 64], box {8, 128, 1}
  CUtensorMap dg_hi, dg_lo;      // dgates image (A operand of the weight-gradient GEMMs) as [T, E, 256], box {8, 128, 1}
  const float *gates, *c_prev, *tanh_c;
  const uint8_t* dones;
  long long *g_b_ih, *g_b_hh;    // fixed-point bias gradient accumulators [256]
  int T, E, N, env_begin;
};

// ---------------------------------------------------------------- PTX helpers local to this kernel
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t out;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(out) : "r"(addr), "r"(rank));
  return out;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta -> shared::cluster bulk copy; completion (complete_tx::bytes) on an mbarrier of the destination CTA
__device__ __forceinline__ void dsmem_bulk_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster)
               : "memory");
}
__device__ __forceinline__ void drain_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kDrainThreads) : "memory"); }
__device__ __forceinline__ void st_shared_v2(uint32_t addr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}
// Shared-memory matrix descriptor, no swizzle (layout_type 0), K-major core matrices of 8 rows x 16 bytes stored as
// 128 contiguous bytes: LBO = byte distance between core matrices adjacent in K, SBO = between 8-row groups.
__device__ __forceinline__ uint64_t make_sdesc_nosw(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// byte offset of (row, unit u) in the no-swizzle h tile: 16-byte chunk u/8 is a 2 KB block of 16 core matrices
__device__ __forceinline__ uint32_t nosw_offset(uint32_t row, uint32_t u) {
  return (u >> 3) * 2048u + (row >> 3) * 128u + (row & 7u) * 16u + (u & 7u) * 2u;
}
template <int N>
__device__ __forceinline__ void tmem_ldN(uint32_t taddr, uint32_t (&r)[N]) {
  static_assert(N == 4 || N == 8 || N == 16, "tmem_ldN");
  if constexpr (N == 4) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
  } else if constexpr (N == 8) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
  } else {
    tmem_ld16_issue(taddr, r);
  }
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float lo16(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float hi16(uint32_t p) { return __uint_as_float(p & 0xffff0000u); }
// ex2.approx / rcp.approx forms: relative error ~1e-6 (sigmoid), absolute error ~3e-7 (tanh) -- far inside the 1e-3
// update tolerance; the exact expf / tanhf / IEEE-division forms are ~39 instructions each and were a third of the kernel
__device__ __forceinline__ float sigm(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_(float x) { return __fdividef(2.0f, 1.0f + __expf(-2.0f * x)) - 1.0f; }
// {8 columns, 128 rows, 1 step} box <-> 2 KB column of core matrices
__device__ __forceinline__ void tma_store_3d(const void* tmap, uint32_t src_smem, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(tmap)), "r"(src_smem), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst_smem, const void* tmap, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tmem_wait_dep(uint32_t (&r)[N]) {
  if constexpr (N == 4) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3])::"memory");
  } else if constexpr (N == 8) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])::"memory");
  } else {
    tmem_ld_wait_dep(r);
  }
}
// UT consecutive bf16 (hi, lo) pairs of a row -> packed words
template <int UT>
__device__ __forceinline__ void split_pack(const float* v, uint32_t* hi, uint32_t* lo) {
#pragma unroll
  for (int i = 0; i < UT / 2; ++i) {
    hi[i] = pack2(v[2 * i], v[2 * i + 1]);
    lo[i] = pack2(v[2 * i] - lo16(hi[i]), v[2 * i + 1] - hi16(hi[i]));
  }
}
// fp32 state in the private layout: chunk q (4 columns) of a [rows_t, ncols] block -> float4 at ((q * rows_t) + row) * 4
__device__ __forceinline__ float4* priv4(float* blk, int q, int rows_t, int row) {
  return reinterpret_cast<float4*>(blk + ((size_t)q * rows_t + row) * 4);
}
__device__ __forceinline__ const float4* priv4(const float* blk, int q, int rows_t, int row) {
  return reinterpret_cast<const float4*>(blk + ((size_t)q * rows_t + row) * 4);
}

// ================================================================ forward
template <int NC>
__global__ void __launch_bounds__(kThreads, 1) lstm_tc_fwd_kernel(const __grid_constant__ FwdArgs P) {
  using C = Cfg<NC>;
  constexpr int HU = C::HU, NG = C::NG, UT = C::UT, WB = C::WB, CH = HU / 8;
  static_assert(UT % 4 == 0, "drain threads move float4 chunks");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t sWx = base, sWh = base + 2 * WB, sX = base + 4 * WB, sH = sX + 2 * kTileBytes;
  const uint32_t sM = sH + 4 * kTileBytes;                 // h_t slice (hi, lo) x 2 buffers: source of the head-input stores
  const uint32_t sBias = sM + 4 * C::kSliceBytes;
  const uint32_t bars = sBias + NG * 4;
  const uint32_t wfull = bars, xfull = bars + 8, xempty = bars + 16, hready0 = bars + 24, gfull0 = bars + 40;
  const uint32_t tmem_slot = bars + 56;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));
  float* bias_s = reinterpret_cast<float*>(gen + (sBias - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t r = NC > 1 ? cluster_ctarank() : 0u;
  const int e0 = (blockIdx.x / NC) * 128;
  const int T = P.T, E = P.E;
  const int rows_t = min(128, E - e0);
  constexpr uint32_t kTmemCols = 2 * NG < 32 ? 32 : 2 * NG;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&P.x_hi); tma_prefetch_desc(&P.x_lo);
    tma_prefetch_desc(&P.wx_hi); tma_prefetch_desc(&P.wx_lo);
    tma_prefetch_desc(&P.wh_hi); tma_prefetch_desc(&P.wh_lo);
    tma_prefetch_desc(&P.xh_hi); tma_prefetch_desc(&P.xh_lo);
    tma_prefetch_desc(&P.hx_hi); tma_prefetch_desc(&P.hx_lo);
    mbar_init(wfull, 1); mbar_init(xfull, 1); mbar_init(xempty, 1);
    for (int b = 0; b < 2; ++b) { mbar_init(hready0 + 8 * b, 1); mbar_init(gfull0 + 8 * b, 1); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  if constexpr (NC > 1) cluster_sync_all();       // peers' barriers exist before anything is copied into them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===== producer: resident weights once, then one x tile (hi + lo) per step =====
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, 4u * WB);
      if constexpr (NC == 1) {
        tma_load_2d(sWx, &P.wx_hi, wfull, 0, 0);
        tma_load_2d(sWx + WB, &P.wx_lo, wfull, 0, 0);
        tma_load_2d(sWh, &P.wh_hi, wfull, 0, 0);
        tma_load_2d(sWh + WB, &P.wh_lo, wfull, 0, 0);
      } else {
        for (int g = 0; g < 4; ++g) {            // rows g*64 + r*HU .. of the [256, k] images -> tile rows g*HU ..
          const int wr = g * kH + (int)r * HU;
          const uint32_t o = (uint32_t)g * HU * 128u;
          tma_load_2d(sWx + o, &P.wx_hi, wfull, 0, wr);
          tma_load_2d(sWx + WB + o, &P.wx_lo, wfull, 0, wr);
          tma_load_2d(sWh + o, &P.wh_hi, wfull, 0, wr);
          tma_load_2d(sWh + WB + o, &P.wh_lo, wfull, 0, wr);
        }
      }
      for (int t = 0; t < T; ++t) {
        mbar_wait(xempty, (t & 1) ^ 1);
        mbar_arrive_expect_tx(xfull, 2u * kTileBytes);
        tma_load_2d(sX, &P.x_hi, xfull, 0, t * E + e0);
        tma_load_2d(sX + kTileBytes, &P.x_lo, xfull, 0, t * E + e0);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: x part of step t+1 goes out right behind the h part of step t =====
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, NG, 0, 0);
      constexpr uint32_t lbo = 2048u, sbo = 128u;      // K-adjacent core matrices 2 KB apart, 8-row groups 128 B apart
      mbar_wait(wfull, 0);
      auto x_mma = [&](int t) {
        mbar_wait(xfull, t & 1);
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(t & 1) * NG;
        for (int k = 0; k < P.kx_steps; ++k) {
          const uint64_t ad = make_sdesc_sw128(sX + k * 32, 16, 1024), ald = make_sdesc_sw128(sX + kTileBytes + k * 32, 16, 1024);
          const uint64_t bd = make_sdesc_sw128(sWx + k * 32, 16, 1024), bld = make_sdesc_sw128(sWx + WB + k * 32, 16, 1024);
          umma_bf16(d, ad, bd, idesc, k == 0 ? 0u : 1u);
          umma_bf16(d, ald, bd, idesc, 1u);
          umma_bf16(d, ad, bld, idesc, 1u);
        }
        umma_commit(xempty);
      };
      x_mma(0);
      for (int t = 0; t < T; ++t) {
        const uint32_t b = (uint32_t)t & 1u;
        mbar_wait(hready0 + 8 * b, (t >> 1) & 1);
        tc_fence_after();
        const uint32_t d = tmem_base + b * NG;
        const uint32_t hh = sH + b * 2 * kTileBytes, hl = hh + kTileBytes;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t ad = make_sdesc_nosw(hh + k * 4096, lbo, sbo), ald = make_sdesc_nosw(hl + k * 4096, lbo, sbo);
          const uint64_t bd = make_sdesc_sw128(sWh + k * 32, 16, 1024), bld = make_sdesc_sw128(sWh + WB + k * 32, 16, 1024);
          umma_bf16(d, ad, bd, idesc, 1u);
          umma_bf16(d, ald, bd, idesc, 1u);
          umma_bf16(d, ad, bld, idesc, 1u);
        }
        umma_commit(gfull0 + 8 * b);
        if (t + 1 < T) x_mma(t + 1);
      }
    }
  } else {
    // ===== drains: thread = (row, UT units of the CTA's slice) for the whole sequence =====
    const int dw = warp - 2, lq = warp & 3, cg = dw >> 2;
    const int row = lq * 32 + lane, e = e0 + row;
    const bool valid = e < E;
    const bool elected = dw == 0 && lane == 0;
    const int ul0 = cg * UT, u0 = (int)r * HU + ul0;
    for (int i = dw * 32 + lane; i < NG; i += kDrainThreads) bias_s[i] = P.bias[(i / HU) * kH + (int)r * HU + (i % HU)];
    float c[UT], h[UT];
    auto seed_from = [&](int t) {
      const size_t sv = ((size_t)t * P.N + P.env_begin + e) * kH + u0;
#pragma unroll
      for (int i = 0; i < UT / 4; ++i) {
        const float4 a = *reinterpret_cast<const float4*>(P.saved_h + sv + 4 * i);
        const float4 b4 = *reinterpret_cast<const float4*>(P.saved_c + sv + 4 * i);
        h[4 * i] = a.x; h[4 * i + 1] = a.y; h[4 * i + 2] = a.z; h[4 * i + 3] = a.w;
        c[4 * i] = b4.x; c[4 * i + 1] = b4.y; c[4 * i + 2] = b4.z; c[4 * i + 3] = b4.w;
      }
    };
    // bf16 (hi, lo) of this thread's UT units into a column-of-core-matrices tile (unit index `u` within that tile)
    auto put_tile = [&](uint32_t tile_hi, uint32_t tile_lo, int u, const float* v) {
      uint32_t hi[UT / 2], lo[UT / 2];
      split_pack<UT>(v, hi, lo);
      const uint32_t off = nosw_offset((uint32_t)row, (uint32_t)u);
      if constexpr (UT == 4) {
        st_shared_v2(tile_hi + off, hi[0], hi[1]);
        st_shared_v2(tile_lo + off, lo[0], lo[1]);
      } else {
#pragma unroll
        for (int q = 0; q < UT / 8; ++q) {
          st_shared_v4(tile_hi + off + q * 2048, hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
          st_shared_v4(tile_lo + off + q * 2048, lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
        }
      }
    };
    // End of step tn-1 (tn = 0: prologue; tn = T: epilogue).  (h, c) as they ENTER step tn go into the operand tile, the
    // h_prev image and the c_prev state; h_{tn-1} (already in the slice tile) goes to the head-input image.
    auto publish = [&](int tn) {
      const uint32_t nb = (uint32_t)tn & 1u;
      const uint32_t tile = sH + nb * 2 * kTileBytes;
      if (tn < T) {
        put_tile(tile, tile + kTileBytes, u0, h);
        if (valid) {
          float* cp = P.c_prev + ((size_t)tn * E + e0) * kH + (size_t)r * rows_t * HU;
#pragma unroll
          for (int i = 0; i < UT / 4; ++i) *priv4(cp, cg * (UT / 4) + i, rows_t, row) = make_float4(c[4 * i], c[4 * i + 1], c[4 * i + 2], c[4 * i + 3]);
        }
      }
      fence_proxy_async_smem();
      if (elected) bulk_wait_read0();           // the previous step's tensor stores have finished reading their tiles
      drain_bar_sync();
      if (elected) {
        if (tn < T) {
          const uint32_t bar = hready0 + 8 * nb;
          const uint32_t src = tile + r * C::kSliceBytes;
          if constexpr (NC > 1) {
            mbar_arrive_expect_tx(bar, (uint32_t)(NC - 1) * 2u * C::kSliceBytes);
            for (uint32_t d = 0; d < (uint32_t)NC; ++d) {
              if (d == r) continue;
              const uint32_t rbar = mapa_shared(bar, d);
              dsmem_bulk_copy(mapa_shared(src, d), src, C::kSliceBytes, rbar);
              dsmem_bulk_copy(mapa_shared(src + kTileBytes, d), src + kTileBytes, C::kSliceBytes, rbar);
            }
          } else {
            mbar_arrive(bar);
          }
#pragma unroll
          for (int q = 0; q < CH; ++q) {            // h_prev part of the [x | h] image, rows of step tn
            tma_store_3d(&P.xh_hi, src + q * 2048, (int)r * HU + 8 * q, e0, tn);
            tma_store_3d(&P.xh_lo, src + kTileBytes + q * 2048, (int)r * HU + 8 * q, e0, tn);
          }
        }
        if (tn > 0) {
          const uint32_t m = sM + (uint32_t)((tn - 1) & 1) * 2 * C::kSliceBytes;
#pragma unroll
          for (int q = 0; q < CH; ++q) {            // h_{tn-1}: input of the MLP head, rows of step tn-1
            tma_store_3d(&P.hx_hi, m + q * 2048, (int)r * HU + 8 * q, e0, tn - 1);
            tma_store_3d(&P.hx_lo, m + C::kSliceBytes + q * 2048, (int)r * HU + 8 * q, e0, tn - 1);
          }
        }
        bulk_commit();
      }
    };
#pragma unroll
    for (int j = 0; j < UT; ++j) { h[j] = 0.0f; c[j] = 0.0f; }
    if (valid) seed_from(0);
    publish(0);
    for (int t = 0; t < T; ++t) {
      const uint32_t b = (uint32_t)t & 1u;
      // a done at t starts a new trajectory at t + 1 (fetched before the wait: the byte is off the critical path)
      const bool reseed = valid && t + 1 < T && P.dones[(size_t)t * P.N + P.env_begin + e] != 0;
      mbar_wait(gfull0 + 8 * b, (t >> 1) & 1);
      tc_fence_after();
      uint32_t rg[4][UT];
      const uint32_t taddr = tmem_base + ((uint32_t)(lq * 32) << 16) + b * NG + (uint32_t)ul0;
#pragma unroll
      for (int g = 0; g < 4; ++g) tmem_ldN<UT>(taddr + g * HU, rg[g]);
#pragma unroll
      for (int g = 0; g < 4; ++g) tmem_wait_dep<UT>(rg[g]);
      tc_fence_before();
      float ga[4][UT], tch[UT];
#pragma unroll
      for (int j = 0; j < UT; ++j) {
        const float ig = sigm(__uint_as_float(rg[0][j]) + bias_s[ul0 + j]);
        const float fg = sigm(__uint_as_float(rg[1][j]) + bias_s[HU + ul0 + j]);
        const float gg = tanh_(__uint_as_float(rg[2][j]) + bias_s[2 * HU + ul0 + j]);
        const float og = sigm(__uint_as_float(rg[3][j]) + bias_s[3 * HU + ul0 + j]);
        c[j] = fg * c[j] + ig * gg;
        tch[j] = tanh_(c[j]);
        h[j] = og * tch[j];
        ga[0][j] = ig; ga[1][j] = fg; ga[2][j] = gg; ga[3][j] = og;
      }
      if (valid) {
        float* gb = P.gates + ((size_t)t * E + e0) * 4 * kH + (size_t)r * rows_t * NG;
        float* tb = P.tanh_c + ((size_t)t * E + e0) * kH + (size_t)r * rows_t * HU;
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
          for (int i = 0; i < UT / 4; ++i)
            *priv4(gb, g * (HU / 4) + cg * (UT / 4) + i, rows_t, row) = make_float4(ga[g][4 * i], ga[g][4 * i + 1], ga[g][4 * i + 2], ga[g][4 * i + 3]);
#pragma unroll
        for (int i = 0; i < UT / 4; ++i)
          *priv4(tb, cg * (UT / 4) + i, rows_t, row) = make_float4(tch[4 * i], tch[4 * i + 1], tch[4 * i + 2], tch[4 * i + 3]);
      }
      const uint32_t m = sM + b * 2 * C::kSliceBytes;
      put_tile(m, m + C::kSliceBytes, ul0, h);
      if (reseed) seed_from(t + 1);
      publish(t + 1);
    }
    if (elected) bulk_wait_all0();
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (NC > 1) cluster_sync_all();       // outgoing copies read this CTA's tiles until the peers' last step
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

// ================================================================ backward (BPTT)
// W_hh [256, 64] (nn.LSTM layout, fp32) -> bf16 (hi, lo) image [64 units n][256 k'] with k' = r*NG + g*HU + j for the
// gate row q = g*64 + r*HU + j: CTA r's K slice of the product dgates . W_hh is then one contiguous column range.
__global__ void __launch_bounds__(256)
lstm_tc_wt_image_kernel(const float* __restrict__ W_hh, int HU, bf16* __restrict__ img /* hi [64*256] then lo */) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 4 * kH * kH) return;
  const int n = idx / (4 * kH), kp = idx % (4 * kH);
  const int NG = 4 * HU, r = kp / NG, g = (kp % NG) / HU, j = kp % HU;
  const float v = W_hh[(size_t)(g * kH + r * HU + j) * kH + n];
  const bf16 h = __float2bfloat16(v);
  img[idx] = h;
  img[idx + 4 * kH * kH] = __float2bfloat16(v - __bfloat162float(h));
}

template <int NC>
__global__ void __launch_bounds__(kThreads, 1) lstm_tc_bwd_kernel(const __grid_constant__ BwdArgs P) {
  using C = Cfg<NC>;
  constexpr int HU = C::HU, NG = C::NG, UT = C::UT, KB = C::KB, CH = HU / 8;
  constexpr uint32_t kBlk = C::kBlkBytes;
  static_assert(NC > 1 && UT % 4 == 0, "the backward kernel exchanges partial sums across a cluster");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t sW = base, sA = sW + 2 * KB * 8192, sStage = sA + 2 * KB * kTileBytes;
  const uint32_t sRed = sStage + 2 * NC * kBlk;
  const uint32_t sDh = sRed + 2 * NC * kBlk;             // dL/dh_t slice from the head (hi, lo) x 2 buffers
  const uint32_t bars = sDh + 4 * C::kSliceBytes;
  const uint32_t wfull = bars, aready = bars + 8, dfull = bars + 16, rready0 = bars + 24, dhfull0 = bars + 40;
  const uint32_t tmem_slot = bars + 56;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t r = cluster_ctarank();
  const int e0 = (blockIdx.x / NC) * 128;
  const int T = P.T, E = P.E;
  const int rows_t = min(128, E - e0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&P.w_hi); tma_prefetch_desc(&P.w_lo);
    tma_prefetch_desc(&P.dh_hi); tma_prefetch_desc(&P.dh_lo);
    tma_prefetch_desc(&P.dg_hi); tma_prefetch_desc(&P.dg_lo);
    mbar_init(wfull, 1); mbar_init(aready, 1); mbar_init(dfull, 1);
    for (int b = 0; b < 2; ++b) { mbar_init(rready0 + 8 * b, 1); mbar_init(dhfull0 + 8 * b, 1); }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    if (lane == 0) {                                   // resident W_hh slice: KB blocks of [64 units, 64 k] (hi, lo)
      mbar_arrive_expect_tx(wfull, 2u * KB * 8192u);
      for (int kb = 0; kb < KB; ++kb) {
        tma_load_2d(sW + kb * 8192, &P.w_hi, wfull, (int)r * NG + kb * 64, 0);
        tma_load_2d(sW + (KB + kb) * 8192, &P.w_lo, wfull, (int)r * NG + kb * 64, 0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, kH, 0, 0);
      mbar_wait(wfull, 0);
      for (int s = 0; s < T - 1; ++s) {                // one product per step t = T-1 .. 1
        mbar_wait(aready, s & 1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < NG / 16; ++k) {            // A: columns of core matrices (2 KB per 8 k); W: SWIZZLE_128B blocks of 64 k
          const uint32_t ah = sA + k * 4096, al = sA + KB * kTileBytes + k * 4096;
          const uint32_t bh = sW + (k >> 2) * 8192 + (k & 3) * 32, bl = bh + KB * 8192;
          const uint64_t ad = make_sdesc_nosw(ah, 2048, 128), ald = make_sdesc_nosw(al, 2048, 128);
          const uint64_t bd = make_sdesc_sw128(bh, 16, 1024), bld = make_sdesc_sw128(bl, 16, 1024);
          umma_bf16(tmem_base, ad, bd, idesc, k ? 1u : 0u);
          umma_bf16(tmem_base, ald, bd, idesc, 1u);
          umma_bf16(tmem_base, ad, bld, idesc, 1u);
        }
        umma_commit(dfull);
      }
    }
  } else {
    const int dw = warp - 2, lq = warp & 3, cg = dw >> 2;
    const int row = lq * 32 + lane, e = e0 + row;
    const bool valid = e < E;
    const bool elected = dw == 0 && lane == 0;
    const int ul0 = cg * UT, u0 = (int)r * HU + ul0;
    float dc[UT], dh_rec[UT], bsum[4][UT];
#pragma unroll
    for (int j = 0; j < UT; ++j) { dc[j] = 0.0f; dh_rec[j] = 0.0f; bsum[0][j] = bsum[1][j] = bsum[2][j] = bsum[3][j] = 0.0f; }
    // dL/dh_t from the head: this CTA's unit slice of step t, TMA -> columns of core matrices, two steps in flight
    auto load_dh = [&](int t, uint32_t buf) {
      const uint32_t bar = dhfull0 + 8 * buf, dst = sDh + buf * 2 * C::kSliceBytes;
      mbar_arrive_expect_tx(bar, 2u * C::kSliceBytes);
#pragma unroll
      for (int q = 0; q < CH; ++q) {
        tma_load_3d(dst + q * 2048, &P.dh_hi, bar, (int)r * HU + 8 * q, e0, t);
        tma_load_3d(dst + C::kSliceBytes + q * 2048, &P.dh_lo, bar, (int)r * HU + 8 * q, e0, t);
      }
    };
    if (elected) {
      load_dh(T - 1, 0);
      if (T > 1) load_dh(T - 2, 1);
    }
    // saved state of a step (private layout), prefetched one step ahead while the previous step's product is in flight
    float ga[4][UT], cp[UT], tch[UT];
    bool carry = false;
    auto load_step = [&](int t) {
      carry = false;
      if (!valid) return;
      const float* gb = P.gates + ((size_t)t * E + e0) * 4 * kH + (size_t)r * rows_t * NG;
      const float* cb = P.c_prev + ((size_t)t * E + e0) * kH + (size_t)r * rows_t * HU;
      const float* tb = P.tanh_c + ((size_t)t * E + e0) * kH + (size_t)r * rows_t * HU;
#pragma unroll
      for (int i = 0; i < UT / 4; ++i) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const float4 v = *priv4(gb, g * (HU / 4) + cg * (UT / 4) + i, rows_t, row);
          ga[g][4 * i] = v.x; ga[g][4 * i + 1] = v.y; ga[g][4 * i + 2] = v.z; ga[g][4 * i + 3] = v.w;
        }
        const float4 a = *priv4(cb, cg * (UT / 4) + i, rows_t, row);
        const float4 b = *priv4(tb, cg * (UT / 4) + i, rows_t, row);
        cp[4 * i] = a.x; cp[4 * i + 1] = a.y; cp[4 * i + 2] = a.z; cp[4 * i + 3] = a.w;
        tch[4 * i] = b.x; tch[4 * i + 1] = b.y; tch[4 * i + 2] = b.z; tch[4 * i + 3] = b.w;
      }
      carry = t < T - 1 && P.dones[(size_t)t * P.N + P.env_begin + e] == 0;   // a done at t cuts the gradient from t+1
    };
    load_step(T - 1);
    for (int t = T - 1, s = 0; t >= 0; --t, ++s) {
      const uint32_t pb = (uint32_t)s & 1u;
      // ---- dgates of step t ----
      mbar_wait(dhfull0 + 8 * pb, (s >> 1) & 1);
      float dhh[UT];
      {
        const uint32_t a = sDh + pb * 2 * C::kSliceBytes + nosw_offset((uint32_t)row, (uint32_t)ul0);
#pragma unroll
        for (int q = 0; q < UT / 4; ++q) {
          uint2 hh, hl;
          asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(hh.x), "=r"(hh.y) : "r"(a + 8 * q) : "memory");
          asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(hl.x), "=r"(hl.y) : "r"(a + C::kSliceBytes + 8 * q) : "memory");
          dhh[4 * q] = lo16(hh.x) + lo16(hl.x); dhh[4 * q + 1] = hi16(hh.x) + hi16(hl.x);
          dhh[4 * q + 2] = lo16(hh.y) + lo16(hl.y); dhh[4 * q + 3] = hi16(hh.y) + hi16(hl.y);
        }
      }
      float pg[4][UT];
#pragma unroll
      for (int j = 0; j < UT; ++j) {
        float pi = 0.0f, pf = 0.0f, pgg = 0.0f, po = 0.0f;
        if (valid) {
          const float dh = dhh[j] + (carry ? dh_rec[j] : 0.0f);
          const float dcn_in = carry ? dc[j] : 0.0f;
          const float ig = ga[0][j], fg = ga[1][j], gg = ga[2][j], og = ga[3][j];
          const float th = tch[j];
          const float d_o = dh * th;
          const float dcn = dcn_in + dh * og * (1.0f - th * th);
          const float di = dcn * gg, dgg = dcn * ig, df = dcn * cp[j];
          dc[j] = dcn * fg;
          pi = di * ig * (1.0f - ig); pf = df * fg * (1.0f - fg); pgg = dgg * (1.0f - gg * gg); po = d_o * og * (1.0f - og);
        }
        pg[0][j] = pi; pg[1][j] = pf; pg[2][j] = pgg; pg[3][j] = po;
        bsum[0][j] += pi; bsum[1][j] += pf; bsum[2][j] += pgg; bsum[3][j] += po;
      }
      // ---- A tile (columns of core matrices, k = g*HU + local unit): MMA operand AND source of the dgates image ----
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint32_t hi[UT / 2], lo[UT / 2];
        split_pack<UT>(pg[g], hi, lo);
        const uint32_t off = nosw_offset((uint32_t)row, (uint32_t)(g * HU + ul0));
        if constexpr (UT == 4) {
          st_shared_v2(sA + off, hi[0], hi[1]);
          st_shared_v2(sA + KB * kTileBytes + off, lo[0], lo[1]);
        } else {
#pragma unroll
          for (int q = 0; q < UT / 8; ++q) {
            st_shared_v4(sA + off + q * 2048, hi[4 * q], hi[4 * q + 1], hi[4 * q + 2], hi[4 * q + 3]);
            st_shared_v4(sA + KB * kTileBytes + off + q * 2048, lo[4 * q], lo[4 * q + 1], lo[4 * q + 2], lo[4 * q + 3]);
          }
        }
      }
      fence_proxy_async_smem();
      drain_bar_sync();
      if (elected) {
        if (t > 0) mbar_arrive(aready);
#pragma unroll
        for (int ch = 0; ch < NG / 8; ++ch) {            // chunk ch = columns g*64 + r*HU + (8ch mod HU) of the dgates image
          const int g = (8 * ch) / HU, ul = (8 * ch) % HU;
          tma_store_3d(&P.dg_hi, sA + ch * 2048, g * kH + (int)r * HU + ul, e0, t);
          tma_store_3d(&P.dg_lo, sA + KB * kTileBytes + ch * 2048, g * kH + (int)r * HU + ul, e0, t);
        }
        bulk_commit();
        if (t >= 2) load_dh(t - 2, pb);                  // everyone has read buffer pb (barrier above)
      }
      if (t == 0) break;
      load_step(t - 1);                                  // overlaps the product
      mbar_wait(dfull, s & 1);
      tc_fence_after();
      // partial dL/dh_{t-1}[row, 16cg .. 16cg+16) -> the block of the CTA that owns those units
      uint32_t rr[16];
      tmem_ld16_issue(tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)(16 * cg), rr);
      tmem_ld_wait_dep(rr);
      tc_fence_before();
      const uint32_t dest = (uint32_t)(16 * cg) / HU, within = (uint32_t)(16 * cg) % HU;
      const uint32_t sp = sStage + (pb * NC + dest) * kBlk + (uint32_t)row * (HU * 4) + within * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) st_shared_v4(sp + 16 * q, rr[4 * q], rr[4 * q + 1], rr[4 * q + 2], rr[4 * q + 3]);
      fence_proxy_async_smem();
      if (elected) bulk_wait_read0();                    // the dgates stores have finished reading the A tile
      drain_bar_sync();
      const uint32_t bar = rready0 + 8 * pb;
      if (elected) {
        mbar_arrive_expect_tx(bar, (uint32_t)(NC - 1) * kBlk);
        for (uint32_t d = 0; d < (uint32_t)NC; ++d) {
          if (d == r) continue;
          dsmem_bulk_copy(mapa_shared(sRed + (pb * NC + r) * kBlk, d), sStage + (pb * NC + d) * kBlk, kBlk, mapa_shared(bar, d));
        }
      }
      mbar_wait(bar, (s >> 1) & 1);
#pragma unroll
      for (int j = 0; j < UT; ++j) dh_rec[j] = 0.0f;
#pragma unroll
      for (uint32_t src = 0; src < (uint32_t)NC; ++src) {     // fixed order: the sum does not depend on arrival order
        const uint32_t blk = (src == r ? sStage : sRed) + (pb * NC + src) * kBlk + (uint32_t)row * (HU * 4) + (uint32_t)ul0 * 4;
#pragma unroll
        for (int q = 0; q < UT / 4; ++q) {
          const uint4 v = ld_shared_v4(blk + 16 * q);
          dh_rec[4 * q] += __uint_as_float(v.x); dh_rec[4 * q + 1] += __uint_as_float(v.y);
          dh_rec[4 * q + 2] += __uint_as_float(v.z); dh_rec[4 * q + 3] += __uint_as_float(v.w);
        }
      }
    }
    if (elected) bulk_wait_all0();
    // bias gradients: sum over the warp's 32 rows, one fixed-point atomic per (warp, gate unit)
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
      for (int j = 0; j < UT; ++j) {
        float a = bsum[g][j];
        for (int off = 16; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
        if (lane == 0 && a != 0.0f) {
          fx_add(P.g_b_ih + g * kH + u0 + j, a, kFxScale);
          fx_add(P.g_b_hh + g * kH + u0 + j, a, kFxScale);
        }
      }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc(tmem_base, 64);
}

}  // namespace lstmtc
}  // namespace urlg
