// Fused multi-layer MLP forward on tcgen05 (K3 / the forward half of K7): ONE persistent kernel runs every Linear+ELU
// layer of the actor and of the critic for a 128-row tile, with the hidden activations never leaving the SM:
//
//   input rows (fp32, optionally gathered by index) --split to bf16 (hi, lo)--> X        [shared, SWIZZLE_128B K-major]
//   layer l:  D_l (TMEM, fp32) = act_{l-1} . W_l^T         weights streamed by TMA through a 3-stage ring
//   drain  :  D_l -> registers -> + bias -> ELU -> (hi, lo) bf16 -> 64-column chunk of the activation ring [shared]
//             each chunk is the K-block of layer l+1's MMAs the moment it is written (and, in update mode, leaves for
//             the saved-activation tensors by a TMA store straight out of the ring slot)
//   last   :  D_L -> + bias -> fp32 outputs; in rollout mode the actor's rows go on to Normal(mu, std) sampling, log-prob
//             and the transition slabs, the critic's to values[t] (what sample_store_kernel did in a separate launch)
//
// so a rollout policy forward (PPO.act: [UPSTREAM rsl_rl v1.0.2] ActorCritic.act / evaluate, shapes
// legged_robot_config.py:247-251) is one launch instead of stage + 4 grouped GEMMs + sample, and the update's forward one
// launch per minibatch that reads the gathered observation rows once and writes each saved activation once.
//
// Warp roles (19 warps): 0 = weight TMA producer, 1 = MMA issuer (one thread) + TMEM owner, 2..17 = drain / input
// staging (4 per TMEM lane quarter x 4 column groups of 16), 18 = activation saver (update mode).
// The order of MMAs and drains per tile is a small table built on the host (schedule_for): producer, issuer and
// drainers walk the same table, so any layer-width combination inside the limits below runs without new device code.
//
// Limits (checked on the host; anything else takes the per-layer GEMM path): <= 4 layers, input width <= 64,
// first hidden width <= 512, other widths <= 256, hidden widths multiples of 16, output width <= 16.
// TMEM: columns [0,256) hold even layers (layer 0 in two alternating 128-column halves when wider than 256),
// columns [256,512) odd layers.  Operand precision: split bf16 (hi + lo, 3 MMAs per product), like tc_gemm.cuh.
#pragma once
#include <cuda.h>
#include "tc_ptx.cuh"
#include "tc_gemm.cuh"
#include "philox.cuh"

namespace urlg {
namespace fused {

using namespace urlg::tc;

constexpr int kMaxLayers = 4, kMaxSteps = 48, kMaxDrains = 24;
constexpr int kActSlots = 3, kWStages = 3;
constexpr int kTileBytes = 128 * 64 * 2;           // one 128-row x 64-column bf16 tile (16 KB)
constexpr int kSlotBytes = 2 * kTileBytes;         // hi + lo
constexpr int kDrainWarps = 16;
constexpr int kThreads = (2 + kDrainWarps + 1) * 32;
constexpr int kNumTfull = 2 + kMaxLayers;          // tfull[0,1]: alternating halves of a chunked layer 0; tfull[2+l]: layer l
constexpr int kSmemBytes = (1 + kActSlots + kWStages) * kSlotBytes + 1024 + 512;

enum : uint8_t { kStepFirstK = 1, kStepAisX = 2, kStepAFirstUse = 4, kStepALastUse = 8, kStepDone = 16 };
enum : uint8_t { kDrainWait = 1, kDrainFinal = 2 };

struct MmaStep {
  uint8_t layer, nt, kb, flags;
  uint16_t tmem_col;      // accumulator column of this (layer, nt)
  uint8_t n16;            // MMA N / 16
  uint8_t done_bar;       // tfull index signalled when flags & kStepDone
};
struct DrainStep {
  uint8_t layer, bar, flags, pad;
  uint16_t tmem_col;      // first of the 64 accumulator columns of this chunk
  uint16_t out_col;       // column of the layer's output the chunk starts at
};

struct NetArgs {
  CUtensorMap w_hi[kMaxLayers], w_lo[kMaxLayers];          // weight images [N_l, kpad_l], K-major, box {64, w_rows[l]}
  CUtensorMap save_hi[kMaxLayers], save_lo[kMaxLayers];    // update mode: [0] = network input image, [l] = hidden act l
  const float* bias[kMaxLayers];
  int dims[kMaxLayers + 1];
  int w_rows[kMaxLayers];
  int num_layers, n_steps, n_drains;
  MmaStep steps[kMaxSteps];
  DrainStep drains[kMaxDrains];
  const float* x;            // fp32 input rows [*, x_cols]; row r of the tile space reads row (gather ? gather[r] : r)
  const int64_t* gather;
  float* x_store;            // optional fp32 copy of the input rows ([M, x_cols]: the rollout's observation slab)
  int x_cols;
  float* out;                // fp32 [M, dims[L]] (ld = dims[L]) or NULL
};

// Rollout epilogue of PPO.act: Normal(mu, std).sample() / log_prob / transition slabs (actor rows), values (critic rows)
struct ActArgs {
  int enabled;
  const float* std_param;    // [A]
  const float* noise;        // optional injected N(0,1) [M, A]
  float *actions, *mu, *sigma, *log_prob, *values;   // storage slabs of step t: [M, A] x3, [M], [M]
  float* actions_out;        // [M, A]
  int A;
  uint32_t seed_lo, seed_hi;
  int64_t env_offset;
  const int64_t* step_base;
  int64_t step_off;
};

struct FusedArgs {
  NetArgs net[2];
  ActArgs act;
  int num_nets, M, tiles_m, save;
};

// ---------------------------------------------------------------- host: schedule
inline int pad16(int x) { return (x + 15) / 16 * 16; }
inline bool fusable(const int* dims, int L) {
  if (L < 1 || L > kMaxLayers) return false;
  if (dims[0] > 64 || dims[L] > 16) return false;
  for (int l = 1; l < L; ++l) {
    if (dims[l] % 16) return false;
    if (dims[l] > (l == 1 ? 512 : 256)) return false;
  }
  return true;
}
// Fills steps / drains / w_rows of a network with dims[0..L].
inline void schedule_for(NetArgs& n) {
  const int L = n.num_layers;
  int NT[kMaxLayers], NKB[kMaxLayers], region[kMaxLayers];
  for (int l = 0; l < L; ++l) {
    const int N = n.dims[l + 1];
    NT[l] = (N + 127) / 128;
    NKB[l] = l == 0 ? 1 : (n.dims[l] + 63) / 64;
    region[l] = (l & 1) ? 256 : 0;
    n.w_rows[l] = N >= 128 ? 128 : pad16(N);
  }
  const bool chunked = NT[0] > 2;                   // layer 0 wider than its 256-column region: alternating halves
  int ns = 0, nd = 0;
  auto step = [&](int l, int nt, int kb, bool a_first, bool a_last) {
    MmaStep& s = n.steps[ns++];
    s.layer = (uint8_t)l; s.nt = (uint8_t)nt; s.kb = (uint8_t)kb;
    const int N = n.dims[l + 1];
    const int ncols = (nt + 1) * 128 <= N ? 128 : pad16(N - nt * 128);
    s.n16 = (uint8_t)(ncols / 16);
    s.tmem_col = (uint16_t)((l == 0 && chunked) ? (nt & 1) * 128 : region[l] + nt * 128);
    s.flags = (kb == 0 ? kStepFirstK : 0) | (l == 0 ? kStepAisX : 0) | (a_first ? kStepAFirstUse : 0) |
              (a_last ? kStepALastUse : 0);
    s.done_bar = 0;
    if (l == 0 && chunked) { s.flags |= kStepDone; s.done_bar = (uint8_t)(nt & 1); }
    else if (kb == NKB[l] - 1 && nt == NT[l] - 1) { s.flags |= kStepDone; s.done_bar = (uint8_t)(2 + l); }
  };
  // layer 0: X is its only K block; when chunked, tiles 2.. are interleaved into layer 1 (one per two K blocks of layer 1:
  // by then the half it overwrites has been drained, which the issuer knows from the activation chunks it waited for)
  const int l0_head = chunked ? 2 : NT[0];
  for (int nt = 0; nt < l0_head; ++nt) step(0, nt, 0, nt == 0, nt == NT[0] - 1);
  int l0_next = l0_head;
  for (int l = 1; l < L; ++l) {
    for (int kb = 0; kb < NKB[l]; ++kb) {
      for (int nt = 0; nt < NT[l]; ++nt) step(l, nt, kb, nt == 0, nt == NT[l] - 1);
      if (l == 1 && (kb & 1) && l0_next < NT[0]) { step(0, l0_next, 0, false, l0_next == NT[0] - 1); ++l0_next; }
    }
  }
  n.n_steps = ns;
  auto drain = [&](int l, int bar, int tcol, int ocol, bool wait, bool fin) {
    DrainStep& d = n.drains[nd++];
    d.layer = (uint8_t)l; d.bar = (uint8_t)bar; d.flags = (wait ? kDrainWait : 0) | (fin ? kDrainFinal : 0); d.pad = 0;
    d.tmem_col = (uint16_t)tcol; d.out_col = (uint16_t)ocol;
  };
  for (int l = 0; l < L; ++l) {
    const int N = n.dims[l + 1];
    if (l == L - 1) { drain(l, 2 + l, region[l], 0, true, true); break; }
    if (l == 0 && chunked) {
      for (int nt = 0; nt < NT[0]; ++nt)
        for (int h = 0; h < 2 && nt * 128 + h * 64 < N; ++h) drain(0, nt & 1, (nt & 1) * 128 + h * 64, nt * 128 + h * 64, h == 0, false);
    } else {
      for (int j = 0; j * 64 < N; ++j) drain(l, 2 + l, region[l] + j * 64, j * 64, j == 0, false);
    }
  }
  n.n_drains = nd;
}

// ---------------------------------------------------------------- device
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

__global__ void __launch_bounds__(kThreads, 1) mlp_fused_fwd_kernel(const __grid_constant__ FusedArgs P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t sX = base, sAct = sX + kSlotBytes, sW = sAct + kActSlots * kSlotBytes;
  const uint32_t bars = sW + kWStages * kSlotBytes;
  const uint32_t wfull0 = bars, wempty0 = wfull0 + 8 * kWStages, afull0 = wempty0 + 8 * kWStages;
  const uint32_t aempty0 = afull0 + 8 * kActSlots, xfull = aempty0 + 8 * kActSlots, xempty = xfull + 8;
  const uint32_t tfull0 = xempty + 8, tmem_slot = tfull0 + 8 * kNumTfull;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_items = P.tiles_m * P.num_nets;     // item = (tile, net): net-major so one CTA alternates as little as possible

  if (warp == 0 && lane == 0) {
    for (int q = 0; q < P.num_nets; ++q)
      for (int l = 0; l < P.net[q].num_layers; ++l) {
        tma_prefetch_desc(&P.net[q].w_hi[l]);
        tma_prefetch_desc(&P.net[q].w_lo[l]);
      }
    for (int s = 0; s < kWStages; ++s) { mbar_init(wfull0 + 8 * s, 1); mbar_init(wempty0 + 8 * s, 1); }
    for (int s = 0; s < kActSlots; ++s) { mbar_init(afull0 + 8 * s, 4); mbar_init(aempty0 + 8 * s, P.save ? 2 : 1); }
    mbar_init(xfull, kDrainWarps);
    mbar_init(xempty, P.save ? 2 : 1);
    for (int b = 0; b < kNumTfull; ++b) mbar_init(tfull0 + 8 * b, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0) {
    // ===== weight producer: one (layer, n-tile, k-block) tile of W (hi + lo) per MMA step, in schedule order =====
    if (lane == 0) {
      uint32_t wc = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const NetArgs& n = P.net[item / P.tiles_m];
        for (int i = 0; i < n.n_steps; ++i, ++wc) {
          const MmaStep st = n.steps[i];
          const uint32_t s = wc % kWStages, ph = (wc / kWStages) & 1;
          mbar_wait(wempty0 + 8 * s, ph ^ 1);
          const uint32_t full = wfull0 + 8 * s;
          const uint32_t rows = (uint32_t)n.w_rows[st.layer];       // TMA always delivers the whole box (OOB rows = 0)
          mbar_arrive_expect_tx(full, 2u * rows * 128u);
          const uint32_t dst = sW + s * kSlotBytes;
          tma_load_2d(dst, &n.w_hi[st.layer], full, st.kb * 64, st.nt * 128);
          tma_load_2d(dst + kTileBytes, &n.w_lo[st.layer], full, st.kb * 64, st.nt * 128);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      uint32_t wc = 0, ac = 0, it = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
        const NetArgs& n = P.net[item / P.tiles_m];
        for (int i = 0; i < n.n_steps; ++i, ++wc) {
          const MmaStep st = n.steps[i];
          const bool is_x = st.flags & kStepAisX;
          const uint32_t slot = ac % kActSlots;
          if (st.flags & kStepAFirstUse) {
            if (is_x) mbar_wait(xfull, it & 1);
            else mbar_wait(afull0 + 8 * slot, (ac / kActSlots) & 1);
            tc_fence_after();
          }
          const uint32_t s = wc % kWStages;
          mbar_wait(wfull0 + 8 * s, (wc / kWStages) & 1);
          tc_fence_after();
          const uint32_t a_hi = is_x ? sX : sAct + slot * kSlotBytes, a_lo = a_hi + kTileBytes;
          const uint32_t b_hi = sW + s * kSlotBytes, b_lo = b_hi + kTileBytes;
          const uint32_t idesc = make_idesc_bf16(128, st.n16 * 16, 0, 0);
          const uint32_t d = tmem_base + st.tmem_col;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t ad = make_sdesc_sw128(a_hi + k * 32, 16, 1024), bd = make_sdesc_sw128(b_hi + k * 32, 16, 1024);
            const uint64_t ald = make_sdesc_sw128(a_lo + k * 32, 16, 1024), bld = make_sdesc_sw128(b_lo + k * 32, 16, 1024);
            umma_bf16(d, ad, bd, idesc, ((st.flags & kStepFirstK) && k == 0) ? 0u : 1u);
            umma_bf16(d, ald, bd, idesc, 1u);
            umma_bf16(d, ad, bld, idesc, 1u);
          }
          umma_commit(wempty0 + 8 * s);
          if (st.flags & kStepALastUse) {
            if (is_x) umma_commit(xempty);
            else { umma_commit(aempty0 + 8 * slot); ++ac; }
          }
          if (st.flags & kStepDone) umma_commit(tfull0 + 8 * st.done_bar);
        }
      }
    }
  } else if (warp < 2 + kDrainWarps) {
    // ===== input staging + accumulator drains =====
    const int dw = warp - 2;
    const int lq = warp & 3;                 // TMEM lane quarter this warp may read (hardware rule: warp id % 4)
    const int cg = dw >> 2;                  // warp group 0..3: one warp per lane quarter; chunks are dealt round-robin
    const int row_in_tile = lq * 32 + lane;  // accumulator row held by this thread
    uint32_t pc = 0, it = 0, tuse = 0;       // tuse: 4 bits per tfull barrier = uses so far (phase)
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
      const int q = item / P.tiles_m, tile = item - q * P.tiles_m;
      const NetArgs& n = P.net[q];
      const int m0 = tile * 128;
      // ---- X: 128 input rows -> (hi, lo) bf16, SWIZZLE_128B K-major; thread = (row, 16-column quarter) ----
      {
        const int t = dw * 32 + lane, r = t >> 2, cq = t & 3;
        const int row = m0 + r;
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0.0f;
        if (row < P.M) {
          const int64_t src = n.gather ? n.gather[row] : (int64_t)row;
          const float* xp = n.x + src * n.x_cols + 16 * cq;
          float* sp = n.x_store ? n.x_store + (int64_t)row * n.x_cols + 16 * cq : nullptr;
          const int nc = min(16, n.x_cols - 16 * cq);
          if (nc == 16 && (n.x_cols & 3) == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 f = __ldg(reinterpret_cast<const float4*>(xp) + i);
              v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w;
              if (sp) reinterpret_cast<float4*>(sp)[i] = f;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) if (i < nc) { v[i] = xp[i]; if (sp) sp[i] = v[i]; }
          }
        }
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          hi[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
          lo[i] = pack_bf16x2(v[2 * i] - bf16lo(hi[i]), v[2 * i + 1] - bf16hi(hi[i]));
        }
        mbar_wait(xempty, (it & 1) ^ 1);            // the previous item's layer-0 MMAs (and saver) are done with X
        const uint32_t o0 = sw128_offset(r, 2 * cq), o1 = sw128_offset(r, 2 * cq + 1);
        st_shared_v4(sX + o0, hi[0], hi[1], hi[2], hi[3]);
        st_shared_v4(sX + o1, hi[4], hi[5], hi[6], hi[7]);
        st_shared_v4(sX + kTileBytes + o0, lo[0], lo[1], lo[2], lo[3]);
        st_shared_v4(sX + kTileBytes + o1, lo[4], lo[5], lo[6], lo[7]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(xfull);
      }
      // ---- drains: the four warp groups (one warp per TMEM lane quarter each) take the 64-column chunks round-robin, so up
      //      to four chunks are in flight; a group waits for an accumulator the first time it reads from it ----
      uint32_t cur_bar = 0, cur_par = 0;
      bool waited = true;
      for (int i = 0; i < n.n_drains; ++i) {
        const DrainStep ds = n.drains[i];
        if (ds.flags & kDrainWait) {
          const uint32_t uses = (tuse >> (4 * ds.bar)) & 15u;
          cur_bar = tfull0 + 8 * ds.bar;
          cur_par = uses & 1u;
          waited = false;
          tuse += 1u << (4 * ds.bar);
          if (((tuse >> (4 * ds.bar)) & 15u) == 14u) tuse -= 14u << (4 * ds.bar);   // keep the nibble (parity preserved)
        }
        const int N = n.dims[ds.layer + 1];
        const float* bias = n.bias[ds.layer];
        if (!(ds.flags & kDrainFinal)) {
          if ((int)(pc & 3u) == cg) {
            if (!waited) { mbar_wait(cur_bar, cur_par); tc_fence_after(); waited = true; }
            const uint32_t slot = pc % kActSlots;
            const uint32_t sb = sAct + slot * kSlotBytes;
            const uint32_t taddr = tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)ds.tmem_col;
            bool slot_ready = false;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t ra[16], rb[16];
              tmem_ld16_issue(taddr + 32 * h, ra);
              tmem_ld16_issue(taddr + 32 * h + 16, rb);
              tmem_ld_wait_dep(ra);
              tmem_ld_wait_dep(rb);
#pragma unroll
              for (int g2 = 0; g2 < 2; ++g2) {
                const int cgi = 2 * h + g2;                              // 16-column group of the chunk
                const int col0 = ds.out_col + 16 * cgi;                  // column of the layer output
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(g2 ? rb[j] : ra[j]);
                if (col0 < N) {                                          // N is a multiple of 16: whole group valid
                  if (((uintptr_t)(bias + col0) & 15) == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                      const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + col0) + j);
                      v[4 * j] += b4.x; v[4 * j + 1] += b4.y; v[4 * j + 2] += b4.z; v[4 * j + 3] += b4.w;
                    }
                  } else {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += __ldg(bias + col0 + j);
                  }
#pragma unroll
                  for (int j = 0; j < 16; ++j) v[j] = elu1(v[j]);
                } else {
#pragma unroll
                  for (int j = 0; j < 16; ++j) v[j] = 0.0f;              // K padding of the next layer
                }
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  hi[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
                  lo[j] = pack_bf16x2(v[2 * j] - bf16lo(hi[j]), v[2 * j + 1] - bf16hi(hi[j]));
                }
                if (!slot_ready) {                                       // consumer MMAs (and saver) released the slot
                  mbar_wait(aempty0 + 8 * slot, ((pc / kActSlots) & 1) ^ 1);
                  slot_ready = true;
                }
                const uint32_t o0 = sw128_offset(row_in_tile, 2 * cgi), o1 = sw128_offset(row_in_tile, 2 * cgi + 1);
                st_shared_v4(sb + o0, hi[0], hi[1], hi[2], hi[3]);
                st_shared_v4(sb + o1, hi[4], hi[5], hi[6], hi[7]);
                st_shared_v4(sb + kTileBytes + o0, lo[0], lo[1], lo[2], lo[3]);
                st_shared_v4(sb + kTileBytes + o1, lo[4], lo[5], lo[6], lo[7]);
              }
            }
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(afull0 + 8 * slot);
          }
          ++pc;
        } else if (cg == 0) {
          // ---- last layer: <= 16 outputs per row; thread = row (group 0) ----
          if (!waited) { mbar_wait(cur_bar, cur_par); tc_fence_after(); waited = true; }
          float v[16];
          tmem_ld16(tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)ds.tmem_col, v);
          const int row = m0 + row_in_tile;
          if (row < P.M) {
#pragma unroll
            for (int j = 0; j < 16; ++j) if (j < N) v[j] += bias[j];
            if (n.out) {
              float* o = n.out + (int64_t)row * N;
#pragma unroll
              for (int j = 0; j < 16; ++j) if (j < N) o[j] = v[j];
            }
            if (P.act.enabled) {
              const ActArgs& a = P.act;
              if (q == 1) {
                a.values[row] = v[0];
              } else {
                // Normal(mu, std).sample() + log_prob, term by term like torch.distributions.Normal ([UPSTREAM] PPO.act)
                const int A = a.A;
                SlotRng rng;
                if (!a.noise) rng.init(a.seed_lo, a.seed_hi, kStreamPolicy, a.env_offset + row, eff_step(a.step_base, a.step_off), nullptr);
                float lp = 0.0f;
                float act[16], sg[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  act[j] = 0.0f; sg[j] = 0.0f;
                  if (j < A) {
                    const float mu = v[j];
                    const float sd = mu * 0.0f + a.std_param[j];
                    float eps;
                    if (a.noise) {
                      eps = a.noise[(int64_t)row * A + j];
                    } else {
                      const float u1 = 1.0f - rng.uniform(2 * j), u2 = rng.uniform(2 * j + 1);   // u1 in (0,1]
                      eps = sqrtf(-2.0f * logf(u1)) * cosf(6.28318530717958647692f * u2);
                    }
                    const float ac = mu + sd * eps;
                    const float d = ac - mu;
                    lp += -(d * d) / (2.0f * (sd * sd)) - logf(sd) - 0.918938533204672741780f;
                    act[j] = ac; sg[j] = sd;
                  }
                }
                float* pa = a.actions + (int64_t)row * A;
                float* pm = a.mu + (int64_t)row * A;
                float* ps = a.sigma + (int64_t)row * A;
                float* po = a.actions_out + (int64_t)row * A;
#pragma unroll
                for (int j = 0; j < 16; ++j) if (j < A) { pa[j] = act[j]; pm[j] = v[j]; ps[j] = sg[j]; po[j] = act[j]; }
                a.log_prob[row] = lp;
              }
            }
          }
          tc_fence_before();
        }
      }
    }
  } else {
    // ===== activation saver (update mode): TMA tensor stores straight out of X / the ring slots =====
    if (lane == 0 && P.save) {
      uint32_t pc = 0, it = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
        const int q = item / P.tiles_m, tile = item - q * P.tiles_m;
        const NetArgs& n = P.net[q];
        const int m0 = tile * 128;
        mbar_wait(xfull, it & 1);
        tma_store_2d(&n.save_hi[0], sX, 0, m0);
        tma_store_2d(&n.save_lo[0], sX + kTileBytes, 0, m0);
        bulk_commit();
        bulk_wait_read0();
        mbar_arrive(xempty);
        for (int i = 0; i < n.n_drains; ++i) {
          const DrainStep ds = n.drains[i];
          if (ds.flags & kDrainFinal) continue;
          const uint32_t slot = pc % kActSlots;
          mbar_wait(afull0 + 8 * slot, (pc / kActSlots) & 1);
          const uint32_t sb = sAct + slot * kSlotBytes;
          tma_store_2d(&n.save_hi[ds.layer + 1], sb, ds.out_col, m0);
          tma_store_2d(&n.save_lo[ds.layer + 1], sb + kTileBytes, ds.out_col, m0);
          bulk_commit();
          bulk_wait_read0();
          mbar_arrive(aempty0 + 8 * slot);
          ++pc;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace fused
}  // namespace urlg
