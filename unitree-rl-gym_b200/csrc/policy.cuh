// Internal layout of a urlg_policy: where the bf16 operand images and activation slabs live inside the
// caller-provided workspace.  Shared by policy_kernels.cu (rollout forward), ppo_kernels.cu (update) and
// lstm_kernels.cu (recurrent variants).
#pragma once
#include "common.cuh"
#include <cuda_bf16.h>

namespace urlg {

typedef __nv_bfloat16 bf16;

inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

struct NetWs {                       // one MLP (actor or critic)
  urlg_mlp_desc d;
  int kpad[URLG_MAX_LAYERS];         // padded input width of layer l (multiple of 8 -> 16-byte bf16 row pitch)
  bf16* w[URLG_MAX_LAYERS];          // [dims[l+1], kpad[l]] image of weight l (pad columns zero)
  bf16* x;                           // [R, kpad[0]] network input
  bf16* h[URLG_MAX_LAYERS];          // h[l], l=1..L-1: [R, dims[l]] post-activation output of layer l-1
  float* out;                        // [R, dims[L]] fp32 output of the last layer
  bf16* dout;                        // [R, opad] gradient wrt the output (opad = round_up(dims[L], 8))
  bf16* dz[URLG_MAX_LAYERS];         // dz[l], l=1..L-1: [R, dims[l]] gradient wrt the pre-activation of layer l-1
  int opad;
};

struct LstmWs {                      // one Memory (nn.LSTM, 1 layer)
  int in, in_pad, H, kcat;           // kcat = in_pad + H: gates = [x | h_prev] * [W_ih | W_hh]^T
  int64_t off[4];                    // w_ih, w_hh, b_ih, b_hh offsets in the flat parameter buffer
  bf16* w;                           // [4H, kcat] image
  float* bias;                       // [4H] = b_ih + b_hh
  bf16* xcat;                        // [T*R, kcat]   (rollout uses the first R rows)
  float* gates;                      // [T*R, 4H] activated gates (i,f,g,o) kept for the backward pass
  float* c;                          // [(T+1)*R, H] cell states (c[0] = initial)
  float* h0;                         // [R, H] initial hidden of the sequence (update mode)
  bf16* dgates;                      // [T*R, 4H]
  float* dh;                         // [R, H] running dL/dh
  float* dc;                         // [R, H] running dL/dc
  float* dhrec;                      // [R, H] dgates * W_hh (fp32 out of the recurrent dgrad)
  bf16* dx_head;                     // [T*R, H] dL/dh_t coming from the MLP head
};

}  // namespace urlg

namespace urlg { namespace fused { struct FusedArgs; } }

struct urlg_policy {
  urlg_policy_desc desc;
  float *params, *grads, *m, *v, *stats;
  urlg::NetWs net[2];                // 0 = actor, 1 = critic
  urlg::LstmWs lstm[2];
  int R;                             // max rows
  int T;                             // max sequence length (recurrent update)
  float* red;                        // small fp32 scratch (head reductions)
  int64_t lo_off;                    // split precision: element offset from a bf16 slab to its low-part twin (0 = off)
  int64_t last_rows;                 // rows of the pending minibatch gradient (set by minibatch_grads)
  // rollout: the actor and critic chains are independent until the sampling kernel, so the critic runs on a side
  // stream forked from / joined to the caller's stream by events (capturable: the side stream joins the graph)
  cudaStream_t side;
  cudaEvent_t ev_fork, ev_join;
  // fused multi-layer forward (mlp_fused.cuh): argument template with the weight / saved-activation tensor maps and the
  // per-network schedules, built once at create; NULL when the layer widths are outside its limits or URLG_FUSED=0
  urlg::fused::FusedArgs* fz;
  // order-independent gradient accumulation: [num_params + 4] 64-bit fixed-point accumulators (library-owned, zero
  // between minibatches); every gradient / loss-sum contribution of a minibatch is an integer atomic add into it, and
  // grads_finalize_kernel converts it into the flat fp32 gradient + the four per-minibatch sums and re-zeroes it
  long long* gacc;
  // split-K weight gradients of the MLPs: every split stores its fp32 partial tile into its own slab of wpart
  // (library-owned) and grads_finalize_kernel sums the slabs in split order -- as reproducible as the integer atomics,
  // without their cost (ncu: 30 % of the backward launches' stall samples sat on the 64-bit REDs).  wpart_off[net][l] =
  // float offset of layer l's slabs (sized for the largest split count at create), wpart_splits = splits of the
  // pending minibatch (0 = this layer accumulated into gacc instead)
  float* wpart;
  int64_t wpart_off[2][URLG_MAX_LAYERS];
  int wpart_cap[2][URLG_MAX_LAYERS];
  int wpart_splits[2][URLG_MAX_LAYERS];
  // data-parallel gradient exchange over peer memory (urlg_policy_bind_peer_exchange): when bound, grads_finalize_kernel
  // publishes [grads | 4 sums] into this rank's exchange half and urlg_ppo_apply's norm kernel sums the ranks' halves
  urlg_peer_exchange peer;
  int peer_bound;
  // single-rank updates: grads_finalize_kernel also produces the gradient norm + scalars (ppo_kernels.cu NormArgs);
  // one partial per block of that kernel (library-owned), norm_done = the pending minibatch's apply may skip its norm launch
  float* norm_partials;
  int norm_partials_cap;
  int norm_done;
};

namespace urlg {
// (hi, lo) split of an fp32 value over two bf16 slabs `lo_off` elements apart (lo_off == 0: plain bf16).
__device__ __forceinline__ void store_split(bf16* p, int64_t lo_off, size_t i, float v) {
  const bf16 h = __float2bfloat16(v);
  p[i] = h;
  if (lo_off) p[i + lo_off] = __float2bfloat16(v - __bfloat162float(h));
}
__device__ __forceinline__ float load_split(const bf16* p, int64_t lo_off, size_t i) {
  float v = __bfloat162float(p[i]);
  if (lo_off) v += __bfloat162float(p[i + lo_off]);
  return v;
}
// Fill the low-part operand pointers of a GEMM whose bf16 operands live in the policy workspace.
inline void with_lo(const urlg_policy* p, urlg_gemm_desc& g) {
  if (!p->lo_off) return;
  g.A_lo = static_cast<const bf16*>(g.A) + p->lo_off;
  g.B_lo = static_cast<const bf16*>(g.B) + p->lo_off;
  if (g.out_bf16) g.out_bf16_lo = static_cast<bf16*>(g.out_bf16) + p->lo_off;
  if (g.dact_ref) g.dact_ref_lo = static_cast<const bf16*>(g.dact_ref) + p->lo_off;
}
// URLG_ONE_STREAM=1 keeps the critic chain on the caller's stream (debugging aid).
inline cudaStream_t side_stream(const urlg_policy* p, cudaStream_t st) {
  static const bool one = []() { const char* e = getenv("URLG_ONE_STREAM"); return e && e[0] == '1'; }();
  return one ? st : p->side;
}
int sync_weights(urlg_policy* p, cudaStream_t st);
int mlp_forward(urlg_policy* p, NetWs& n, int64_t rows, cudaStream_t st, int max_ctas = 0);
int mlp_backward(urlg_policy* p, NetWs& n, int64_t rows, bf16* dx_out, cudaStream_t st);
int mlp_forward_pair(urlg_policy* p, NetWs& a, NetWs& c, int64_t rows, cudaStream_t st);
int mlp_backward_pair(urlg_policy* p, NetWs& a, NetWs& c, int64_t rows, bf16* dxa, bf16* dxc, cudaStream_t st);
// mlp_fused.cu
int fused_build(urlg_policy* p);
void fused_free(urlg_policy* p);
// rollout (PPO.act of a feed-forward policy): obs / critic obs rows -> actions, log-prob, values, transition slabs of step t
int fused_act(urlg_policy* p, const float* obs, const float* cobs, const urlg_storage* s, int t, const float* noise,
              float* actions_out, const int64_t* step_base, int64_t step_off, cudaStream_t st);
// update forward: rows gathered by `indices` from the storage's observation tensors; saves x and the hidden activations
// into the policy workspace (what the backward GEMMs read), outputs into net.out
int fused_update_forward(urlg_policy* p, const urlg_storage* s, const int64_t* indices, int64_t rows, cudaStream_t st);
// plain forward of one network on fp32 rows (act_inference / evaluate)
int fused_forward_one(urlg_policy* p, int which, const float* x, int64_t rows, float* out, cudaStream_t st);
int stage_obs(const float* src, int64_t rows, int cols, float* store, bf16* dst, int64_t lo_off, int kpad, int ld, cudaStream_t st);
}
