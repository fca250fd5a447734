// Host side of the fused multi-layer forward (mlp_fused.cuh): builds the argument template of a policy once (tensor
// maps of the weight images and of the saved-activation slabs, the MMA / drain schedules) and launches the kernel for
// the three callers: PPO.act (rollout), PPO.update's forward, and single-network inference.
#include "common.cuh"
#include "policy.cuh"
#include "mlp_fused.cuh"
#include <new>

namespace urlg {
namespace tc {
int make_tmap_bf16(CUtensorMap* out, const void* ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner, int box_outer,
                   CUtensorMapSwizzle swizzle);
bool timing_on();
int timing_begin(cudaStream_t st);
int timing_end(cudaStream_t st, double flops, double bytes);
}

using namespace urlg::fused;

static bool fused_enabled() {
  static const bool on = []() { const char* e = getenv("URLG_FUSED"); return !(e && e[0] == '0'); }();
  return on;
}

int fused_build(urlg_policy* p) {
  p->fz = nullptr;
  if (!fused_enabled() || p->desc.recurrent || !p->lo_off) return URLG_OK;       // split-bf16 feed-forward policies
  for (int q = 0; q < 2; ++q)
    if (!fusable(p->net[q].d.dims, p->net[q].d.num_layers)) return URLG_OK;
  FusedArgs* F = new (std::nothrow) FusedArgs();
  if (!F) return fail(URLG_ENOMEM, "fused_build: out of host memory");
  for (int q = 0; q < 2; ++q) {
    NetWs& w = p->net[q];
    NetArgs& n = F->net[q];
    const int L = w.d.num_layers;
    n.num_layers = L;
    for (int l = 0; l <= L; ++l) n.dims[l] = w.d.dims[l];
    schedule_for(n);
    if (n.n_steps > kMaxSteps || n.n_drains > kMaxDrains) { delete F; return URLG_OK; }
    for (int l = 0; l < L; ++l) {
      n.bias[l] = p->params + w.d.b_off[l];
      int rc = tc::make_tmap_bf16(&n.w_hi[l], w.w[l], w.d.dims[l], w.d.dims[l + 1], w.kpad[l], 64, n.w_rows[l],
                                  CU_TENSOR_MAP_SWIZZLE_128B);
      if (!rc) rc = tc::make_tmap_bf16(&n.w_lo[l], w.w[l] + p->lo_off, w.d.dims[l], w.d.dims[l + 1], w.kpad[l], 64,
                                       n.w_rows[l], CU_TENSOR_MAP_SWIZZLE_128B);
      // saved activations: [0] the network input image x [R, kpad0], [l] hidden activation h[l] [R, dims[l]]
      const bf16* sv = l == 0 ? w.x : w.h[l];
      const int ld = l == 0 ? w.kpad[0] : w.d.dims[l];
      const int inner = l == 0 ? w.kpad[0] : w.d.dims[l];
      if (!rc) rc = tc::make_tmap_bf16(&n.save_hi[l], sv, inner, p->R, ld, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
      if (!rc) rc = tc::make_tmap_bf16(&n.save_lo[l], sv + p->lo_off, inner, p->R, ld, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) { delete F; return rc; }
    }
    n.x = nullptr; n.gather = nullptr; n.x_store = nullptr; n.x_cols = w.d.dims[0]; n.out = nullptr;
  }
  F->act.enabled = 0;
  static bool attr_set = false;
  if (!attr_set) {
    URLG_CUDA(cudaFuncSetAttribute(mlp_fused_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  p->fz = F;
  return URLG_OK;
}

void fused_free(urlg_policy* p) {
  delete p->fz;
  p->fz = nullptr;
}

static int fused_launch(FusedArgs& A, int64_t rows, cudaStream_t st) {
  A.M = (int)rows;
  A.tiles_m = (int)((rows + 127) / 128);
  const int items = A.tiles_m * A.num_nets;
  const int grid = items < kNumSMs ? items : kNumSMs;
  const bool timed = tc::timing_on();
  if (timed) if (int rc = tc::timing_begin(st)) return rc;
  URLG_CUDA(launch_pdl(mlp_fused_fwd_kernel, dim3(grid), dim3(kThreads), (size_t)kSmemBytes, st, A));
  URLG_LAUNCH_CHECK();
  if (timed) {
    // algorithmic work of the launch: 2 M N K per layer; bytes = input rows + weights (hi+lo) + outputs (+ saved activations)
    double fl = 0.0, by = 0.0;
    for (int q = 0; q < A.num_nets; ++q) {
      const NetArgs& n = A.net[q];
      by += 4.0 * rows * n.x_cols * (n.x_store ? 2.0 : 1.0) + 4.0 * rows * n.dims[n.num_layers];
      for (int l = 0; l < n.num_layers; ++l) {
        fl += 2.0 * rows * n.dims[l] * n.dims[l + 1];
        by += 4.0 * n.dims[l] * n.dims[l + 1];
        if (A.save) by += 4.0 * rows * n.dims[l];
      }
    }
    if (int rc = tc::timing_end(st, fl, by)) return rc;
  }
  return URLG_OK;
}

int fused_act(urlg_policy* p, const float* obs, const float* cobs, const urlg_storage* s, int t, const float* noise,
              float* actions_out, const int64_t* step_base, int64_t step_off, cudaStream_t st) {
  FusedArgs A = *p->fz;                        // launch-local copy of the template (thread-safe; 11 KB)
  const urlg_policy_desc& d = p->desc;
  const int64_t N = s->N;
  A.num_nets = 2;
  A.save = 0;
  A.net[0].x = obs; A.net[0].x_store = s->observations + (size_t)t * N * s->num_obs;
  A.net[1].x = cobs;
  A.net[1].x_store = s->privileged_observations ? s->privileged_observations + (size_t)t * N * s->num_priv : nullptr;
  A.net[0].out = nullptr; A.net[1].out = nullptr;
  ActArgs& a = A.act;
  a.enabled = 1;
  a.std_param = p->params + d.std_off;
  a.noise = noise;
  const size_t row0 = (size_t)t * N;
  a.actions = s->actions + row0 * d.num_actions;
  a.mu = s->mu + row0 * d.num_actions;
  a.sigma = s->sigma + row0 * d.num_actions;
  a.log_prob = s->actions_log_prob + row0;
  a.values = s->values + row0;
  a.actions_out = actions_out;
  a.A = d.num_actions;
  a.seed_lo = d.seed_lo; a.seed_hi = d.seed_hi; a.env_offset = d.env_offset;
  a.step_base = step_base; a.step_off = step_off;
  return fused_launch(A, N, st);
}

int fused_update_forward(urlg_policy* p, const urlg_storage* s, const int64_t* indices, int64_t rows, cudaStream_t st) {
  FusedArgs A = *p->fz;
  A.num_nets = 2;
  A.save = 1;
  A.act.enabled = 0;
  A.net[0].x = s->observations; A.net[0].gather = indices; A.net[0].out = p->net[0].out;
  A.net[1].x = s->privileged_observations ? s->privileged_observations : s->observations;
  A.net[1].gather = indices; A.net[1].out = p->net[1].out;
  return fused_launch(A, rows, st);
}

int fused_forward_one(urlg_policy* p, int which, const float* x, int64_t rows, float* out, cudaStream_t st) {
  FusedArgs A = *p->fz;
  if (which == 1) A.net[0] = A.net[1];
  A.num_nets = 1;
  A.save = 0;
  A.act.enabled = 0;
  A.net[0].x = x; A.net[0].gather = nullptr; A.net[0].x_store = nullptr; A.net[0].out = out;
  return fused_launch(A, rows, st);
}

}  // namespace urlg
