// Host side of the tcgen05 LSTM recurrence (lstm_tc.cuh): tensor maps, cluster launches, the W_hh^T image of the
// backward pass.  Called from lstm_kernels.cu (lstm_update_forward / lstm_update_backward) when the Memory fits the
// kernel's limits: hidden size 64, input width <= 64, split-bf16 operands.  URLG_LSTM_TC = 0 keeps the CUDA-core
// recurrence (default: clusters of 4 CTAs).
#include "common.cuh"
#include "policy.cuh"
#include "lstm_tc.cuh"

namespace urlg {
namespace tc {
int make_tmap_bf16(CUtensorMap* out, const void* ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner, int box_outer,
                   CUtensorMapSwizzle swizzle);
int make_tmap_bf16_3d(CUtensorMap* out, const void* ptr, int64_t inner, int64_t rows, int64_t steps, int64_t ld, int box_inner,
                      int box_rows);
}

using namespace urlg::lstmtc;

static int tc_mode() {
  static const int m = []() {
    const char* e = getenv("URLG_LSTM_TC");
    return (e && e[0] == '0') ? 0 : 4;
  }();
  return m;
}

// Cluster size of the tensor-core recurrence for this Memory, or 0 when it is outside the kernel's limits.
int lstm_tc_clusters(const urlg_policy* p, const LstmWs& w) {
  if (!p->lo_off || w.H != kH || w.in_pad > 64 || (w.kcat & 7) || (int64_t)p->R * kH * 4 < 4 * kH * kH * 4) return 0;
  return tc_mode();
}

template <int NC, typename Args>
static int launch_cluster(void (*kernel)(Args), int tiles, int smem, cudaStream_t st, const Args& a) {
  static bool attr_set = false;              // one instance per (NC, kernel): set the dynamic smem limit once
  if (!attr_set) {
    URLG_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(tiles * NC); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = NC; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  URLG_CUDA(cudaLaunchKernelEx(&cfg, kernel, a));
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

int lstm_tc_forward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int E, const float* saved_h,
                    const float* saved_c, cudaStream_t st) {
  LstmWs& w = p->lstm[k];
  NetWs& n = p->net[k];
  constexpr int NC = 4, HU = kH / NC;
  const int T = s->T;
  FwdArgs a = {};
  int rc = tc::make_tmap_bf16(&a.x_hi, w.xcat, w.in_pad, (int64_t)T * E, w.kcat, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16(&a.x_lo, w.xcat + p->lo_off, w.in_pad, (int64_t)T * E, w.kcat, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16(&a.wx_hi, w.w, w.in_pad, 4 * kH, w.kcat, 64, HU, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16(&a.wx_lo, w.w + p->lo_off, w.in_pad, 4 * kH, w.kcat, 64, HU, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16(&a.wh_hi, w.w + w.in_pad, kH, 4 * kH, w.kcat, 64, HU, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16(&a.wh_lo, w.w + p->lo_off + w.in_pad, kH, 4 * kH, w.kcat, 64, HU, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.xh_hi, w.xcat + w.in_pad, kH, E, T, w.kcat, 8, 128);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.xh_lo, w.xcat + p->lo_off + w.in_pad, kH, E, T, w.kcat, 8, 128);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.hx_hi, n.x, kH, E, T, n.kpad[0], 8, 128);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.hx_lo, n.x + p->lo_off, kH, E, T, n.kpad[0], 8, 128);
  if (rc) return rc;
  a.bias = w.bias;
  a.gates = w.gates; a.c_prev = w.c; a.tanh_c = w.c + (size_t)p->R * kH;
  a.dones = s->dones; a.saved_h = saved_h; a.saved_c = saved_c;
  a.T = T; a.E = E; a.N = s->N; a.env_begin = env_begin;
  a.kx_steps = (w.in_pad + 15) / 16;
  return launch_cluster<NC>(lstm_tc_fwd_kernel<NC>, (E + 127) / 128, Cfg<NC>::kFwdSmem, st, a);
}

int lstm_tc_backward(urlg_policy* p, int k, const urlg_storage* s, int env_begin, int E, cudaStream_t st) {
  LstmWs& w = p->lstm[k];
  constexpr int NC = 4, HU = kH / NC;
  const int T = s->T;
  // W_hh^T image (hi, lo) in the [R, H] fp32 scratch of the per-step path (64 KB, checked by lstm_tc_clusters)
  bf16* img = reinterpret_cast<bf16*>(w.dhrec);
  lstm_tc_wt_image_kernel<<<(4 * kH * kH + 255) / 256, 256, 0, st>>>(p->params + w.off[1], HU, img);
  URLG_LAUNCH_CHECK();
  BwdArgs a = {};
  int rc = tc::make_tmap_bf16(&a.w_hi, img, 4 * kH, kH, 4 * kH, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16(&a.w_lo, img + 4 * kH * kH, 4 * kH, kH, 4 * kH, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.dh_hi, w.dx_head, kH, E, T, kH, 8, 128);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.dh_lo, w.dx_head + p->lo_off, kH, E, T, kH, 8, 128);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.dg_hi, w.dgates, 4 * kH, E, T, 4 * kH, 8, 128);
  if (!rc) rc = tc::make_tmap_bf16_3d(&a.dg_lo, w.dgates + p->lo_off, 4 * kH, E, T, 4 * kH, 8, 128);
  if (rc) return rc;
  a.gates = w.gates; a.c_prev = w.c; a.tanh_c = w.c + (size_t)p->R * kH;
  a.dones = s->dones;
  a.g_b_ih = p->gacc + w.off[2]; a.g_b_hh = p->gacc + w.off[3];
  a.T = T; a.E = E; a.N = s->N; a.env_begin = env_begin;
  return launch_cluster<NC>(lstm_tc_bwd_kernel<NC>, (E + 127) / 128, Cfg<NC>::kBwdSmem, st, a);
}

}  // namespace urlg
