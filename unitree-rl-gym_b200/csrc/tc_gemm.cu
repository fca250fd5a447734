// Host side of the tcgen05 GEMM: tensor-map construction (driver entry point resolved at run time, so
// the library still loads on a box without libcuda) and the launch.
#include "common.cuh"
#include "tc_gemm.cuh"
#include <cudaTypedefs.h>
#include <cstdlib>
#include <utility>
#include <vector>

namespace urlg {
namespace tc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 row-major matrix [outer, inner] with row pitch ld (elements); box {box_inner, box_outer}; zero OOB fill.
int make_tmap_bf16(CUtensorMap* out, const void* ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner, int box_outer,
                   CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(URLG_ECUDA, "cuTensorMapEncodeTiled entry point unavailable");
  if (((uintptr_t)ptr & 15) || ((ld * 2) & 15))
    return fail(URLG_EINVAL, "tensor map: base must be 16-byte aligned and row pitch a multiple of 16 bytes (ld=%lld)", (long long)ld);
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(URLG_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return URLG_OK;
}

// bf16 tensor [steps, rows, inner] (row pitch ld elements, step pitch rows*ld): box {box_inner, box_rows, 1}, no swizzle.
// The LSTM recurrence moves per-step [128 rows, 8 columns] core-matrix columns with it; rows past `rows` are clipped
// (stores) / zero-filled (loads) per step, so a ragged last tile never touches the next step's rows.
int make_tmap_bf16_3d(CUtensorMap* out, const void* ptr, int64_t inner, int64_t rows, int64_t steps, int64_t ld, int box_inner,
                      int box_rows) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(URLG_ECUDA, "cuTensorMapEncodeTiled entry point unavailable");
  if (((uintptr_t)ptr & 15) || ((ld * 2) & 15))
    return fail(URLG_EINVAL, "tensor map: base must be 16-byte aligned and row pitch a multiple of 16 bytes (ld=%lld)", (long long)ld);
  cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)steps};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)rows * (cuuint64_t)ld * 2};
  cuuint32_t box[3] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(URLG_ECUDA, "cuTensorMapEncodeTiled (3-D) failed with CUresult %d", (int)r);
  return URLG_OK;
}

// Optional per-launch timing (bench.py's roofline leg): CUDA events recorded on the launching stream around every
// tensor-core launch (GEMM or fused forward) while enabled -- never under graph capture -- summed on read.  One launch
// per bracket, in situ: the stream stays busy (the host enqueues faster than these kernels run), so a bracket is the
// kernel plus its inter-kernel gap with the caches in the state the real step leaves them.
static bool g_timing = false;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_events;
static double g_flops = 0.0, g_bytes = 0.0;
static size_t g_used = 0;

int launch_group_impl(const urlg_gemm_desc* const* ds, int n, cudaStream_t st);

bool timing_on() { return g_timing; }
int timing_begin(cudaStream_t st) {
  if (g_used == g_events.size()) {
    cudaEvent_t a, b;
    URLG_CUDA(cudaEventCreate(&a));
    URLG_CUDA(cudaEventCreate(&b));
    g_events.emplace_back(a, b);
  }
  URLG_CUDA(cudaEventRecord(g_events[g_used].first, st));
  return URLG_OK;
}
int timing_end(cudaStream_t st, double flops, double bytes) {
  URLG_CUDA(cudaEventRecord(g_events[g_used].second, st));
  ++g_used;
  g_flops += flops;
  g_bytes += bytes;
  return URLG_OK;
}

// n independent problems in one launch (n <= kMaxGroup); see GroupArgs in tc_gemm.cuh.
int launch_gemm_group(const urlg_gemm_desc* const* ds, int n, cudaStream_t st) {
  if (!g_timing) return launch_group_impl(ds, n, st);
  if (int rc = timing_begin(st)) return rc;
  const int rc = launch_group_impl(ds, n, st);
  double fl = 0.0, by = 0.0;
  for (int i = 0; ds && i < n; ++i) {
    const urlg_gemm_desc* d = ds[i];
    if (!d) continue;
    fl += 2.0 * d->M * d->N * d->K;
    // algorithmic HBM bytes of the launch: each operand element and each output element once (bf16, x2 when split into hi+lo)
    const double parts = (d->A_lo && d->B_lo) ? 2.0 : 1.0, mn = (double)d->M * d->N;
    by += 2.0 * parts * ((double)d->M * d->K + (double)d->N * d->K) + (d->out_f32 ? 4.0 * mn : 0.0) +
          (d->out_bf16 ? 2.0 * parts * mn : 0.0) + (d->dact_ref ? 2.0 * parts * mn : 0.0);
  }
  if (int rc2 = timing_end(st, fl, by)) return rc2;
  return rc;
}

int launch_gemm(const urlg_gemm_desc* d, cudaStream_t st) { return launch_gemm_group(&d, 1, st); }

// K splits cover whole k-blocks: the count actually used for a requested k_splits (and the K range of one split).
int effective_splits(int K, int k_splits, int* k_per_split) {
  int splits = k_splits > 0 ? k_splits : 1;
  const int kps = ((K + splits - 1) / splits + kBlockK - 1) / kBlockK * kBlockK;
  if (k_per_split) *k_per_split = kps;
  return (K + kps - 1) / kps;
}

// Fills one problem's kernel arguments and its 8 tensor maps.
static int build_problem(const urlg_gemm_desc* d, GemmArgs& g, CUtensorMap* tm, int group_n) {
  if (!d || !d->A || !d->B || d->M <= 0 || d->N <= 0 || d->K <= 0) return fail(URLG_EINVAL, "urlg_tc_gemm: bad argument");
  CUtensorMap &ta = tm[kTmA], &tb = tm[kTmB], &tal = tm[kTmAlo], &tbl = tm[kTmBlo], &to = tm[kTmO], &tol = tm[kTmOlo],
              &tr = tm[kTmR], &trl = tm[kTmRlo];
  g.M = d->M; g.N = d->N; g.K = d->K;
  g.a_mn = d->a_mn ? 1 : 0; g.b_mn = d->b_mn ? 1 : 0;
  int bn = d->block_n;
  if (bn <= 0) {
    // auto: the widest tile that still gives every SM a tile (the epilogue, not the MMA, is the cost here)
    bn = g.b_mn ? ((d->N + 63) / 64) * 64 : ((d->N + 15) / 16) * 16;
    if (bn > 128) bn = 128;
    const int tm = (d->M + kBlockM - 1) / kBlockM;
    const int min_bn = g.b_mn ? 64 : 32;
    while (bn > min_bn && (bn % 2 == 0) && ((bn / 2) % (g.b_mn ? 64 : 16) == 0) &&
           tm * ((d->N + bn - 1) / bn) * group_n < (kNumSMs * 3) / 4)     // the group's problems share the SMs
      bn /= 2;
  }
  if (bn > 256) bn = 256;
  if ((bn & 15) || (g.b_mn && (bn & 63))) return fail(URLG_EINVAL, "urlg_tc_gemm: block_n=%d invalid", bn);
  g.block_n = bn;
  int kps = 0;
  const int splits = effective_splits(d->K, d->k_splits, &kps);
  g.k_per_split = kps;
  g.out_part = d->out_part;
  if (splits > 1 && !(d->flags & kFlagAtomic) && !d->out_part) return fail(URLG_EINVAL, "urlg_tc_gemm: k_splits > 1 needs the atomic flag");
  g.out_f32 = d->out_f32; g.ld_f32 = d->ld_f32;
  g.out_bf16 = reinterpret_cast<__nv_bfloat16*>(d->out_bf16); g.ld_bf16 = d->ld_bf16;
  g.bias = d->bias;
  g.dact_ref = reinterpret_cast<const __nv_bfloat16*>(d->dact_ref); g.ld_ref = d->ld_ref;
  g.colsum = d->colsum;
  g.out_fx = reinterpret_cast<long long*>(d->out_fx);
  g.colsum_fx = reinterpret_cast<long long*>(d->colsum_fx);
  g.flags = d->flags;
  g.split = (d->A_lo && d->B_lo) ? 1 : 0;
  if ((d->A_lo != nullptr) != (d->B_lo != nullptr)) return fail(URLG_EINVAL, "urlg_tc_gemm: A_lo and B_lo go together");
  if (g.split && bn > 128) { bn = 128; g.block_n = bn; }       // two operand sets per stage: keep >= 2-3 stages in 227 KB
  g.out_lo = reinterpret_cast<__nv_bfloat16*>(d->out_bf16_lo);
  g.dact_lo = reinterpret_cast<const __nv_bfloat16*>(d->dact_ref_lo);
  // bf16 outputs go through shared staging + TMA tensor stores whenever a tensor map can describe them
  g.tma_out = (g.out_bf16 && (((uintptr_t)g.out_bf16 | (uintptr_t)g.out_lo) & 15) == 0 && (g.ld_bf16 & 7) == 0) ? 1 : 0;
  g.tma_ref = (g.tma_out && g.dact_ref && (((uintptr_t)g.dact_ref | (uintptr_t)g.dact_lo) & 15) == 0 && (g.ld_ref & 7) == 0) ? 1 : 0;
  g.stages = gemm_stages_for(bn, g.split, g.tma_out, g.tma_ref);
  int rc;
  if (g.tma_out) {
    rc = make_tmap_bf16(&to, g.out_bf16, d->N, d->M, g.ld_bf16, 16, 32, CU_TENSOR_MAP_SWIZZLE_32B);
    if (rc) return rc;
    if (g.out_lo) {
      rc = make_tmap_bf16(&tol, g.out_lo, d->N, d->M, g.ld_bf16, 16, 32, CU_TENSOR_MAP_SWIZZLE_32B);
      if (rc) return rc;
    } else {
      tol = to;
    }
  }
  if (!g.a_mn) rc = make_tmap_bf16(&ta, d->A, d->K, d->M, d->lda, kBlockK, kBlockM);
  else rc = make_tmap_bf16(&ta, d->A, d->M, d->K, d->lda, 64, kBlockK);
  if (rc) return rc;
  if (!g.b_mn) rc = make_tmap_bf16(&tb, d->B, d->K, d->N, d->ldb, kBlockK, bn);
  else rc = make_tmap_bf16(&tb, d->B, d->N, d->K, d->ldb, 64, kBlockK);
  if (rc) return rc;
  if (g.split) {
    if (!g.a_mn) rc = make_tmap_bf16(&tal, d->A_lo, d->K, d->M, d->lda, kBlockK, kBlockM);
    else rc = make_tmap_bf16(&tal, d->A_lo, d->M, d->K, d->lda, 64, kBlockK);
    if (rc) return rc;
    if (!g.b_mn) rc = make_tmap_bf16(&tbl, d->B_lo, d->K, d->N, d->ldb, kBlockK, bn);
    else rc = make_tmap_bf16(&tbl, d->B_lo, d->N, d->K, d->ldb, 64, kBlockK);
    if (rc) return rc;
  } else {
    tal = ta; tbl = tb;
  }
  if (!g.tma_out) { to = ta; tol = ta; }
  if (g.tma_ref) {
    rc = make_tmap_bf16(&tr, g.dact_ref, d->N, d->M, g.ld_ref, 16, 32, CU_TENSOR_MAP_SWIZZLE_32B);
    if (rc) return rc;
    if (g.dact_lo) {
      rc = make_tmap_bf16(&trl, g.dact_lo, d->N, d->M, g.ld_ref, 16, 32, CU_TENSOR_MAP_SWIZZLE_32B);
      if (rc) return rc;
    } else {
      trl = tr;
    }
  } else {
    tr = ta; trl = ta;
  }
  g.tiles_m = (d->M + kBlockM - 1) / kBlockM;
  g.tiles_n = (d->N + bn - 1) / bn;
  g.splits = splits;
  return URLG_OK;
}

template <int G>
static int launch_group_g(const urlg_gemm_desc* const* ds, int n, cudaStream_t st) {
  GroupArgs<G> P;                           // 1.2 KB of kernel parameters per problem, copied at launch (per call: thread-safe)
  P.num = n;
  int64_t tiles = 0;
  int bn_max = 0, max_ctas = 0;
  P.any_tma_out = P.any_tma_ref = 0;
  for (int i = 0; i < n; ++i) {
    if (int rc = build_problem(ds[i], P.g[i], P.tm[i], n)) return rc;
    const GemmArgs& g = P.g[i];
    if (i && g.split != P.g[0].split) return fail(URLG_EINVAL, "urlg_tc_gemm: grouped problems must share the operand mode (hi+lo or plain)");
    tiles += (int64_t)g.tiles_m * g.tiles_n * g.splits;
    if (tiles > 0x7fffffff) return fail(URLG_EINVAL, "urlg_tc_gemm: too many tiles");
    P.tile_end[i] = (int)tiles;
    bn_max = g.block_n > bn_max ? g.block_n : bn_max;
    P.any_tma_out |= g.tma_out;
    P.any_tma_ref |= g.tma_ref;
    if (ds[i]->max_ctas > 0 && (max_ctas == 0 || ds[i]->max_ctas < max_ctas)) max_ctas = ds[i]->max_ctas;
  }
  for (int i = n; i < G; ++i) P.tile_end[i] = (int)tiles;
  P.split = P.g[0].split;
  P.stages = gemm_stages_for(bn_max, P.split, P.any_tma_out, P.any_tma_ref);
  P.bytes_b_max = stage_bytes_b(bn_max);
  P.acc_cols = (int)tmem_cols_for(bn_max);
  static bool attr_set = false;
  if (!attr_set) {
    URLG_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);           // persistent: one CTA per SM walks the tiles
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  // launched with programmatic stream serialization (see the kernel): back-to-back kernels overlap launch + set-up
  // of the next with the tail of the previous (common.cuh)
  URLG_CUDA(launch_pdl(gemm_bf16_kernel<G>, dim3(grid), dim3(kGemmThreads),
                       gemm_smem_bytes(bn_max, P.split, P.stages, P.any_tma_out, P.any_tma_ref), st, P));
  URLG_LAUNCH_CHECK();
  return URLG_OK;
}

int launch_group_impl(const urlg_gemm_desc* const* ds, int n, cudaStream_t st) {
  if (!ds || n < 1 || n > kMaxGroup) return fail(URLG_EINVAL, "urlg_tc_gemm: 1..%d problems per launch", kMaxGroup);
  if (n == 1) return launch_group_g<1>(ds, n, st);
  if (n == 2) return launch_group_g<2>(ds, n, st);
  return launch_group_g<4>(ds, n, st);
}

}  // namespace tc
}  // namespace urlg

extern "C" int urlg_gemm_timing(int enable) {
  using namespace urlg::tc;
  g_timing = enable != 0;
  if (g_timing) { g_used = 0; g_flops = 0.0; g_bytes = 0.0; }
  return URLG_OK;
}

extern "C" int urlg_gemm_timing_read(double* total_ms, int64_t* launches, double* flops, double* bytes) {
  using namespace urlg::tc;
  double ms = 0.0;
  for (size_t i = 0; i < g_used; ++i) {
    URLG_CUDA(cudaEventSynchronize(g_events[i].second));
    float t = 0.0f;
    URLG_CUDA(cudaEventElapsedTime(&t, g_events[i].first, g_events[i].second));
    ms += t;
  }
  if (total_ms) *total_ms = ms;
  if (launches) *launches = (int64_t)g_used;
  if (flops) *flops = g_flops;
  if (bytes) *bytes = g_bytes;
  return URLG_OK;
}

extern "C" int urlg_tc_gemm_splits(int K, int k_splits) {
  if (K <= 0) return urlg::fail(URLG_EINVAL, "urlg_tc_gemm_splits: K must be positive");
  return urlg::tc::effective_splits(K, k_splits, nullptr);
}

extern "C" int urlg_tc_gemm(const urlg_gemm_desc* d, void* stream) {
  return urlg::tc::launch_gemm(d, (cudaStream_t)stream);
}

extern "C" int urlg_tc_gemm_group(const urlg_gemm_desc* descs, int n, void* stream) {
  if (!descs || n < 1 || n > urlg::tc::kMaxGroup)
    return urlg::fail(URLG_EINVAL, "urlg_tc_gemm_group: 1..%d problems per launch", urlg::tc::kMaxGroup);
  const urlg_gemm_desc* ptr[urlg::tc::kMaxGroup];
  for (int i = 0; i < n; ++i) ptr[i] = descs + i;
  return urlg::tc::launch_gemm_group(ptr, n, (cudaStream_t)stream);
}
