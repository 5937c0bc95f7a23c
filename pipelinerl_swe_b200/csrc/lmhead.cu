// K1 / K2: lm_head forward (fused online log-softmax / entropy / gather) and backward
// (recompute -> dlogits -> dH, dW), plus the plain GEMM entry used by the tests.
// Replaces /root/reference/pipelinerl/finetune/rl/__init__.py:197-208 and its autograd.
#include "gemm_sm100.cuh"
#include <cudaTypedefs.h>
#include <mutex>

namespace b200grpo {

// ---- TMA descriptors (driver entry point fetched through the runtime: no -lcuda) --------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

static int load_encode_fn() {
  static std::once_flag once;
  std::call_once(once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      g_encode = (EncodeTiledFn)fn;
  });
  if (!g_encode) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return B200GRPO_ECUDA; }
  return 0;
}

// bf16 matrix, `inner` contiguous elements per row, `outer` rows, ld elements between rows;
// box = {64 inner elements (128 B, SWIZZLE_128B), box_outer rows}; out-of-range reads give 0.
static int make_tmap(CUtensorMap* tm, const void* ptr, long long inner, long long outer, long long ld, int box_outer) {
  if (int e = load_encode_fn()) return e;
  B200_REQUIRE(((uintptr_t)ptr & 15) == 0, "TMA operand must be 16-byte aligned");
  B200_REQUIRE((ld * 2) % 16 == 0, "TMA operand leading dimension %lld must be a multiple of 8 elements", ld);
  B200_REQUIRE(inner > 0 && outer > 0 && box_outer >= 1 && box_outer <= 256, "TMA operand shape");
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with %d (inner %lld outer %lld ld %lld)", (int)r, inner, outer, ld); return B200GRPO_ECUDA; }
  return 0;
}

// ---- work split: items = m_blocks x n_splits dealt round-robin to the resident clusters -----
static int choose_splits(int m_blocks, int n_blocks, int n_clusters, int forced) {
  if (forced > 0) return forced < n_blocks ? forced : n_blocks;
  int best = 1;
  double best_eff = -1.0;
  const int hi = n_blocks < 256 ? n_blocks : 256;
  for (int s = 1; s <= hi; ++s) {
    const long long items = (long long)m_blocks * s;
    const long long waves = (items + n_clusters - 1) / n_clusters;
    const int tiles_max = (n_blocks + s - 1) / s;
    const double eff = (double)m_blocks * n_blocks / ((double)n_clusters * waves * tiles_max);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
  }
  return best;
}

constexpr int BN = 256;
constexpr int STAGES_1CTA = 4, STAGES_2CTA = 6;

template <int kEpi, int kAMajor, int kBMajor, int kCtas>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st) {
  constexpr int kStages = kCtas == 1 ? STAGES_1CTA : STAGES_2CTA;
  using L = GemmSmem<BN, kCtas, kStages>;
  auto kern = gemm_sm100_kernel<kEpi, kAMajor, kBMajor, kCtas, BN, kStages>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    configured = true;
  }
  const int n_items = p.m_blocks * p.n_splits;
  int clusters = sm_count() / kCtas;
  if (clusters > n_items) clusters = n_items;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * kCtas);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = L::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
  return 0;
}

// Operand maps for  D[M,N] = A . B^T  under the four layout combinations.
static int make_ab_maps(CUtensorMap* ta, CUtensorMap* tb, const void* A, long long lda, int a_mn, const void* B, long long ldb,
                        int b_mn, int M, int N, int K, int ctas) {
  int e;
  if (!a_mn) e = make_tmap(ta, A, K, M, lda, GEMM_BLOCK_M);       // [M rows, K cols]
  else e = make_tmap(ta, A, M, K, lda, GEMM_BLOCK_K);             // [K rows, M cols], 64x64 atoms
  if (e) return e;
  if (!b_mn) e = make_tmap(tb, B, K, N, ldb, BN / ctas);
  else e = make_tmap(tb, B, N, K, ldb, GEMM_BLOCK_K);
  return e;
}

static void fill_tiling(GemmParams& p, int M, int N, int K, int ctas, int forced_splits) {
  p.M = M; p.N = N; p.K = K;
  p.m_blocks = ceil_div(M, GEMM_BLOCK_M * ctas);
  p.n_blocks = ceil_div(N, BN);
  int clusters = sm_count() / ctas;
  if (clusters < 1) clusters = 1;
  p.n_splits = choose_splits(p.m_blocks, p.n_blocks, clusters, forced_splits);
}

template <int kEpi>
static int dispatch(int a_mn, int b_mn, int ctas, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st) {
#define B200_CASE(AM, BM, C) \
  if (a_mn == AM && b_mn == BM && ctas == C) return launch_gemm<kEpi, AM, BM, C>(ta, tb, p, st);
  if constexpr (kEpi == EPI_STORE) {
    B200_CASE(0, 0, 1) B200_CASE(0, 0, 2) B200_CASE(0, 1, 1) B200_CASE(0, 1, 2)
    B200_CASE(1, 0, 1) B200_CASE(1, 0, 2) B200_CASE(1, 1, 1) B200_CASE(1, 1, 2)
  } else {
    B200_CASE(0, 0, 1) B200_CASE(0, 0, 2)
  }
#undef B200_CASE
  set_error("unsupported GEMM variant a_mn=%d b_mn=%d ctas=%d", a_mn, b_mn, ctas);
  return B200GRPO_EINVAL;
}

static int ctas_for(int M) {
  (void)M;
  return get_tunable(TUNE_TWO_CTA) ? 2 : 1;
}

// ---- small kernels around the GEMMs -----------------------------------------------------
// merge per-split partials [S][M][4] -> stats [M][4]  (online-softmax merge; order fixed)
__global__ void __launch_bounds__(256) fwd_merge_kernel(const float4* __restrict__ part, int S, int M, float4* __restrict__ stats) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= M) return;
  float m = -INFINITY;
  for (int s = 0; s < S; ++s) m = fmaxf(m, part[(long long)s * M + row].x);
  float sum = 0.f, dot = 0.f, tgt = 0.f;
  for (int s = 0; s < S; ++s) {
    const float4 q = part[(long long)s * M + row];
    const float w = __expf(q.x - m);
    sum = fmaf(q.y, w, sum); dot = fmaf(q.z, w, dot); tgt += q.w;
  }
  stats[row] = make_float4(m, sum, dot, tgt);
}

// merge per-rank stats [R][M][4] and finish: lse, logprob, entropy  (rl/__init__.py:202-208)
__global__ void __launch_bounds__(256) finalize_kernel(const float4* __restrict__ stats, int R, int M, float* logprob,
                                                       float* entropy, float* lse_out) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= M) return;
  float m = -INFINITY;
  for (int r = 0; r < R; ++r) m = fmaxf(m, stats[(long long)r * M + row].x);
  float sum = 0.f, dot = 0.f, tgt = 0.f;
  for (int r = 0; r < R; ++r) {
    const float4 q = stats[(long long)r * M + row];
    const float w = expf(q.x - m);
    sum = fmaf(q.y, w, sum); dot = fmaf(q.z, w, dot); tgt += q.w;
  }
  const float lse = m + logf(sum);
  logprob[row] = tgt - lse;            // gather(log_softmax(z), target)
  entropy[row] = lse - dot / sum;      // -(p * logp).sum()
  if (lse_out) lse_out[row] = lse;
}

}  // namespace b200grpo

using namespace b200grpo;

static size_t dlogits_ld(int Vloc) { return align_up((size_t)Vloc, 64); }

extern "C" size_t b200grpo_lmhead_fwd_workspace(int M, int H, int Vloc) {
  (void)H;
  GemmParams p;
  fill_tiling(p, M, Vloc, 64, ctas_for(M), get_tunable(TUNE_FWD_SPLITS));
  return align_up((size_t)p.n_splits * M * 16, 256);
}

extern "C" int b200grpo_lmhead_fwd(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                                   int M, int H, int Vloc, int64_t v_offset, float inv_temperature, float* stats,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(hidden && weight && targets && stats && workspace, "lmhead_fwd: null pointer");
  B200_REQUIRE(M >= 1 && H >= 8 && Vloc >= 1 && H % 8 == 0, "lmhead_fwd: M=%d H=%d Vloc=%d (H must be a multiple of 8)", M, H, Vloc);
  B200_REQUIRE(inv_temperature > 0.f, "lmhead_fwd: temperature must be positive");
  const int ctas = ctas_for(M);
  GemmParams p = {};
  fill_tiling(p, M, Vloc, H, ctas, get_tunable(TUNE_FWD_SPLITS));
  const size_t need = align_up((size_t)p.n_splits * M * 16, 256);
  if (workspace_bytes < need) { set_error("lmhead_fwd: workspace %zu < %zu", workspace_bytes, need); return B200GRPO_EWORKSPACE; }
  p.targets = (const long long*)targets; p.v_offset = v_offset; p.inv_temp = inv_temperature; p.part = (float*)workspace;
  CUtensorMap ta, tb;
  if (int e = make_ab_maps(&ta, &tb, hidden, ldh, 0, weight, ldw, 0, M, Vloc, H, ctas)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  { ProfScope ps(PROF_FWD_GEMM, st); if (int e = dispatch<EPI_FWD_STATS>(0, 0, ctas, ta, tb, p, st)) return e; }
  ProfScope ps(PROF_FWD_MERGE, st);
  fwd_merge_kernel<<<ceil_div(M, 256), 256, 0, st>>>((const float4*)workspace, p.n_splits, M, (float4*)stats);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_lmhead_finalize(const float* stats, int R, int M, float* logprob, float* entropy, float* lse, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(stats && logprob && entropy && R >= 1 && M >= 1, "lmhead_finalize: bad argument");
  ProfScope ps(PROF_FINALIZE, (cudaStream_t)stream);
  finalize_kernel<<<ceil_div(M, 256), 256, 0, (cudaStream_t)stream>>>((const float4*)stats, R, M, logprob, entropy, lse);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" size_t b200grpo_lmhead_bwd_workspace(int M, int H, int Vloc) {
  (void)H;
  return align_up((size_t)M * dlogits_ld(Vloc) * 2, 256);
}

extern "C" int b200grpo_lmhead_bwd(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                                   const float* lse, const float* entropy, const float* coef_lp, const float* coef_ent,
                                   const float* grad_scale, int M, int H, int Vloc, int64_t v_offset, float inv_temperature,
                                   void* grad_hidden, int64_t ldgh, void* grad_weight, int64_t ldgw, int grad_weight_is_f32,
                                   int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(hidden && weight && targets && lse && entropy && coef_lp && coef_ent && workspace, "lmhead_bwd: null pointer");
  B200_REQUIRE(M >= 1 && H >= 8 && Vloc >= 1 && H % 8 == 0, "lmhead_bwd: M=%d H=%d Vloc=%d", M, H, Vloc);
  B200_REQUIRE(!accumulate || grad_weight_is_f32, "lmhead_bwd: accumulate needs an f32 grad_weight");
  const size_t ldd = dlogits_ld(Vloc);
  const size_t need = align_up((size_t)M * ldd * 2, 256);
  if (workspace_bytes < need) { set_error("lmhead_bwd: workspace %zu < %zu", workspace_bytes, need); return B200GRPO_EWORKSPACE; }
  const int ctas = ctas_for(M);
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap ta, tb;
  {  // 1. recompute logits tile by tile -> bf16 dlogits
    GemmParams p = {};
    fill_tiling(p, M, Vloc, H, ctas, 0);
    p.targets = (const long long*)targets; p.v_offset = v_offset; p.inv_temp = inv_temperature;
    p.lse = lse; p.entropy = entropy; p.coef_lp = coef_lp; p.coef_ent = coef_ent; p.grad_scale = grad_scale;
    p.dlogits = (__nv_bfloat16*)workspace; p.ldd = (long long)ldd;
    if (int e = make_ab_maps(&ta, &tb, hidden, ldh, 0, weight, ldw, 0, M, Vloc, H, ctas)) return e;
    ProfScope ps(PROF_BWD_DLOGITS, st);
    if (int e = dispatch<EPI_BWD_DLOGITS>(0, 0, ctas, ta, tb, p, st)) return e;
  }
  if (grad_hidden) {  // 2. dH[M,H] = dlogits[M,V] . W[V,H]   (B is MN-major: vocab rows are the K axis)
    GemmParams p = {};
    fill_tiling(p, M, H, Vloc, ctas, 0);
    p.C = grad_hidden; p.ldc = ldgh; p.c_is_f32 = 0; p.accumulate = 0;
    if (int e = make_ab_maps(&ta, &tb, workspace, (long long)ldd, 0, weight, ldw, 1, M, H, Vloc, ctas)) return e;
    ProfScope ps(PROF_BWD_DH, st);
    if (int e = dispatch<EPI_STORE>(0, 1, ctas, ta, tb, p, st)) return e;
  }
  if (grad_weight) {  // 3. dW[V,H] = dlogits^T[V,M] . hidden[M,H]   (both operands MN-major: tokens are the K axis)
    GemmParams p = {};
    fill_tiling(p, Vloc, H, M, ctas, 0);
    p.C = grad_weight; p.ldc = ldgw; p.c_is_f32 = grad_weight_is_f32; p.accumulate = accumulate;
    if (int e = make_ab_maps(&ta, &tb, workspace, (long long)ldd, 1, hidden, ldh, 1, Vloc, H, M, ctas)) return e;
    ProfScope ps(PROF_BWD_DW, st);
    if (int e = dispatch<EPI_STORE>(1, 1, ctas, ta, tb, p, st)) return e;
  }
  return B200GRPO_OK;
}

extern "C" int b200grpo_gemm_bf16(const void* A, int64_t lda, int a_mn_major, const void* B, int64_t ldb, int b_mn_major, void* C,
                                  int64_t ldc, int c_is_f32, int accumulate, int M, int N, int K, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(A && B && C && M >= 1 && N >= 1 && K >= 1, "gemm: bad argument");
  B200_REQUIRE(!accumulate || c_is_f32, "gemm: accumulate needs f32 C");
  const int ctas = ctas_for(M);
  GemmParams p = {};
  fill_tiling(p, M, N, K, ctas, 0);
  p.C = C; p.ldc = ldc; p.c_is_f32 = c_is_f32; p.accumulate = accumulate;
  CUtensorMap ta, tb;
  if (int e = make_ab_maps(&ta, &tb, A, lda, a_mn_major, B, ldb, b_mn_major, M, N, K, ctas)) return e;
  ProfScope ps(PROF_GEMM, (cudaStream_t)stream);
  return dispatch<EPI_STORE>(a_mn_major ? 1 : 0, b_mn_major ? 1 : 0, ctas, ta, tb, p, (cudaStream_t)stream);
}
