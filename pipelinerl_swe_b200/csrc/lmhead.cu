// K1 / K2: lm_head forward (fused online log-softmax / entropy / gather) and backward
// (recompute -> dlogits -> dH, dW), plus the plain GEMM entry used by the tests.
// Replaces /root/reference/pipelinerl/finetune/rl/__init__.py:197-208 and its autograd.
#include "gemm_sm100.cuh"
#include <atomic>
#include <mutex>
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

// bf16 panels [n_panels][rows][256]: dims {256, rows, n_panels}; box = {64, box_rows, 1}
static int make_tmap_panel(CUtensorMap* tm, const void* ptr, long long rows, long long n_panels, int box_rows) {
  if (int e = load_encode_fn()) return e;
  B200_REQUIRE(((uintptr_t)ptr & 255) == 0, "panel operand must be 256-byte aligned");
  cuuint64_t gdim[3] = {(cuuint64_t)PANEL_W, (cuuint64_t)rows, (cuuint64_t)n_panels};
  cuuint64_t gstride[2] = {(cuuint64_t)PANEL_W * 2, (cuuint64_t)rows * PANEL_W * 2};
  cuuint32_t box[3] = {64u, (cuuint32_t)box_rows, 1u};
  cuuint32_t estr[3] = {1u, 1u, 1u};
  const CUresult r = g_encode(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstride, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (panel) failed with %d (rows %lld panels %lld)", (int)r, rows, n_panels); return B200GRPO_ECUDA; }
  return 0;
}

// ---- tiling --------------------------------------------------------------------------------
constexpr int BN = 256;
#ifndef B200_STAGES_2CTA
#define B200_STAGES_2CTA 6
#endif
constexpr int STAGES_1CTA = 4, STAGES_2CTA = B200_STAGES_2CTA, STAGES_BIGM = 4;   // 32 KB, 32 KB(pair: 2 x 16+16) and 48 KB stages

// k-splits for an EPI_STORE GEMM: smallest split count whose wave efficiency is within 1% of the
// best reachable with <= 8 splits (each split costs an f32 slab round trip through HBM).
static int choose_ksplits(int tiles, int clusters, int kb_total, int forced) {
  if (forced > 0) return forced < kb_total ? forced : kb_total;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= 8; ++s) {
    if (s > 1 && kb_total / s < 64) break;
    const long long items = (long long)tiles * s;
    const long long waves = (items + clusters - 1) / clusters;
    const double eff = (double)items / (double)(waves * clusters);
    if (eff > best_eff + 0.01) { best_eff = eff; best = s; }
  }
  return best;
}

// ---- soft K-lockstep: arrival counters of one launch ----------------------------------------------------------------------
// A ring of counters in device memory that belongs to the library (no caller workspace to size, nothing allocated): each
// launch takes the next slice and zeroes it on its own stream right before the kernel.  A slice comes round again after
// LOCKSTEP_RING / (counters per launch) launches -- hundreds of GEMMs later; should a launch that old still be running, its
// pacing degrades to the time-out (lockstep_pace), nothing else.
constexpr int LOCKSTEP_RING = 1 << 16;
static __device__ __align__(256) unsigned g_lockstep_ring[LOCKSTEP_RING];
static std::atomic<unsigned> g_lockstep_cursor{0};
static void* g_lockstep_addr[64] = {};
static std::mutex g_lockstep_mu;

// Pacing is optional: when the slice cannot be set up (symbol look-up refused, e.g. first use inside a stream capture) the
// launch simply runs un-paced -- never an error of the GEMM.
static bool lockstep_slice(int n, cudaStream_t st, unsigned** out) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { (void)cudaGetLastError(); return false; }
  void* base;
  {
    std::lock_guard<std::mutex> lk(g_lockstep_mu);
    if (!g_lockstep_addr[dev] && cudaGetSymbolAddress(&g_lockstep_addr[dev], g_lockstep_ring) != cudaSuccess) {
      g_lockstep_addr[dev] = nullptr;
      (void)cudaGetLastError();
      return false;
    }
    base = g_lockstep_addr[dev];
  }
  unsigned start = g_lockstep_cursor.fetch_add((unsigned)n) % LOCKSTEP_RING;
  if (start + (unsigned)n > LOCKSTEP_RING) start = 0;   // do not straddle the end
  unsigned* slice = (unsigned*)base + start;
  if (cudaMemsetAsync(slice, 0, (size_t)n * sizeof(unsigned), st) != cudaSuccess) { (void)cudaGetLastError(); return false; }
  *out = slice;
  return true;
}

template <int kEpi, int kAMajor, int kBMajor, int kCtas, bool kSave, bool kAPanel, int kMRep = 1>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p_in, cudaStream_t st) {
  GemmParams p = p_in;
  constexpr int kStages = kMRep == 2 ? STAGES_BIGM : kCtas == 1 ? STAGES_1CTA : STAGES_2CTA;
  using L = GemmSmem<BN, kCtas, kStages, kMRep>;
  static_assert(BN == PANEL_W, "panels are one N tile wide");
  static_assert(L::TOTAL <= 227 * 1024, "shared memory budget");
  auto kern = gemm_sm100_kernel<kEpi, kAMajor, kBMajor, kCtas, BN, kStages, kSave, kAPanel, kMRep>;
  static PerDeviceOnce configured;
  if (configured.first()) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
  }
  B200_REQUIRE((long long)p.m_blocks * p.n_blocks * p.k_splits < (1ll << 30), "gemm: too many tiles");
  const long long n_items = p.n_items;
  int clusters = sm_count() / kCtas;
  if (clusters > n_items) clusters = (int)n_items;
  // soft K-lockstep (tunables lockstep_fwd / _dh / _dw = k-blocks per epoch, lockstep_slack = epochs of lead allowed)
  p.sync = nullptr;
  {
    const int epoch_kb = get_tunable(kEpi != EPI_STORE ? TUNE_LOCKSTEP_FWD : (kAMajor == MAJOR_K ? TUNE_LOCKSTEP_DH : TUNE_LOCKSTEP_DW));
    const int slack = get_tunable(TUNE_LOCKSTEP_SLACK) < 1 ? 1 : get_tunable(TUNE_LOCKSTEP_SLACK);
    const long long n_full = n_items < p.tail_first ? n_items : p.tail_first;
    if (epoch_kb > 0 && p.k_splits <= 1 && clusters > 1 && n_full > 0) {
      const int epochs = ceil_div(p.kb_total, epoch_kb);
      const long long waves = (n_full + clusters - 1) / clusters;
      if (epochs > slack && waves * epochs <= LOCKSTEP_RING / 4) {
        if (lockstep_slice((int)(waves * epochs), st, &p.sync)) {
          p.sync_kb = epoch_kb; p.sync_epochs = epochs; p.sync_full = (int)n_full; p.sync_slack = slack;
        } else {
          p.sync = nullptr;
        }
      }
    }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * kCtas);
  cfg.blockDim = dim3(GEMM_THREADS + 128 * (kMRep - 1));   // one more epilogue warpgroup per extra 128 rows
  cfg.dynamicSmemBytes = L::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p)); count_launch();
  return 0;
}

// Operand maps for  D[M,N] = A . B^T  under the four layout combinations.
static int make_ab_maps(CUtensorMap* ta, CUtensorMap* tb, const void* A, long long lda, int a_mn, const void* B, long long ldb,
                        int b_mn, int M, int N, int K, int ctas, int mrep = 1) {
  int e;
  if (!a_mn) e = make_tmap(ta, A, K, M, lda, GEMM_BLOCK_M * mrep);  // [M rows, K cols]
  else e = make_tmap(ta, A, M, K, lda, GEMM_BLOCK_K);             // [K rows, M cols], 64x64 atoms
  if (e) return e;
  if (!b_mn) e = make_tmap(tb, B, K, N, ldb, BN / ctas);
  else e = make_tmap(tb, B, N, K, ldb, GEMM_BLOCK_K);
  return e;
}

static int clusters_for(int ctas) {
  int c = sm_count() / ctas;
  return c < 1 ? 1 : c;
}

// group_m <= 0: all row blocks in one group (row block fastest over the whole matrix)
static unsigned long long hint_code(int h) {  // tunable encoding: 0 normal, 1 evict-first, 2 evict-last
  return h == 1 ? ptx::L2_EVICT_FIRST : h == 2 ? ptx::L2_EVICT_LAST : ptx::L2_EVICT_NORMAL;
}

static void fill_tiling(GemmParams& p, int M, int N, int K, int ctas, int group_m, int k_splits, int group_n = 0,
                        int hint_a = 0, int hint_b = 0, int mrep = 1) {
  p.M = M; p.N = N; p.K = K;
  p.hint_a = hint_code(hint_a); p.hint_b = hint_code(hint_b);
  p.m_blocks = ceil_div(M, GEMM_BLOCK_M * mrep * ctas);
  p.n_blocks = ceil_div(N, BN);
  p.kb_total = ceil_div(K, GEMM_BLOCK_K);
  p.group_m = (group_m <= 0 || group_m > p.m_blocks) ? p.m_blocks : group_m;
  p.k_splits = k_splits < 1 ? 1 : k_splits;
  p.group_n = (group_n <= 0 || group_n > p.n_blocks) ? p.n_blocks : group_n;
  p.n_items = p.m_blocks * p.n_blocks * p.k_splits;
  p.tail_first = 0x7fffffff; p.tail_split = 0;
  p.tile_rows = GEMM_BLOCK_M * mrep * ctas;
}

// Tail split of an EPI_STORE GEMM: tiles = q * clusters + r; the r tiles of the last partial wave are cut into
// s = clusters / r K-pieces each so that the last wave costs 1/s of a full one.  Returns s (0 = not worth it).
constexpr int TAIL_MIN_KB = 8;   // fewest k-blocks in a piece
static int tail_split_for(int tiles, int clusters, int kb_total) {
  const int r = tiles % clusters;
  if (r == 0) return 0;
  int s = clusters / r;
  if (s > kb_total / TAIL_MIN_KB) s = kb_total / TAIL_MIN_KB;
  return s >= 2 ? s : 0;
}
static size_t tail_slab_bytes(int tiles, int clusters, int kb_total, int tile_rows) {
  const int s = tail_split_for(tiles, clusters, kb_total);
  return s ? align_up((size_t)(tiles % clusters) * s * tile_rows * PANEL_W * 4, 256) : 0;
}

template <int kEpi, bool kSave = false, bool kAPanel = false>
static int dispatch(int a_mn, int b_mn, int ctas, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t st,
                    int mrep = 1) {
#define B200_CASE(AM, BM, C) \
  if (a_mn == AM && b_mn == BM && ctas == C && mrep == 1) return launch_gemm<kEpi, AM, BM, C, kSave, kAPanel>(ta, tb, p, st);
#define B200_BIG(AM, BM) \
  if (a_mn == AM && b_mn == BM && ctas == 2 && mrep == 2) return launch_gemm<kEpi, AM, BM, 2, kSave, kAPanel, 2>(ta, tb, p, st);
  if constexpr (kEpi == EPI_STORE && kAPanel) {   // the two backward GEMMs read dlogits panels
    B200_CASE(0, 1, 1) B200_CASE(0, 1, 2) B200_CASE(1, 1, 1) B200_CASE(1, 1, 2)
    B200_BIG(0, 1) B200_BIG(1, 1)
  } else if constexpr (kEpi == EPI_STORE) {
    B200_CASE(0, 0, 1) B200_CASE(0, 0, 2) B200_CASE(0, 1, 1) B200_CASE(0, 1, 2)
    B200_CASE(1, 0, 1) B200_CASE(1, 0, 2) B200_CASE(1, 1, 1) B200_CASE(1, 1, 2)
    B200_BIG(0, 0) B200_BIG(0, 1) B200_BIG(1, 0) B200_BIG(1, 1)
  } else if constexpr (kEpi == EPI_FWD_STATS) {
    B200_CASE(0, 0, 1) B200_CASE(0, 0, 2) B200_BIG(0, 0)
  } else {
    B200_CASE(0, 0, 1) B200_CASE(0, 0, 2)
  }
#undef B200_CASE
#undef B200_BIG
  set_error("unsupported GEMM variant a_mn=%d b_mn=%d ctas=%d rows/CTA=%d", a_mn, b_mn, ctas, 128 * mrep);
  return B200GRPO_EINVAL;
}

static int ctas_for(int M) {
  (void)M;
  return get_tunable(TUNE_TWO_CTA) ? 2 : 1;
}

// 256 rows per CTA (pair tile 512 x 256) when the tunable asks for it and CTA pairs are in use
static int mrep_for(int tunable, int ctas) { return (ctas == 2 && get_tunable(tunable)) ? 2 : 1; }

// ---- small kernels around the GEMMs -----------------------------------------------------
// Per-row reference r of the forward's exponentials (units of z/T): the largest of the target logit (when the target lies in
// this rank's vocab rows) and four sampled logits.  Any value works for correctness -- e = exp(z/T - r) is kept in f32 / bf16,
// whose exponent range is +-88 in these units, and every later use re-scales by exp(r - lse) -- r only has to sit within
// that distance of the row's log-sum-exp, and a sampled token's own logit does (its log-prob would be < -80 otherwise; an
// overflow surfaces as a non-finite log-prob, i.e. the reference's own assert at rl/__init__.py:209).
// One warp per row: the hidden row is read once (16-byte vectors), dotted with up to five weight rows.
constexpr int RR_COLS = 4;
__global__ void __launch_bounds__(256) row_ref_kernel(const __nv_bfloat16* __restrict__ hidden, long long ldh,
                                                      const __nv_bfloat16* __restrict__ weight, long long ldw,
                                                      const long long* __restrict__ targets, int M, int H, int Vloc, long long v_offset,
                                                      float inv_temp, float* __restrict__ row_ref) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const long long t = targets[row] - v_offset;
  const bool has_t = t >= 0 && t < Vloc;
  const __nv_bfloat16* wr[RR_COLS + 1];
#pragma unroll
  for (int c = 0; c < RR_COLS; ++c) wr[c] = weight + (long long)((long long)c * Vloc / RR_COLS) * ldw;
  wr[RR_COLS] = weight + (has_t ? t : 0) * ldw;
  float acc[RR_COLS + 1];
#pragma unroll
  for (int c = 0; c <= RR_COLS; ++c) acc[c] = 0.f;
  const __nv_bfloat16* h = hidden + (long long)row * ldh;
  const bool vec = ((ldh | ldw) & 7) == 0 && (((uintptr_t)hidden | (uintptr_t)weight) & 15) == 0;
  if (vec) {
    for (int k = lane * 8; k < H; k += 256) {
      const uint4 hv = __ldg(reinterpret_cast<const uint4*>(h + k));
      const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&hv);
#pragma unroll
      for (int c = 0; c <= RR_COLS; ++c) {
        const uint4 wv = __ldg(reinterpret_cast<const uint4*>(wr[c] + k));
        const __nv_bfloat162* wp = reinterpret_cast<const __nv_bfloat162*>(&wv);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 a = __bfloat1622float2(hp[q]), w = __bfloat1622float2(wp[q]);
          acc[c] = fmaf(a.x, w.x, fmaf(a.y, w.y, acc[c]));
        }
      }
    }
  } else {
    for (int k = lane; k < H; k += 32) {
      const float a = __bfloat162float(h[k]);
#pragma unroll
      for (int c = 0; c <= RR_COLS; ++c) acc[c] = fmaf(a, __bfloat162float(wr[c][k]), acc[c]);
    }
  }
  float r = -INFINITY;
#pragma unroll
  for (int c = 0; c <= RR_COLS; ++c) {
    const float d = warp_sum(acc[c]);
    if (c < RR_COLS || has_t) r = fmaxf(r, d);
  }
  if (lane == 0) row_ref[row] = r * inv_temp;
}

// merge the per-tile partials [n_blocks][M][4] -> stats [M][4] = (r, sum e, sum e*z, z_target): every tile of a row shares the
// reference r, so the merge is three plain sums in a fixed order (deterministic).
// Block = 32 rows x 8 tile slices: a warp reads 512 contiguous bytes per tile, 4 loads in flight.
constexpr int MG_SLICES = 8;
// kMulticast: `stats` is an NVSwitch multicast address (this rank's slot of a symmetric [R][M][4] buffer): the merged row is
// stored with multimem.st, so the merge IS the all-gather of the vocab-parallel exchange -- every rank's copy of the slot is
// written by the switch, no collective kernel follows (the caller only synchronises).
template <bool kMulticast>
__global__ void __launch_bounds__(32 * MG_SLICES) fwd_merge_kernel(const float4* __restrict__ part, const float* __restrict__ row_ref, int S,
                                                                   int M, float4* __restrict__ stats) {
  const int rx = threadIdx.x & 31, sy = threadIdx.x >> 5;
  const int row = blockIdx.x * 32 + rx;
  float sum = 0.f, dot = 0.f, tgt = 0.f;
  if (row < M) {
    int s = sy;
    for (; s + 3 * MG_SLICES < S; s += 4 * MG_SLICES) {
      float4 q[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) q[j] = __ldg(&part[(long long)(s + j * MG_SLICES) * M + row]);
#pragma unroll
      for (int j = 0; j < 4; ++j) { sum += q[j].y; dot += q[j].z; tgt += q[j].w; }
    }
    for (; s < S; s += MG_SLICES) {
      const float4 q = __ldg(&part[(long long)s * M + row]);
      sum += q.y; dot += q.z; tgt += q.w;
    }
  }
  __shared__ float4 sh[MG_SLICES][32];
  sh[sy][rx] = make_float4(0.f, sum, dot, tgt);
  __syncthreads();
  if (sy == 0 && row < M) {
    float S2 = 0.f, D2 = 0.f, T2 = 0.f;
#pragma unroll
    for (int j = 0; j < MG_SLICES; ++j) { const float4 q = sh[j][rx]; S2 += q.y; D2 += q.z; T2 += q.w; }
    const float M2 = row_ref[row];
    if constexpr (kMulticast) {
      asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(stats + row), "f"(M2), "f"(S2), "f"(D2), "f"(T2) : "memory");
    } else {
      stats[row] = make_float4(M2, S2, D2, T2);
    }
  }
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

// Saved forward exponentials -> dlogits, in place (the recompute-free backward WITH an entropy term; without one the
// backward never forms dlogits at all, see lmhead_bwd_impl).  z holds E = exp(logit/T - r) as bf16 in [n_blocks][M][256]
// panels; p = E * exp(r - lse).  HBM-bound: 4 B/element (bf16 in + bf16 out), 16-byte vectors, 32 threads per panel row,
// fully contiguous traversal.
constexpr int DL_THREADS = 256;
constexpr int DL_ROWS = 32;   // panel rows per block: 32 rows x 256 columns = 16 KB in, 16 KB out
__global__ void __launch_bounds__(DL_THREADS) dlogits_from_saved_kernel(
    __nv_bfloat16* __restrict__ z, const float* __restrict__ row_ref, int M, int N, const long long* __restrict__ targets,
    long long v_offset, const float* __restrict__ lse, const float* __restrict__ entropy, const float* __restrict__ coef_lp,
    const float* __restrict__ coef_ent, const float* __restrict__ grad_scale, float inv_temp) {
  __shared__ float s_a[DL_ROWS], s_e[DL_ROWS], s_ent[DL_ROWS], s_shift[DL_ROWS], s_c[DL_ROWS];
  __shared__ int s_t[DL_ROWS];
  const int nb = blockIdx.y, row0 = blockIdx.x * DL_ROWS;
  if (threadIdx.x < DL_ROWS) {   // per-row constants, loaded once per block
    const int row = row0 + threadIdx.x;
    if (row < M) {
      const float gs = (grad_scale ? *grad_scale : 1.0f) * inv_temp;
      s_a[threadIdx.x] = coef_lp[row] * gs;
      s_e[threadIdx.x] = coef_ent[row] * gs;
      s_ent[threadIdx.x] = entropy[row];
      const float d = row_ref[row] - lse[row];             // log p = log E + (r - lse)
      s_shift[threadIdx.x] = d * LOG2E_F;
      s_c[threadIdx.x] = __expf(d);
      const long long tl = targets[row] - v_offset - (long long)nb * PANEL_W;
      s_t[threadIdx.x] = (tl >= 0 && tl < PANEL_W) ? (int)tl : -1000;
    }
  }
  __syncthreads();
  const int sub = threadIdx.x & 31, wr = threadIdx.x >> 5;   // 8 columns per thread; warp = one panel row
  const int col = nb * PANEL_W + sub * 8;
  if (col >= N) return;                                      // tail columns stay zero
  const int n_valid = min(8, N - col);
  __nv_bfloat16* base = z + ((long long)nb * M + row0) * PANEL_W + sub * 8;
  uint4 raw[DL_ROWS / 8];
#pragma unroll
  for (int j = 0; j < DL_ROWS / 8; ++j) {                    // 4 independent 16-byte loads in flight
    const int r = wr + 8 * j;
    if (row0 + r < M) raw[j] = *reinterpret_cast<const uint4*>(base + (long long)r * PANEL_W);
  }
#pragma unroll
  for (int j = 0; j < DL_ROWS / 8; ++j) {
    const int r = wr + 8 * j;
    if (row0 + r >= M) continue;
    const float a = s_a[r], e = s_e[r], ent = s_ent[r], shift2 = s_shift[r], c = s_c[r];
    const int t = s_t[r] - sub * 8;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw[j]);
    float d[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 y = __bfloat1622float2(h[q]);
#pragma unroll
      for (int w = 0; w < 2; ++w) {
        const float ev = w ? y.y : y.x;
        const float pr = ev * c;
        const float lp = (__log2f(ev) + shift2) * LN2_F;                      // log p (-inf when E underflowed to 0)
        d[2 * q + w] = pr > 0.f ? -pr * fmaf(e, lp + ent, a) : 0.f;          // -p * (a + e*(logp + H))
      }
    }
    if ((unsigned)t < 8u || n_valid < 8) {   // rare: the row's target column or the vocab tail is in this vector
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (k == t) d[k] += a;
        if (k >= n_valid) d[k] = 0.f;
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) h[q] = __floats2bfloat162_rn(d[2 * q], d[2 * q + 1]);
    *reinterpret_cast<uint4*>(base + (long long)r * PANEL_W) = raw[j];
  }
}

// ---- the entropy-free backward never forms dlogits: with coef_ent == 0, dlogits = a * (onehot - p) and p = E * exp(r - lse),
// so   dH = diag(s) . (E . W) + a * W[target]          dW = E^T . (diag(s) . hidden) + scatter(a * hidden at the target rows)
// with the per-row scale s = -a * exp(r - lse).  The two GEMMs read the forward's E panels as they are; what is left of the
// 5 GB dlogits pass is this row kernel (scale + scaled copy of hidden, 2 x 58 MB at C3) and a sparse add of M rows into dW.
__global__ void __launch_bounds__(256) bwd_rows_kernel(const __nv_bfloat16* __restrict__ hidden, long long ldh, const float* __restrict__ row_ref,
                                                       const float* __restrict__ lse, const float* __restrict__ coef_lp,
                                                       const float* __restrict__ grad_scale, float inv_temp, int M, int H,
                                                       const long long* __restrict__ targets, long long v_offset, int Vloc,
                                                       float* __restrict__ row_scale, float* __restrict__ row_coef, int* __restrict__ row_key,
                                                       __nv_bfloat16* __restrict__ hs, __nv_bfloat16* __restrict__ panels) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const float a = coef_lp[row] * (grad_scale ? *grad_scale : 1.0f) * inv_temp;
  const float sc = a != 0.f ? -a * __expf(row_ref[row] - lse[row]) : 0.f;
  if (lane == 0) {
    const long long t = targets[row] - v_offset;          // key of the onehot term: local target row, or -1 when there is none
    bool has = a != 0.f && t >= 0 && t < Vloc;
    float coef = a;
    // dlogit[target] = a (1 - p) is a difference of nearly equal numbers for a confident token.  Adding the onehot terms
    // (a * W[target] in dH's epilogue, a * hidden[t] in the sparse dW kernel) AFTER GEMMs whose operands were rounded to bf16
    // keeps 2^-9 of 1 where 2^-9 of 1 - p is needed (dW 1.5x off at p = 0.998).  For p > 1/2 the onehot is therefore folded
    // into the target's saved exponential here, in f32, BEFORE the GEMMs: s (E - exp(lse - r)) = a (1 - p), and only the
    // difference is rounded.  For p <= 1/2 the onehot terms stay exact f32 additions after the GEMMs.
    if (panels && has) {
      __nv_bfloat16* e = panels + ((long long)(t / PANEL_W) * M + row) * PANEL_W + (t % PANEL_W);
      const float ev = __bfloat162float(*e), big = __expf(lse[row] - row_ref[row]);
      if (ev > 0.5f * big) {
        *e = __float2bfloat16_rn(ev - big);
        has = false; coef = 0.f;
      }
    }
    row_scale[row] = sc; row_coef[row] = coef;
    row_key[row] = has ? (int)t : -1;
  }
  if (!hs) return;
  const __nv_bfloat16* h = hidden + (long long)row * ldh;
  __nv_bfloat16* o = hs + (long long)row * H;
  if ((ldh & 7) == 0 && ((uintptr_t)hidden & 15) == 0) {
    for (int k = lane * 8; k < H; k += 256) {
      uint4 v = sc != 0.f ? __ldg(reinterpret_cast<const uint4*>(h + k)) : make_uint4(0, 0, 0, 0);
      __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&v);
      if (sc != 0.f) {
#pragma unroll
        for (int q = 0; q < 4; ++q) { const float2 f = __bfloat1622float2(p[q]); p[q] = __floats2bfloat162_rn(f.x * sc, f.y * sc); }
      }
      *reinterpret_cast<uint4*>(o + k) = v;
    }
  } else {
    for (int k = lane; k < H; k += 32) o[k] = __float2bfloat16_rn(sc != 0.f ? __bfloat162float(h[k]) * sc : 0.f);
  }
}

// dW[v, :] += sum over the tokens t with key(t) == v of a_t * hidden[t, :]   (the onehot term of the entropy-free backward;
// key = local target row, -1 for tokens without one).  Deterministic and atomic-free: one warp per token row scans the key
// column (32 KB at T = 8192: L1-resident after the first touch; four 16-byte loads in flight per lane, 512 keys per trip);
// the FIRST token of every distinct key is its leader and folds all tokens of that key, in token order, into an f32 row in
// shared memory, then updates the gradient row once -- every element of dW is touched at most once per launch whatever the
// multiplicity of a token in the micro-batch (hot tokens repeat hundreds of times in real text).  row_key is padded to a
// multiple of 512 entries with -2 by bwd_rows_kernel's launch (see lmhead_bwd_impl).
constexpr int DWT_TRIP = 512;
__global__ void __launch_bounds__(128) dw_target_kernel(const __nv_bfloat16* __restrict__ hidden, long long ldh, const int* __restrict__ row_key,
                                                        const float* __restrict__ row_coef, int M, int H, void* grad_weight,
                                                        long long ldgw, int gw_is_f32) {
  extern __shared__ __align__(16) unsigned char dwt_smem[];
  const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * warps + warp;
  if (row >= M) return;
  const int key = row_key[row];
  if (key < 0) return;
  float* acc = reinterpret_cast<float*>(dwt_smem) + (size_t)warp * H;
  const bool vec = (ldh & 7) == 0 && ((uintptr_t)hidden & 15) == 0 && (H & 7) == 0;
  bool leader = false;
  const int M_pad = (M + DWT_TRIP - 1) / DWT_TRIP * DWT_TRIP;
  for (int j0 = 0; j0 < M_pad; j0 += DWT_TRIP) {           // warp-uniform control flow throughout
    int4 k4[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) k4[u] = __ldg(reinterpret_cast<const int4*>(row_key + j0 + u * 128) + lane);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const unsigned m0 = __ballot_sync(0xffffffffu, k4[u].x == key), m1 = __ballot_sync(0xffffffffu, k4[u].y == key);
      const unsigned m2 = __ballot_sync(0xffffffffu, k4[u].z == key), m3 = __ballot_sync(0xffffffffu, k4[u].w == key);
      unsigned any = m0 | m1 | m2 | m3;
      while (any) {                                        // matches in increasing token order: lane-major, then the four keys
        const int l = __ffs(any) - 1;
        any &= any - 1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const unsigned mq = q == 0 ? m0 : q == 1 ? m1 : q == 2 ? m2 : m3;
          if (!((mq >> l) & 1u)) continue;
          const int jj = j0 + u * 128 + l * 4 + q;
          if (!leader) {
            if (jj != row) return;                         // an earlier token owns this key
            if (vec) for (int k = lane * 4; k < H; k += 128) *reinterpret_cast<float4*>(acc + k) = make_float4(0.f, 0.f, 0.f, 0.f);
            else for (int k = lane; k < H; k += 32) acc[k] = 0.f;
            leader = true;
          }
          const float a = __ldg(row_coef + jj);
          const __nv_bfloat16* h = hidden + (long long)jj * ldh;
          if (vec) {   // 16-byte loads, four in flight per lane
            for (int k0 = lane * 8; k0 < H; k0 += 4 * 256) {
              uint4 hv[4];
#pragma unroll
              for (int w = 0; w < 4; ++w)
                if (k0 + w * 256 < H) hv[w] = __ldg(reinterpret_cast<const uint4*>(h + k0 + w * 256));
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                const int k = k0 + w * 256;
                if (k >= H) break;
                const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&hv[w]);
                float4 x = *reinterpret_cast<float4*>(acc + k), y = *reinterpret_cast<float4*>(acc + k + 4);
                const float2 f0 = __bfloat1622float2(hp[0]), f1 = __bfloat1622float2(hp[1]), f2 = __bfloat1622float2(hp[2]),
                             f3 = __bfloat1622float2(hp[3]);
                x.x = fmaf(a, f0.x, x.x); x.y = fmaf(a, f0.y, x.y); x.z = fmaf(a, f1.x, x.z); x.w = fmaf(a, f1.y, x.w);
                y.x = fmaf(a, f2.x, y.x); y.y = fmaf(a, f2.y, y.y); y.z = fmaf(a, f3.x, y.z); y.w = fmaf(a, f3.y, y.w);
                *reinterpret_cast<float4*>(acc + k) = x; *reinterpret_cast<float4*>(acc + k + 4) = y;
              }
            }
          } else {
            for (int k = lane * 2; k < H; k += 64) {
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(h + k));
              acc[k] = fmaf(a, f.x, acc[k]);
              acc[k + 1] = fmaf(a, f.y, acc[k + 1]);
            }
          }
        }
      }
    }
  }
  if (!leader) return;
  __syncwarp();
  if (gw_is_f32) {
    float* g = reinterpret_cast<float*>(grad_weight) + (long long)key * ldgw;
    if ((ldgw & 3) == 0 && ((uintptr_t)grad_weight & 15) == 0 && (H & 3) == 0) {
      for (int k0 = lane * 4; k0 < H; k0 += 4 * 128) {     // four independent 16-byte read-modify-writes per lane
        float4 gv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) if (k0 + u * 128 < H) gv[u] = *reinterpret_cast<const float4*>(g + k0 + u * 128);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int k = k0 + u * 128;
          if (k >= H) break;
          const float4 x = *reinterpret_cast<const float4*>(acc + k);
          *reinterpret_cast<float4*>(g + k) = make_float4(gv[u].x + x.x, gv[u].y + x.y, gv[u].z + x.z, gv[u].w + x.w);
        }
      }
    } else {
      for (int k = lane * 2; k < H; k += 64) { g[k] += acc[k]; g[k + 1] += acc[k + 1]; }
    }
  } else {
    __nv_bfloat16* g = reinterpret_cast<__nv_bfloat16*>(grad_weight) + (long long)key * ldgw;
    for (int k = lane * 2; k < H; k += 64) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(g + k));
      *reinterpret_cast<__nv_bfloat162*>(g + k) = __floats2bfloat162_rn(f.x + acc[k], f.y + acc[k + 1]);
    }
  }
}

// sum the k-split slabs [S][M][N] (f32) into C (bf16 or f32, optionally accumulating)
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ slab, int S, int M, int N, void* C,
                                                            long long ldc, int c_is_f32, int accumulate) {
  const long long quads = (long long)M * (N / 4);
  const long long plane = (long long)M * N;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (long long)gridDim.x * blockDim.x) {
    const int row = (int)(i / (N / 4)), c4 = (int)(i - (long long)row * (N / 4)) * 4;
    float4 acc = __ldg(reinterpret_cast<const float4*>(slab + (long long)row * N + c4));
    for (int s = 1; s < S; ++s) {
      const float4 x = __ldg(reinterpret_cast<const float4*>(slab + s * plane + (long long)row * N + c4));
      acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
    }
    if (c_is_f32) {
      float* o = reinterpret_cast<float*>(C) + (long long)row * ldc + c4;
      if (accumulate) { acc.x += o[0]; acc.y += o[1]; acc.z += o[2]; acc.w += o[3]; }
      o[0] = acc.x; o[1] = acc.y; o[2] = acc.z; o[3] = acc.w;
    } else {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(C) + (long long)row * ldc + c4;
      o[0] = __float2bfloat16_rn(acc.x); o[1] = __float2bfloat16_rn(acc.y);
      o[2] = __float2bfloat16_rn(acc.z); o[3] = __float2bfloat16_rn(acc.w);
    }
  }
}

// sum the K-pieces of the tail tiles: block = 4 tile rows x 64 float4 columns; grid (tail tiles, tile_rows / 4)
constexpr int TAIL_RED_ROWS = 4;
__global__ void __launch_bounds__(256) tail_reduce_kernel(const GemmParams p) {
  int m_blk, nb;
  decode_tile(p, p.tail_first + blockIdx.x, m_blk, nb);
  const int r_local = blockIdx.y * TAIL_RED_ROWS + (threadIdx.x >> 6), c_local = (threadIdx.x & 63) * 4;
  const int row = m_blk * p.tile_rows + r_local, col = nb * PANEL_W + c_local;
  if (row >= p.M || col >= p.N) return;                // N % 4 == 0: a quad is inside or outside as a whole
  const float* src = p.slab + ((long long)blockIdx.x * p.tail_split * p.tile_rows + r_local) * PANEL_W + c_local;
  const long long piece = (long long)p.tile_rows * PANEL_W;
  float4 acc = __ldg(reinterpret_cast<const float4*>(src));
  for (int s = 1; s < p.tail_split; ++s) {
    const float4 x = __ldg(reinterpret_cast<const float4*>(src + s * piece));
    acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
  }
  if (p.row_scale || p.row_coef) {   // the row epilogue the full tiles apply in the GEMM (store_chunk)
    const RowEpi re = load_row_epi(p, row);
    acc.x *= re.scale; acc.y *= re.scale; acc.z *= re.scale; acc.w *= re.scale;
    if (re.src) {
      acc.x = fmaf(re.coef, __bfloat162float(re.src[col]), acc.x); acc.y = fmaf(re.coef, __bfloat162float(re.src[col + 1]), acc.y);
      acc.z = fmaf(re.coef, __bfloat162float(re.src[col + 2]), acc.z); acc.w = fmaf(re.coef, __bfloat162float(re.src[col + 3]), acc.w);
    }
  }
  if (p.c_is_f32) {
    float* o = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col;
    if (p.accumulate) { acc.x += o[0]; acc.y += o[1]; acc.z += o[2]; acc.w += o[3]; }
    o[0] = acc.x; o[1] = acc.y; o[2] = acc.z; o[3] = acc.w;
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col;
    o[0] = __float2bfloat16_rn(acc.x); o[1] = __float2bfloat16_rn(acc.y);
    o[2] = __float2bfloat16_rn(acc.z); o[3] = __float2bfloat16_rn(acc.w);
  }
}

// optional row epilogue of a store GEMM (see GemmParams)
struct RowEpiArgs {
  const float* row_scale = nullptr;
  const float* row_coef = nullptr;
  const __nv_bfloat16* gather_src = nullptr;
  long long gather_ld = 0;
  int gather_rows = 0;
  const long long* targets = nullptr;
  long long v_offset = 0;
};

// C[M,N] (+)= A . B^T with optional k-splits; slab workspace is needed when k_splits > 1.
static size_t store_gemm_slab_bytes(int M, int N, int K, int ctas, int forced_ksplit, int mrep = 1) {
  if (N % 4 != 0) return 0;
  if (ctas != 2) mrep = 1;
  const int tiles = ceil_div(M, GEMM_BLOCK_M * mrep * ctas) * ceil_div(N, BN);
  if (forced_ksplit <= 0 && get_tunable(TUNE_TAIL_SPLIT))
    return tail_slab_bytes(tiles, clusters_for(ctas), ceil_div(K, GEMM_BLOCK_K), GEMM_BLOCK_M * mrep * ctas);
  const int s = choose_ksplits(tiles, clusters_for(ctas), ceil_div(K, GEMM_BLOCK_K), forced_ksplit);
  return s > 1 ? align_up((size_t)s * M * N * 4, 256) : 0;
}

// a_panel_rows > 0: A is a dlogits panel buffer [ceil(cols/256)][a_panel_rows][256] (lda unused)
static int run_store_gemm(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, void* C, long long ldc,
                          int c_is_f32, int accumulate, int M, int N, int K, int ctas, int group_m, int forced_ksplit,
                          void* slab, size_t slab_bytes, cudaStream_t st, int a_panel_rows = 0, int group_n = 0,
                          int hint_a = 0, int hint_b = 0, int mrep = 1, int zigzag = 0, const RowEpiArgs* epi = nullptr) {
  GemmParams p = {};
  p.zigzag = zigzag;
  if (epi) {
    p.row_scale = epi->row_scale; p.row_coef = epi->row_coef; p.gather_src = epi->gather_src; p.gather_ld = epi->gather_ld;
    p.gather_rows = epi->gather_rows; p.targets = epi->targets; p.v_offset = epi->v_offset;
    if (forced_ksplit > 1) forced_ksplit = 0;   // the uniform split-K reduce does not carry the row epilogue: tail split instead
  }
  if (ctas != 2) mrep = 1;   // the 256-row variant exists for CTA pairs only
  const int tiles = ceil_div(M, GEMM_BLOCK_M * mrep * ctas) * ceil_div(N, BN);
  // forced_ksplit >= 1: uniform k-splits (1 = none); <= 0: automatic = tail split of the last partial wave
  // (or, with the tail_split tunable off, the uniform split count with the best wave efficiency)
  // (a row epilogue travels with the tail split only -- tail_reduce_kernel applies it, splitk_reduce_kernel does not -- so it
  // takes the tail split whatever the tunable says; found by the tunable-randomising fuzz: dH came back un-scaled)
  const bool want_tail = forced_ksplit <= 0 && (get_tunable(TUNE_TAIL_SPLIT) || epi != nullptr);
  int ks = (N % 4 == 0 && slab && !want_tail) ? choose_ksplits(tiles, clusters_for(ctas), ceil_div(K, GEMM_BLOCK_K), forced_ksplit) : 1;
  if (epi) ks = 1;
  if (ks > 1 && slab_bytes < (size_t)ks * M * N * 4) ks = 1;
  fill_tiling(p, M, N, K, ctas, group_m, ks, group_n, hint_a, hint_b, mrep);
  int tail = 0;
  if (want_tail && slab && N % 4 == 0) {
    const int clusters = clusters_for(ctas);
    tail = tail_split_for(tiles, clusters, p.kb_total);
    if (tail && slab_bytes < tail_slab_bytes(tiles, clusters, p.kb_total, p.tile_rows)) tail = 0;
    if (tail) {
      const int r = tiles % clusters;
      p.tail_first = tiles - r;
      p.tail_split = tail;
      p.n_items = p.tail_first + r * tail;
    }
  }
  p.C = C; p.ldc = ldc; p.c_is_f32 = c_is_f32; p.accumulate = accumulate; p.slab = (float*)slab;
  CUtensorMap ta, tb;
  if (a_panel_rows > 0) {
    const int cols = a_mn ? M : K;  // the panelled (vocab) axis
    if (int e = make_tmap_panel(&ta, A, a_panel_rows, ceil_div(cols, PANEL_W), a_mn ? GEMM_BLOCK_K : GEMM_BLOCK_M * mrep)) return e;
    if (int e = (!b_mn ? make_tmap(&tb, B, K, N, ldb, BN / ctas) : make_tmap(&tb, B, N, K, ldb, GEMM_BLOCK_K))) return e;
    if (int e = dispatch<EPI_STORE, false, true>(a_mn ? 1 : 0, b_mn ? 1 : 0, ctas, ta, tb, p, st, mrep)) return e;
  } else {
    if (int e = make_ab_maps(&ta, &tb, A, lda, a_mn, B, ldb, b_mn, M, N, K, ctas, mrep)) return e;
    if (int e = dispatch<EPI_STORE>(a_mn ? 1 : 0, b_mn ? 1 : 0, ctas, ta, tb, p, st, mrep)) return e;
  }
  if (tail) {
    dim3 grid((p.n_items - p.tail_first) / tail, p.tile_rows / TAIL_RED_ROWS);   // x: tail tiles
    tail_reduce_kernel<<<grid, 256, 0, st>>>(p); count_launch();
    B200_CUDA_OK(cudaGetLastError());
  }
  if (ks > 1) {
    const long long quads = (long long)M * (N / 4);
    long long grid = (quads + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (grid > cap) grid = cap;
    splitk_reduce_kernel<<<(int)grid, 256, 0, st>>>((const float*)slab, ks, M, N, C, ldc, c_is_f32, accumulate); count_launch();
    B200_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

}  // namespace b200grpo

using namespace b200grpo;

static size_t panel_bytes(int M, int Vloc) { return align_up((size_t)ceil_div(Vloc, PANEL_W) * M * PANEL_W * 2, 256); }
// zero the last panel when the vocab is not a multiple of 256: its unwritten tail columns are K-padding of dH
static int clear_tail_panel(void* panels, int M, int Vloc, cudaStream_t st) {
  if (Vloc % PANEL_W == 0) return 0;
  const size_t one = (size_t)M * PANEL_W * 2;
  B200_CUDA_OK(cudaMemsetAsync((char*)panels + (size_t)(ceil_div(Vloc, PANEL_W) - 1) * one, 0, one, st));
  return 0;
}
static size_t part_bytes(int M, int Vloc) { return align_up((size_t)ceil_div(Vloc, BN) * M * 16, 256); }
static size_t ref_bytes(int M) { return align_up((size_t)M * 4, 256); }

// forward scratch: [per-tile partials][row references]
extern "C" size_t b200grpo_lmhead_fwd_workspace(int M, int H, int Vloc) {
  (void)H;
  return part_bytes(M, Vloc) + ref_bytes(M);
}

// saved-forward state for the recompute-free backward: [per-tile partials][row references r][bf16 E = exp(logit/T - r) panels]
extern "C" size_t b200grpo_lmhead_saved_bytes(int M, int H, int Vloc) {
  (void)H;
  return part_bytes(M, Vloc) + ref_bytes(M) + panel_bytes(M, Vloc);
}

static int lmhead_fwd_impl(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                           int M, int H, int Vloc, int64_t v_offset, float inv_temperature, float* stats, bool stats_multicast,
                           void* saved, void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(hidden && weight && targets && stats && (workspace || saved), "lmhead_fwd: null pointer");
  B200_REQUIRE(M >= 1 && H >= 8 && Vloc >= 1 && H % 8 == 0, "lmhead_fwd: M=%d H=%d Vloc=%d (H must be a multiple of 8)", M, H, Vloc);
  B200_REQUIRE(inv_temperature > 0.f, "lmhead_fwd: temperature must be positive");
  B200_REQUIRE(!saved || ((uintptr_t)saved & 255) == 0, "lmhead_fwd: saved buffer must be 256-byte aligned");
  const int ctas = ctas_for(M);
  const int mrep = mrep_for(TUNE_FWD_BIG_M, ctas);
  GemmParams p = {};
  fill_tiling(p, M, Vloc, H, ctas, max(1, get_tunable(TUNE_FWD_GROUP_M) / mrep), 1, 0, get_tunable(TUNE_HINT_RESIDENT),
              get_tunable(TUNE_HINT_STREAM), mrep);
  const size_t n_part = part_bytes(M, Vloc), need = n_part + ref_bytes(M);
  if (!saved && workspace_bytes < need) { set_error("lmhead_fwd: workspace %zu < %zu", workspace_bytes, need); return B200GRPO_EWORKSPACE; }
  p.targets = (const long long*)targets; p.v_offset = v_offset; p.inv_temp = inv_temperature;
  char* base = (char*)(saved ? saved : workspace);
  p.part = (float*)base;
  float* row_ref = (float*)(base + n_part);
  p.row_ref = row_ref;
  cudaStream_t st = (cudaStream_t)stream;
  if (saved) {
    p.zsave = (__nv_bfloat16*)(base + need);
    if (int e = clear_tail_panel(p.zsave, M, Vloc, st)) return e;
  }
  CUtensorMap ta, tb;
  if (int e = make_ab_maps(&ta, &tb, hidden, ldh, 0, weight, ldw, 0, M, Vloc, H, ctas, mrep)) return e;
  {
    ProfScope ps(PROF_FWD_GEMM, st);
    row_ref_kernel<<<ceil_div(M, 8), 256, 0, st>>>((const __nv_bfloat16*)hidden, ldh, (const __nv_bfloat16*)weight, ldw,
                                                   (const long long*)targets, M, H, Vloc, v_offset, inv_temperature, row_ref);
    count_launch();
    const int e = saved ? dispatch<EPI_FWD_STATS, true>(0, 0, ctas, ta, tb, p, st, mrep)
                        : dispatch<EPI_FWD_STATS, false>(0, 0, ctas, ta, tb, p, st, mrep);
    if (e) return e;
  }
  ProfScope ps(PROF_FWD_MERGE, st);
  if (stats_multicast) fwd_merge_kernel<true><<<ceil_div(M, 32), 32 * MG_SLICES, 0, st>>>((const float4*)p.part, row_ref, p.n_blocks, M, (float4*)stats);
  else fwd_merge_kernel<false><<<ceil_div(M, 32), 32 * MG_SLICES, 0, st>>>((const float4*)p.part, row_ref, p.n_blocks, M, (float4*)stats);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_lmhead_fwd(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                                   int M, int H, int Vloc, int64_t v_offset, float inv_temperature, float* stats, void* saved,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  return lmhead_fwd_impl(hidden, ldh, weight, ldw, targets, M, H, Vloc, v_offset, inv_temperature, stats, false, saved, workspace,
                         workspace_bytes, stream);
}

extern "C" int b200grpo_lmhead_fwd_multicast(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                                             int M, int H, int Vloc, int64_t v_offset, float inv_temperature, void* stats_mc, void* saved,
                                             void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(((uintptr_t)stats_mc & 15) == 0, "lmhead_fwd_multicast: the multicast slot must be 16-byte aligned");
  return lmhead_fwd_impl(hidden, ldh, weight, ldw, targets, M, H, Vloc, v_offset, inv_temperature, (float*)stats_mc, true, saved,
                         workspace, workspace_bytes, stream);
}

extern "C" int b200grpo_lmhead_finalize(const float* stats, int R, int M, float* logprob, float* entropy, float* lse, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(stats && logprob && entropy && R >= 1 && M >= 1, "lmhead_finalize: bad argument");
  ProfScope ps(PROF_FINALIZE, (cudaStream_t)stream);
  finalize_kernel<<<ceil_div(M, 256), 256, 0, (cudaStream_t)stream>>>((const float4*)stats, R, M, logprob, entropy, lse); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

// backward scratch: [dlogits panels (recompute mode only)][row scale, row coefficient, scaled hidden (saved mode: the
// entropy-free path)][tail-split / split-K slab]
static size_t key_bytes(int M) { return align_up((size_t)ceil_div(M, DWT_TRIP) * DWT_TRIP * 4, 256); }   // padded for the 512-key trips of dw_target_kernel
static size_t rows_scratch_bytes(int M, int H) { return 2 * align_up((size_t)M * 4, 256) + key_bytes(M) + align_up((size_t)M * H * 2, 256); }

extern "C" size_t b200grpo_lmhead_bwd_workspace(int M, int H, int Vloc, int with_saved) {
  const int ctas = ctas_for(M);
  size_t slab = store_gemm_slab_bytes(M, H, Vloc, ctas, get_tunable(TUNE_DH_KSPLIT), mrep_for(TUNE_DH_BIG_M, ctas));
  size_t slab_t = tail_slab_bytes(ceil_div(M, GEMM_BLOCK_M * mrep_for(TUNE_DH_BIG_M, ctas) * ctas) * ceil_div(H, BN), clusters_for(ctas),
                                  ceil_div(Vloc, GEMM_BLOCK_K), GEMM_BLOCK_M * mrep_for(TUNE_DH_BIG_M, ctas) * ctas);   // the row-epilogue path forces the tail split
  if (slab_t > slab) slab = slab_t;
  const size_t slab_dw = store_gemm_slab_bytes(Vloc, H, M, ctas, get_tunable(TUNE_DW_KSPLIT), mrep_for(TUNE_DW_BIG_M, ctas));
  if (slab_dw > slab) slab = slab_dw;   // the two GEMMs run back to back on one stream and share it
  return (with_saved ? rows_scratch_bytes(M, H) : panel_bytes(M, Vloc)) + slab + 256;
}

// stages: B200GRPO_BWD_DLOGITS (prepare: per-row scales + scaled hidden for the entropy-free path; dlogits in place from the
// saved exponentials when there is an entropy term; or the recompute GEMM into the workspace), _DH, _DW.  A caller that
// splits the backward in two calls passes the same saved / workspace buffers to both and B200GRPO_BWD_DLOGITS only to the first.
// coef_ent == nullptr declares "no gradient through the entropy" (entropy_bonus == 0): with saved exponentials the backward
// then never forms dlogits (see bwd_rows_kernel).
static int lmhead_bwd_impl(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                           const float* lse, const float* entropy, const float* coef_lp, const float* coef_ent,
                           const float* grad_scale, int M, int H, int Vloc, int64_t v_offset, float inv_temperature,
                           void* saved, void* grad_hidden, int64_t ldgh, void* grad_weight, int64_t ldgw,
                           int grad_weight_is_f32, int accumulate, int stages, void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  const bool form = stages & B200GRPO_BWD_DLOGITS;
  if (!(stages & B200GRPO_BWD_DH)) grad_hidden = nullptr;
  if (!(stages & B200GRPO_BWD_DW)) grad_weight = nullptr;
  const bool rows_path = saved && !coef_ent && get_tunable(TUNE_ROWS_PATH);   // entropy-free + saved exponentials: GEMMs straight on E
  const bool fold_onehot = rows_path && get_tunable(TUNE_ROWS_PATH) >= 2;     // confident targets: onehot folded into the saved exponential (bwd_rows_kernel)
  B200_REQUIRE(hidden && weight && workspace && targets, "lmhead_bwd: null pointer");
  B200_REQUIRE(!form || (lse && coef_lp && ((saved && !coef_ent) || (entropy && coef_ent))), "lmhead_bwd: null per-token input");
  B200_REQUIRE(saved || coef_ent || !form, "lmhead_bwd: the recompute backward needs coef_ent (pass zeros)");
  B200_REQUIRE(M >= 1 && H >= 8 && Vloc >= 1 && H % 8 == 0, "lmhead_bwd: M=%d H=%d Vloc=%d", M, H, Vloc);
  B200_REQUIRE(!accumulate || grad_weight_is_f32, "lmhead_bwd: accumulate needs an f32 grad_weight");
  const size_t need = b200grpo_lmhead_bwd_workspace(M, H, Vloc, saved != nullptr);
  if (workspace_bytes < need) { set_error("lmhead_bwd: workspace %zu < %zu", workspace_bytes, need); return B200GRPO_EWORKSPACE; }
  const int ctas = ctas_for(M);
  cudaStream_t st = (cudaStream_t)stream;
  __nv_bfloat16* dlog;
  char* slab;
  float *row_scale = nullptr, *row_coef = nullptr;
  int* row_key = nullptr;
  __nv_bfloat16* hs = nullptr;
  const float* row_ref = nullptr;
  if (saved) {
    row_ref = (const float*)((char*)saved + part_bytes(M, Vloc));
    dlog = (__nv_bfloat16*)((char*)saved + part_bytes(M, Vloc) + ref_bytes(M));
    row_scale = (float*)workspace;
    row_coef = (float*)((char*)workspace + align_up((size_t)M * 4, 256));
    row_key = (int*)((char*)workspace + 2 * align_up((size_t)M * 4, 256));
    hs = (__nv_bfloat16*)((char*)workspace + 2 * align_up((size_t)M * 4, 256) + key_bytes(M));
    slab = (char*)workspace + rows_scratch_bytes(M, H);
  } else {
    dlog = (__nv_bfloat16*)workspace;
    slab = (char*)workspace + panel_bytes(M, Vloc);
  }
  if (saved && !coef_ent && !rows_path) {   // A/B switch: the entropy path with zero entropy coefficients (row_coef region as zeros)
    if (form) B200_CUDA_OK(cudaMemsetAsync(row_coef, 0, (size_t)M * 4, st));
    coef_ent = row_coef;
    if (!entropy) entropy = row_coef;
  }
  if (rows_path && form) {   // 1a. per-row scale s = -a exp(r - lse), coefficient a, scaled copy of hidden (operand B of dW)
    ProfScope ps(PROF_DLOGITS_EW, st);
    B200_CUDA_OK(cudaMemsetAsync(row_key + M, 0xfe, key_bytes(M) - (size_t)M * 4, st));   // pad keys: 0xfefefefe < 0, never a match
    bwd_rows_kernel<<<ceil_div(M, 8), 256, 0, st>>>((const __nv_bfloat16*)hidden, ldh, row_ref, lse, coef_lp, grad_scale, inv_temperature,
                                                    M, H, (const long long*)targets, v_offset, Vloc, row_scale, row_coef, row_key, hs,
                                                    fold_onehot ? dlog : nullptr);
    count_launch();
    B200_CUDA_OK(cudaGetLastError());
  } else if (saved && form) {  // 1b. entropy term present: saved exponentials -> dlogits in place (no recompute GEMM)
    ProfScope ps(PROF_DLOGITS_EW, st);
    dim3 grid(ceil_div(M, DL_ROWS), ceil_div(Vloc, PANEL_W));
    dlogits_from_saved_kernel<<<grid, DL_THREADS, 0, st>>>(dlog, row_ref, M, Vloc, (const long long*)targets, v_offset, lse,
                                                           entropy, coef_lp, coef_ent, grad_scale, inv_temperature); count_launch();
    B200_CUDA_OK(cudaGetLastError());
  } else if (form) {      // 1c. recompute logits tile by tile -> bf16 dlogits
    if (int e = clear_tail_panel(dlog, M, Vloc, st)) return e;
    GemmParams p = {};
    fill_tiling(p, M, Vloc, H, ctas, get_tunable(TUNE_FWD_GROUP_M), 1, 0, get_tunable(TUNE_HINT_RESIDENT), get_tunable(TUNE_HINT_STREAM));
    p.targets = (const long long*)targets; p.v_offset = v_offset; p.inv_temp = inv_temperature;
    p.lse = lse; p.entropy = entropy; p.coef_lp = coef_lp; p.coef_ent = coef_ent; p.grad_scale = grad_scale;
    p.dlogits = dlog;
    CUtensorMap ta, tb;
    if (int e = make_ab_maps(&ta, &tb, hidden, ldh, 0, weight, ldw, 0, M, Vloc, H, ctas)) return e;
    ProfScope ps(PROF_BWD_DLOGITS, st);
    if (int e = dispatch<EPI_BWD_DLOGITS>(0, 0, ctas, ta, tb, p, st)) return e;
  }
  const size_t slab_bytes = (char*)workspace + workspace_bytes - slab;
  if (grad_hidden) {  // 2. dH[M,H] = dlogits[M,V] . W[V,H]   (B is MN-major: vocab rows are the K axis)
    ProfScope ps(PROF_BWD_DH, st);
    RowEpiArgs epi;   // rows path: dH = diag(s) . (E . W) + a * W[target]
    epi.row_scale = row_scale; epi.row_coef = row_coef; epi.gather_src = (const __nv_bfloat16*)weight; epi.gather_ld = ldw;
    epi.gather_rows = Vloc; epi.targets = (const long long*)targets; epi.v_offset = v_offset;
    if (int e = run_store_gemm(dlog, 0, 0, weight, ldw, 1, grad_hidden, ldgh, 0, 0, M, H, Vloc, ctas,
                               get_tunable(TUNE_DH_GROUP_M), get_tunable(TUNE_DH_KSPLIT), slab, slab_bytes, st, M, 0,
                               get_tunable(TUNE_HINT_STREAM), get_tunable(TUNE_HINT_STREAM), mrep_for(TUNE_DH_BIG_M, ctas), 0,
                               rows_path ? &epi : nullptr)) return e;
  }
  if (grad_weight) {  // 3. dW[V,H] = dlogits^T[V,M] . hidden[M,H]   (both operands MN-major: tokens are the K axis)
    ProfScope ps(PROF_BWD_DW, st);
    const void* b_op = rows_path ? (const void*)hs : hidden;        // rows path: dW = E^T . (diag(s) . hidden) + onehot term
    if (int e = run_store_gemm(dlog, 0, 1, b_op, rows_path ? (long long)H : ldh, 1, grad_weight, ldgw, grad_weight_is_f32, accumulate, Vloc, H, M,
                               ctas, get_tunable(TUNE_DW_GROUP_M), get_tunable(TUNE_DW_KSPLIT), slab, slab_bytes, st, M, get_tunable(TUNE_DW_GROUP_N),
                               get_tunable(TUNE_HINT_STREAM), get_tunable(TUNE_HINT_RESIDENT), mrep_for(TUNE_DW_BIG_M, ctas), get_tunable(TUNE_DW_ZIGZAG))) return e;
    if (rows_path) {
      int warps = 4;                                          // f32 row per warp in shared memory: 4 x H x 4 B (57 KB at H = 3584: three blocks per SM)
      const size_t cap = 200 * 1024;
      while (warps > 1 && (size_t)warps * H * 4 > cap) warps >>= 1;
      B200_REQUIRE((size_t)warps * H * 4 <= cap, "lmhead_bwd: hidden size %d too large for the target-row kernel", H);
      static PerDeviceOnce configured;
      if (configured.first()) {
        B200_CUDA_OK(cudaFuncSetAttribute(dw_target_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap));
      }
      dw_target_kernel<<<ceil_div(M, warps), warps * 32, (size_t)warps * H * 4, st>>>(
          (const __nv_bfloat16*)hidden, ldh, row_key, row_coef, M, H, grad_weight, ldgw, grad_weight_is_f32); count_launch();
      B200_CUDA_OK(cudaGetLastError());
    }
  }
  return B200GRPO_OK;
}

extern "C" int b200grpo_lmhead_bwd(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                                   const float* lse, const float* entropy, const float* coef_lp, const float* coef_ent,
                                   const float* grad_scale, int M, int H, int Vloc, int64_t v_offset, float inv_temperature,
                                   void* saved, void* grad_hidden, int64_t ldgh, void* grad_weight, int64_t ldgw,
                                   int grad_weight_is_f32, int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
  return lmhead_bwd_impl(hidden, ldh, weight, ldw, targets, lse, entropy, coef_lp, coef_ent, grad_scale, M, H, Vloc, v_offset,
                         inv_temperature, saved, grad_hidden, ldgh, grad_weight, ldgw, grad_weight_is_f32, accumulate,
                         B200GRPO_BWD_DLOGITS | B200GRPO_BWD_DH | B200GRPO_BWD_DW, workspace, workspace_bytes, stream);
}

extern "C" int b200grpo_lmhead_bwd_staged(const void* hidden, int64_t ldh, const void* weight, int64_t ldw, const int64_t* targets,
                                          const float* lse, const float* entropy, const float* coef_lp, const float* coef_ent,
                                          const float* grad_scale, int M, int H, int Vloc, int64_t v_offset, float inv_temperature,
                                          void* saved, void* grad_hidden, int64_t ldgh, void* grad_weight, int64_t ldgw,
                                          int grad_weight_is_f32, int accumulate, int stages, void* workspace,
                                          size_t workspace_bytes, void* stream) {
  B200_REQUIRE(stages > 0 && stages < 8, "lmhead_bwd_staged: stages=%d", stages);
  return lmhead_bwd_impl(hidden, ldh, weight, ldw, targets, lse, entropy, coef_lp, coef_ent, grad_scale, M, H, Vloc, v_offset,
                         inv_temperature, saved, grad_hidden, ldgh, grad_weight, ldgw, grad_weight_is_f32, accumulate, stages,
                         workspace, workspace_bytes, stream);
}

extern "C" int b200grpo_gemm_bf16(const void* A, int64_t lda, int a_mn_major, const void* B, int64_t ldb, int b_mn_major, void* C,
                                  int64_t ldc, int c_is_f32, int accumulate, int M, int N, int K, int group_m, int k_splits,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(A && B && C && M >= 1 && N >= 1 && K >= 1, "gemm: bad argument");
  B200_REQUIRE(!accumulate || c_is_f32, "gemm: accumulate needs f32 C");
  B200_REQUIRE(k_splits <= 1 || (workspace && workspace_bytes >= (size_t)k_splits * M * N * 4 && N % 4 == 0),
               "gemm: k_splits=%d needs a workspace of k_splits*M*N*4 bytes and N %% 4 == 0", k_splits);
  const int ctas = ctas_for(M);
  ProfScope ps(PROF_GEMM, (cudaStream_t)stream);
  return run_store_gemm(A, lda, a_mn_major ? 1 : 0, B, ldb, b_mn_major ? 1 : 0, C, ldc, c_is_f32, accumulate, M, N, K, ctas, group_m,
                        k_splits, k_splits != 1 ? workspace : nullptr, workspace_bytes, (cudaStream_t)stream, 0, 0, 0, 0,
                        mrep_for(TUNE_GEMM_BIG_M, ctas));
}
