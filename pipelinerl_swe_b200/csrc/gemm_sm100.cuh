// Persistent warp-specialised tcgen05 GEMM mainloop for sm_100a with pluggable epilogues.
//
//   D[M,N] (f32 in TMEM) = A[M,K] . B[N,K]^T          bf16 operands, fp32 accumulate
//
// One CTA owns 128 * kMRep rows x BLOCK_N columns of a tile (a cta_group::2 pair: twice the rows):
//   warp 0 lane 0 : TMA producer of operand A (cp.async.bulk.tensor -> 128B-swizzled smem ring)
//   warp 3 lane 0 : TMA producer of operand B
//   warp 1 lane 0 : MMA issuer    (tcgen05.mma; kMRep MMAs per k-step share one B tile in shared memory)
//   warp 2        : TMEM allocator
//   warps 4..7    : epilogue of accumulator rows 0-127   (tcgen05.ld: thread == row, 32 columns per load)
//   warps 8..11   : epilogue of accumulator rows 128-255 (kMRep == 2 only)
// Pipelines: smem full/empty mbarriers (TMA <-> MMA) and TMEM full/empty mbarriers (MMA <-> epilogue).
// kMRep == 1: two accumulator stages, the epilogue of tile i overlaps the MMAs of tile i+1.
// kMRep == 2: one accumulator stage (all 512 TMEM columns), 25 % fewer operand bytes per FLOP into the SM;
//             for long-K GEMMs with a light epilogue (dH).
//
// Work decomposition: items dealt round-robin to the resident clusters.  An item is an output tile, a
// (tile, k-split) pair (uniform split-K), or -- tail split -- a K-piece of one of the tiles of the last partial
// wave (tiles % clusters), cut so that the last wave costs 1/pieces of a full one (decode_item).  Tiles are
// rasterised in groups of `group_m` row blocks (row block fastest inside a group) so that the clusters running
// together share operand tiles through L2:
//   lm_head fwd / recompute : group_m = 16 row blocks -> the 74 live tiles read the same 4-5 weight tiles
//                             while their slice of hidden stays L2 resident;
//   dW (tokens are K)       : group_m = 1 -> one dlogits column slab is shared by the 14 n-tiles of its row
//                             block (fewest bytes in flight per wave, profiles/r01/tile_variants_ab.md);
//   dH (vocab is K)         : square-ish patches (both operands stream, nothing is resident).
// Every epilogue is stateless per tile (the forward writes per-tile softmax partials that a merge kernel
// combines), which is what makes the tile order free.  Split pieces (EPI_STORE only) go to f32 slabs summed in
// fixed order by a second kernel (deterministic).
#pragma once
#include "common.cuh"
#include "sm100_ptx.cuh"
#include <type_traits>

namespace b200grpo {

enum { EPI_STORE = 0, EPI_FWD_STATS = 1, EPI_BWD_DLOGITS = 2 };
enum { MAJOR_K = 0, MAJOR_MN = 1 };

constexpr int GEMM_BLOCK_M = 128;   // rows per CTA (TMEM lanes)
constexpr int GEMM_BLOCK_K = 64;    // one 128-byte swizzle atom of bf16
constexpr int GEMM_UMMA_K = 16;
constexpr int GEMM_THREADS = 256;
constexpr int GEMM_EPI_WARP0 = 4;
constexpr float LOG2E_F = 1.4426950408889634f;
constexpr float LN2_F = 0.6931471805599453f;
constexpr int PANEL_W = 256;        // saved logits / dlogits are stored in [M x 256] column panels (== BLOCK_N)

struct GemmParams {
  int M, N, K;
  int m_blocks, n_blocks, k_splits, group_m, group_n, kb_total;
  int n_items;              // work items of the launch: full tiles x k_splits, or full-K tiles + tail pieces
  int tail_first;           // tail split: items >= tail_first are K-pieces of the last (tiles % clusters) tiles
  int tail_split;           //   pieces per tail tile (>= 2 when the tail split is on, else 0)
  int tile_rows;            //   rows of one tile (all CTAs of the cluster)
  unsigned long long hint_a, hint_b;   // L2 eviction policy of the A / B operand loads
  int zigzag;               // odd tiles of a cluster walk K backwards: the operand rows touched last by one wave are the first
                            // the next wave asks L2 for (K-long operands that do not fit L2 whole, i.e. dW)
  // soft K-lockstep (optional): the clusters of a wave pace their operand loads on shared progress counters so that tiles
  // which share operand rows ask L2 for them within its residency window (see launch_gemm)
  unsigned* sync;           // [waves][sync_epochs] arrival counters, zeroed before the launch; nullptr = off
  int sync_kb;              //   k-blocks per epoch
  int sync_epochs;          //   epochs per full-K tile
  int sync_full;            //   full-K items of the launch (items [0, sync_full) are dealt wave by wave)
  int sync_slack;           //   epochs a cluster may lead the slowest cluster of its wave by
  // EPI_STORE
  void* C;
  long long ldc;
  int c_is_f32, accumulate;
  float* slab;              // [k_splits][M][N] f32 partial sums when k_splits > 1;
                            // [tail piece][tile_rows][256] when the tail split is on
  // EPI_STORE, optional row epilogue (the entropy-free backward: dH = diag(row_scale) . (E . W) + row_coef * W[target]):
  const float* row_scale;          // [M] multiplies the accumulator row (nullptr: 1)
  const float* row_coef;           // [M] coefficient of the gathered row (nullptr: no gather)
  const __nv_bfloat16* gather_src; // [*, gather_ld] rows indexed by targets[row] - v_offset (skipped when outside [0, gather_rows))
  long long gather_ld;
  int gather_rows;
  // EPI_FWD_STATS / EPI_BWD_DLOGITS
  const long long* targets;
  long long v_offset;
  float inv_temp;
  const float* row_ref;     // EPI_FWD_STATS: [M] per-row reference r (in units of z/T); exponentials are taken relative to it
  float* part;              // [n_blocks][M][4]  per-tile (unused, sum e, sum e*z/T.., z_target|0) with e = exp(z/T - r)
  __nv_bfloat16* zsave;     // optional panels [n_blocks][M][256]: E = exp(z/T - r) saved for a recompute-free backward
  const float *lse, *entropy, *coef_lp, *coef_ent, *grad_scale;
  __nv_bfloat16* dlogits;   // panels [n_blocks][M][256]
};

template <int BLOCK_N, int kCtas, int kStages, int kMRep = 1>
struct GemmSmem {
  static constexpr int LOAD_N = BLOCK_N / kCtas;
  static constexpr int A_BYTES = GEMM_BLOCK_M * kMRep * GEMM_BLOCK_K * 2;
  static constexpr int B_BYTES = LOAD_N * GEMM_BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = kStages * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + (2 * kStages + 4) * 8 + 16 + 1024;  // + alignment slack
};

struct Item { int m_blk, nb, ks, kb0, kb1, slot; };

// tile number -> (row block, column block).  Order: column group of group_n tile columns (slowest) > row group of
// group_m tile rows > tile column inside the group > tile row (fastest)
__host__ __device__ __forceinline__ void decode_tile(const GemmParams& p, int t, int& m_blk, int& nb) {
  const int per_ngroup = p.m_blocks * p.group_n;
  const int ng = t / per_ngroup;
  t -= ng * per_ngroup;
  const int n_first = ng * p.group_n;
  const int gn = p.group_n < p.n_blocks - n_first ? p.group_n : p.n_blocks - n_first;
  const int per_group = p.group_m * gn;
  const int g = t / per_group, w = t - g * per_group;
  const int m_first = g * p.group_m;
  const int gm = p.group_m < p.m_blocks - m_first ? p.group_m : p.m_blocks - m_first;
  const int nb_local = w / gm;
  nb = n_first + nb_local;
  m_blk = m_first + (w - nb_local * gm);
}

// Work items: [0, tiles * k_splits) are (k-split, tile) pairs, k-split slowest.  With the tail split on
// (k_splits == 1), items [0, tail_first) are full-K tiles -- whole waves of the persistent grid -- and item
// tail_first + j is K-piece j % tail_split of tile tail_first + j / tail_split: the last partial wave is cut
// along K so that every cluster gets an equal share; pieces land in the slab and tail_reduce_kernel sums them.
__device__ __forceinline__ Item decode_item(const GemmParams& p, int item) {
  Item it;
  if (item >= p.tail_first) {
    const int j = item - p.tail_first;
    const int tt = j / p.tail_split, piece = j - tt * p.tail_split;
    decode_tile(p, p.tail_first + tt, it.m_blk, it.nb);
    it.ks = 0;
    it.slot = j;
    it.kb0 = (int)((long long)piece * p.kb_total / p.tail_split);
    it.kb1 = (int)((long long)(piece + 1) * p.kb_total / p.tail_split);
    return it;
  }
  const int tiles = p.m_blocks * p.n_blocks;
  it.ks = item / tiles;
  it.slot = -1;
  decode_tile(p, item - it.ks * tiles, it.m_blk, it.nb);
  it.kb0 = (int)((long long)it.ks * p.kb_total / p.k_splits);
  it.kb1 = (int)((long long)(it.ks + 1) * p.kb_total / p.k_splits);
  return it;
}

// ---------------------------------------------------------------------------------------
// Epilogues.  Each thread owns one accumulator row; v[] holds 32 consecutive columns.
// ---------------------------------------------------------------------------------------
// MUFU.EX2 directly: 2^x with ~2^-22 relative error, ex2(-inf) = +0 (no range fix-up code)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct FwdRowState { float neg2, s, u, tgt; int tcol; };

// Forward epilogue, ONE pass: exponentials are taken relative to a per-row reference r known before the GEMM (a cheap lower
// bound of the row maximum: the largest of the target logit and a few sampled logits, lmhead.cu row_ref_kernel), so no tile
// maximum is needed: e = exp(z/T - r) accumulates sum e and sum e*z, the target logit is picked, and (optionally) e is saved
// as bf16 -- bf16 has f32's exponent range, so e keeps 2^-9 relative precision whatever its magnitude, and every quantity
// the backward needs is a ROW-scaled multiple of it: p = e * exp(r - lse).
template <bool kSave>
__device__ __forceinline__ void fwd_chunk(FwdRowState& st, const uint32_t (&v)[32], int col0, int n_valid, float scale2,
                                          __nv_bfloat16* zout) {  // zout: this row's 32 columns in the panel
  float s = 0.f, u = 0.f;
  uint32_t packed[16];
  if (n_valid >= 32) {
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      const float a0 = __uint_as_float(v[j]), a1 = __uint_as_float(v[j + 1]);
      const float e0 = fast_exp2(fmaf(a0, scale2, st.neg2)), e1 = fast_exp2(fmaf(a1, scale2, st.neg2));
      s += e0; u = fmaf(e0, a0, u);
      s += e1; u = fmaf(e1, a1, u);
      if constexpr (kSave) {
        const __nv_bfloat162 b = __floats2bfloat162_rn(e0, e1);
        packed[j >> 1] = *reinterpret_cast<const uint32_t*>(&b);
      }
    }
  } else {  // vocab tail: columns past N contribute nothing and are saved as zeros (they are K-padding of the dH GEMM)
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
      const float a0 = __uint_as_float(v[j]), a1 = __uint_as_float(v[j + 1]);
      const float e0 = j < n_valid ? fast_exp2(fmaf(a0, scale2, st.neg2)) : 0.f;
      const float e1 = j + 1 < n_valid ? fast_exp2(fmaf(a1, scale2, st.neg2)) : 0.f;
      s += e0; u = fmaf(e0, j < n_valid ? a0 : 0.f, u);
      s += e1; u = fmaf(e1, j + 1 < n_valid ? a1 : 0.f, u);
      if constexpr (kSave) {
        const __nv_bfloat162 b = __floats2bfloat162_rn(e0, e1);
        packed[j >> 1] = *reinterpret_cast<const uint32_t*>(&b);
      }
    }
  }
  st.s += s; st.u += u;
  const int t = st.tcol - col0;
  if (t >= 0 && t < 32) {
#pragma unroll
    for (int j = 0; j < 32; ++j) if (j == t) st.tgt = __uint_as_float(v[j]);
  }
  if constexpr (kSave) {
    uint4* o = reinterpret_cast<uint4*>(zout);
#pragma unroll
    for (int q = 0; q < 4; ++q) o[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
  }
}

struct BwdRowConst { float lse2, a, e, ent; int tcol; };

__device__ __forceinline__ void bwd_chunk(const BwdRowConst& rc, const uint32_t (&v)[32], int col0, int n_valid,
                                          float scale2, __nv_bfloat16* out) {  // out: this row's 32 columns in the panel
  uint32_t packed[16];
  const int t = rc.tcol - col0;
#pragma unroll
  for (int j = 0; j < 32; j += 2) {
    float d[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const float y2 = fmaf(__uint_as_float(v[j + q]), scale2, -rc.lse2);  // log2 p
      const float pr = fast_exp2(y2);
      float g = -pr * fmaf(rc.e, fmaf(y2, LN2_F, rc.ent), rc.a);           // -p*(a + e*(logp + H))
      if (j + q == t) g += rc.a;
      d[q] = g;
    }
    const __nv_bfloat162 b = __floats2bfloat162_rn(d[0], d[1]);
    packed[j >> 1] = *reinterpret_cast<const uint32_t*>(&b);
  }
  if (n_valid < 32) {  // vocab tail: columns past N stay zero (K-padding of the dH GEMM)
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j >= n_valid) packed[j >> 1] &= (j & 1) ? 0x0000ffffu : 0xffff0000u;
  }
  uint4* o = reinterpret_cast<uint4*>(out);
#pragma unroll
  for (int q = 0; q < 4; ++q) o[q] = make_uint4(packed[4 * q], packed[4 * q + 1], packed[4 * q + 2], packed[4 * q + 3]);
}

// Row epilogue of the entropy-free backward: x <- row_scale[row] * x + row_coef[row] * gather_src[target(row)][col]
struct RowEpi { float scale, coef; const __nv_bfloat16* src; };
__device__ __forceinline__ RowEpi load_row_epi(const GemmParams& p, int row) {
  RowEpi r;
  r.scale = p.row_scale ? __ldg(p.row_scale + row) : 1.0f;
  r.coef = 0.f; r.src = nullptr;
  if (p.row_coef) {
    const float c = __ldg(p.row_coef + row);
    const long long t = p.targets[row] - p.v_offset;
    if (c != 0.f && t >= 0 && t < p.gather_rows) { r.coef = c; r.src = p.gather_src + t * p.gather_ld; }
  }
  return r;
}
__device__ __forceinline__ void apply_row_epi(const RowEpi& r, uint32_t (&v)[32], int col0, int n_valid) {
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) * r.scale);
  if (r.src) {
    if (n_valid >= 32 && ((reinterpret_cast<uintptr_t>(r.src + col0) & 15) == 0)) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 w = __ldg(reinterpret_cast<const uint4*>(r.src + col0) + q);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&w);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 f = __bfloat1622float2(h[k]);
          v[8 * q + 2 * k] = __float_as_uint(fmaf(r.coef, f.x, __uint_as_float(v[8 * q + 2 * k])));
          v[8 * q + 2 * k + 1] = __float_as_uint(fmaf(r.coef, f.y, __uint_as_float(v[8 * q + 2 * k + 1])));
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < n_valid) v[j] = __float_as_uint(fmaf(r.coef, __bfloat162float(r.src[col0 + j]), __uint_as_float(v[j])));
    }
  }
}

// slot >= 0: tail piece, written whole (32 columns) at tile-local coordinates (r_local, c_local)
__device__ __forceinline__ void store_chunk(const GemmParams& p, uint32_t (&v)[32], int row, int col0, int n_valid, int ks,
                                            int slot, int r_local, int c_local, const RowEpi& re) {
  if (slot < 0 && p.k_splits <= 1 && (p.row_scale || p.row_coef)) apply_row_epi(re, v, col0, n_valid);   // pieces: applied by the reduce kernel
  if (slot >= 0) {
    float4* o4 = reinterpret_cast<float4*>(p.slab + ((long long)slot * p.tile_rows + r_local) * PANEL_W + c_local);
#pragma unroll
    for (int q = 0; q < 8; ++q)
      o4[q] = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
  } else if (p.k_splits > 1) {  // f32 partial of this k-split; reduced (and cast / accumulated) by splitk_reduce_kernel
    float* o = p.slab + ((long long)ks * p.M + row) * p.N + col0;
    if (n_valid >= 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        o4[q] = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3]));
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) if (j < n_valid) o[j] = __uint_as_float(v[j]);
    }
  } else if (p.c_is_f32) {
    float* o = reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col0;
    if (n_valid >= 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 x = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]),
                                     __uint_as_float(v[4 * q + 3]));
        // accumulate: the add happens in L2 (every element is touched once per launch, so the result does not
        // depend on ordering); the SM never waits for the old gradient to come back
        if (p.accumulate) ptx::red_add_v4_f32(o4 + q, x); else o4[q] = x;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < n_valid) o[j] = (p.accumulate ? o[j] : 0.f) + __uint_as_float(v[j]);
    }
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col0;
    if (n_valid >= 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      uint4* o4 = reinterpret_cast<uint4*>(o);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t w[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const __nv_bfloat162 b = __floats2bfloat162_rn(__uint_as_float(v[8 * q + 2 * h]), __uint_as_float(v[8 * q + 2 * h + 1]));
          w[h] = *reinterpret_cast<const uint32_t*>(&b);
        }
        o4[q] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < n_valid) o[j] = __float2bfloat16_rn(__uint_as_float(v[j]));
    }
  }
}

// Soft K-lockstep of the producers of one wave.  Every cluster announces each epoch (sync_kb k-blocks) of its tile on a counter
// shared by the wave, and does not start loading epoch e before all clusters of the wave have STARTED epoch e - sync_slack:
// the clusters then walk K within sync_slack + 1 epochs of one another, so an operand row fetched from DRAM for one tile is
// still in L2 when the other tiles of its row / column block ask for it.  Pacing only: the wait gives up after LOCKSTEP_WAIT_NS
// (a cluster that is not resident yet, a concurrent kernel holding SMs) -- results never depend on it and nothing can hang.
constexpr unsigned long long LOCKSTEP_WAIT_NS = 20000ull;
__device__ __forceinline__ void lockstep_pace(const GemmParams& p, int wave, int epoch, int n_clusters) {
  unsigned* c = p.sync + (long long)wave * p.sync_epochs + epoch;
  atomicAdd(c, 1u);
  if (epoch < p.sync_slack) return;
  const int rem = p.sync_full - wave * n_clusters;
  const unsigned need = (unsigned)(rem < n_clusters ? rem : n_clusters);
  const volatile unsigned* w = c - p.sync_slack;
  if (*w >= need) return;
  const unsigned long long t0 = ptx::globaltimer_ns();
  while (*w < need && ptx::globaltimer_ns() - t0 < LOCKSTEP_WAIT_NS) __nanosleep(64);
}

// ---------------------------------------------------------------------------------------
// The kernel.  kMRep = 1: 128 rows per CTA, two accumulator stages (epilogue overlaps the next
// tile's MMAs).  kMRep = 2: 256 rows per CTA = two 128-row MMAs per k-step that share one B tile
// in shared memory (25 % fewer operand bytes per FLOP from L2), one accumulator stage (all 512
// TMEM columns); used where the K loop is long enough to make the un-overlapped epilogue cheap.
// ---------------------------------------------------------------------------------------
template <int kEpi, int kAMajor, int kBMajor, int kCtas, int BLOCK_N, int kStages, bool kSave = false, bool kAPanel = false,
          int kMRep = 1>
__global__ void __launch_bounds__(GEMM_THREADS + 128 * (kMRep - 1), 1)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, const GemmParams p) {
  using L = GemmSmem<BLOCK_N, kCtas, kStages, kMRep>;
  constexpr int LOAD_N = L::LOAD_N;
  constexpr int CTA_M = GEMM_BLOCK_M * kMRep;                // rows of the output tile owned by one CTA
  constexpr int UMMA_M = GEMM_BLOCK_M * kCtas;
  constexpr int ACC_COLS = BLOCK_N * kMRep;                  // TMEM columns of one accumulator stage
  constexpr int ACC_STAGES = 512 / ACC_COLS >= 2 ? 2 : 1;
  constexpr uint32_t TMEM_COLS = ACC_STAGES * ACC_COLS <= 32 ? 32 : ACC_STAGES * ACC_COLS <= 64 ? 64 : ACC_STAGES * ACC_COLS <= 128 ? 128
                                 : ACC_STAGES * ACC_COLS <= 256 ? 256 : 512;
  static_assert(ACC_STAGES * ACC_COLS <= 512, "accumulators must fit the 512 TMEM columns");
  static_assert(BLOCK_N % 32 == 0 && LOAD_N % 64 == 0, "tile shape");
  static_assert(kMRep == 1 || kMRep == 2, "one or two 128-row MMAs per CTA");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t pad = (1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u;
  uint8_t* smem = smem_raw + pad;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = kCtas > 1 ? ptx::cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
  const int cluster_id = blockIdx.x / kCtas, n_clusters = gridDim.x / kCtas;
  const int n_items = p.n_items;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tm_a);
    ptx::prefetch_tensormap(&tm_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) { ptx::mbar_init(&full_bar[i], 2 * kCtas); ptx::mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(&tmem_full[i], 1); ptx::mbar_init(&tmem_empty[i], 128 * kCtas * kMRep); }
    ptx::fence_mbar_init();
  }
  if (kCtas > 1) ptx::cluster_sync();   // peer barriers initialised before anyone signals them
  if (warp == 2) ptx::tmem_alloc<kCtas>(tmem_ptr, TMEM_COLS);
  ptx::tc_fence_before();
  if (kCtas > 1) ptx::cluster_sync(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if ((warp == 0 || warp == 3) && lane == 0) {
    // ===================== TMA producers =====================
    // Two of them: warp 0 copies operand A, warp 3 operand B, each with its own walk over the items and the ring.  One
    // thread issuing every copy of a k-block executes ~110 instructions per k-block for dW's four copies (coordinates,
    // an ELECT / R2UR broadcast sequence per TMA operand, barrier traffic) -- about the 512 cycles the k-block's MMAs
    // take, so the producer, not the tensor pipe, set the pace there.  Split in two, each stays well ahead.
    auto produce = [&](auto a_side) {
      constexpr bool kA = decltype(a_side)::value;
      uint32_t stage = 0, phase = 0, tile_iter = 0;
      for (int item = cluster_id; item < n_items; item += n_clusters, ++tile_iter) {
        const Item it = decode_item(p, item);
        const int m0 = (it.m_blk * kCtas + (int)cta_rank) * CTA_M;
        const int n0 = it.nb * BLOCK_N + (int)cta_rank * LOAD_N;
        const bool rev = p.zigzag && (tile_iter & 1);
        const bool pace = kA && p.sync != nullptr && leader && it.slot < 0;
        int pace_left = 0, pace_epoch = 0;          // k-blocks until the next epoch mark (no division in the loop)
        int kb = rev ? it.kb1 - 1 : it.kb0;
        const int kb_step = rev ? -1 : 1;
        for (int kbi = it.kb0; kbi < it.kb1; ++kbi, kb += kb_step) {
          if (pace) {
            if (pace_left == 0) {
              lockstep_pace(p, (int)tile_iter, pace_epoch++, n_clusters);
              pace_left = p.sync_kb;
            }
            --pace_left;
          }
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1, 1);
          uint8_t* sa = smem + stage * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], (kA ? L::A_BYTES : L::B_BYTES) * kCtas);
          else ptx::mbar_arrive_cluster(&full_bar[stage], 0);
          const int k0 = kb * GEMM_BLOCK_K;
          auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1) {
            const uint64_t pol = tm == &tm_a ? p.hint_a : p.hint_b;
            if constexpr (kCtas == 1) ptx::tma_load_2d_hint(dst, tm, &full_bar[stage], c0, c1, pol);
            else ptx::tma_load_2d_pair_hint(dst, tm, &full_bar[stage], c0, c1, pol);
          };
          auto load3 = [&](void* dst, const CUtensorMap* tm, int c0, int c1, int c2) {
            if constexpr (kCtas == 1) ptx::tma_load_3d_hint(dst, tm, &full_bar[stage], c0, c1, c2, p.hint_a);
            else ptx::tma_load_3d_pair_hint(dst, tm, &full_bar[stage], c0, c1, c2, p.hint_a);
          };
          if constexpr (kA) {
            if constexpr (kAPanel) {  // A lives in 256-column panels: element (r, c) at panel c/256, row r, column c%256
              if constexpr (kAMajor == MAJOR_K) {
                load3(sa, &tm_a, k0 % PANEL_W, m0, k0 / PANEL_W);        // box {64 k, CTA_M rows, 1}
              } else {
#pragma unroll
                for (int a = 0; a < CTA_M / 64; ++a)
                  load3(sa + a * 8192, &tm_a, (m0 + a * 64) % PANEL_W, k0, (m0 + a * 64) / PANEL_W);  // box {64 m, 64 k, 1}
              }
            } else if constexpr (kAMajor == MAJOR_K) {
              load(sa, &tm_a, k0, m0);                                   // box {64 k, CTA_M rows}
            } else {
#pragma unroll
              for (int a = 0; a < CTA_M / 64; ++a) load(sa + a * 8192, &tm_a, m0 + a * 64, k0);  // box {64 m, 64 k}
            }
          } else {
            if constexpr (kBMajor == MAJOR_K) {
              load(sb, &tm_b, k0, n0);                                   // box {64 k, LOAD_N rows}
            } else {
#pragma unroll
              for (int a = 0; a < LOAD_N / 64; ++a) load(sb + a * 8192, &tm_b, n0 + a * 64, k0);
            }
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    };
    if (warp == 0) produce(std::true_type{}); else produce(std::false_type{});
  } else if (warp == 1 && lane == 0 && leader) {
    // ===================== MMA issuer (leader CTA, one thread) =====================
    constexpr uint32_t idesc = ptx::idesc_bf16_f32(UMMA_M, BLOCK_N, kAMajor, kBMajor);
    constexpr uint32_t A_LBO = kAMajor == MAJOR_K ? 0 : GEMM_BLOCK_K * 128, B_LBO = kBMajor == MAJOR_K ? 0 : GEMM_BLOCK_K * 128;
    constexpr uint32_t A_KSTEP = (kAMajor == MAJOR_K ? GEMM_UMMA_K * 2 : GEMM_UMMA_K * 128) >> 4;
    constexpr uint32_t B_KSTEP = (kBMajor == MAJOR_K ? GEMM_UMMA_K * 2 : GEMM_UMMA_K * 128) >> 4;
    constexpr uint32_t A_REP = (GEMM_BLOCK_M * GEMM_BLOCK_K * 2) >> 4;   // the second 128 rows start 16 KB further (both layouts)
    uint32_t stage = 0, phase = 0, tile_iter = 0;
    for (int item = cluster_id; item < n_items; item += n_clusters, ++tile_iter) {
      const Item it = decode_item(p, item);
      const uint32_t as = tile_iter % ACC_STAGES, aphase = (tile_iter / ACC_STAGES) & 1;
      ptx::mbar_wait(&tmem_empty[as], aphase ^ 1, 2);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * ACC_COLS;
      for (int kb = it.kb0; kb < it.kb1; ++kb) {
        ptx::mbar_wait(&full_bar[stage], phase, 3);
        ptx::tc_fence_after();
        const uint32_t sa = ptx::smem_u32(smem + stage * L::STAGE_BYTES), sb = sa + L::A_BYTES;
        const uint64_t da = ptx::smem_desc_sw128(sa, A_LBO, 1024), db = ptx::smem_desc_sw128(sb, B_LBO, 1024);
#pragma unroll
        for (int k = 0; k < GEMM_BLOCK_K / GEMM_UMMA_K; ++k) {
#pragma unroll
          for (int rep = 0; rep < kMRep; ++rep)
            ptx::umma_bf16<kCtas>(d_tmem + rep * BLOCK_N, da + (uint64_t)(k * A_KSTEP + rep * A_REP), db + (uint64_t)(k * B_KSTEP), idesc,
                                  (kb > it.kb0) || (k != 0));
        }
        ptx::umma_commit<kCtas>(&empty_bar[stage]);              // frees the smem slot when the MMAs retire
        if (kb == it.kb1 - 1) ptx::umma_commit<kCtas>(&tmem_full[as]);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
    if (kCtas > 1 && tile_iter > 0) {  // drain: peer epilogue arrivals must land before teardown
      const uint32_t last = tile_iter - 1;
      ptx::mbar_wait(&tmem_empty[last % ACC_STAGES], (last / ACC_STAGES) & 1, 4);
      if (ACC_STAGES > 1 && tile_iter > 1) ptx::mbar_wait(&tmem_empty[(last - 1) % ACC_STAGES], ((last - 1) / ACC_STAGES) & 1, 5);
    }
  } else if (warp >= GEMM_EPI_WARP0) {
    // ===================== epilogue (4 warps per 128-row accumulator, thread == TMEM lane == tile row) ============
    // kMRep == 2: warps 4-7 drain the first 128 rows (TMEM columns [0, BLOCK_N)), warps 8-11 the second 128 rows
    // concurrently -- the epilogue is not overlapped with the next tile's MMAs there, so its latency is what is exposed.
    const int quad = warp & 3;
    const int rep = (warp - GEMM_EPI_WARP0) >> 2;
    const int row_in_tile = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const float scale2 = p.inv_temp * LOG2E_F;
    uint32_t tile_iter = 0;
    for (int item = cluster_id; item < n_items; item += n_clusters, ++tile_iter) {
      const Item it = decode_item(p, item);
      const int n0 = it.nb * BLOCK_N;
      const uint32_t as = tile_iter % ACC_STAGES, aphase = (tile_iter / ACC_STAGES) & 1;
      {
        const int row = (it.m_blk * kCtas + (int)cta_rank) * CTA_M + rep * GEMM_BLOCK_M + row_in_tile;
        const bool row_ok = row < p.M;
        FwdRowState fs;
        BwdRowConst bc;
        RowEpi re = {1.0f, 0.f, nullptr};
        if constexpr (kEpi == EPI_FWD_STATS) {
          fs.neg2 = row_ok ? -p.row_ref[row] * LOG2E_F : 0.f; fs.s = 0.f; fs.u = 0.f; fs.tgt = 0.f;
          const long long t = row_ok ? p.targets[row] - p.v_offset : -1;
          fs.tcol = (t >= n0 && t < min(p.N, n0 + BLOCK_N)) ? (int)t : -1000000;
        }
        if constexpr (kEpi == EPI_BWD_DLOGITS) {
          const float gs = (p.grad_scale ? *p.grad_scale : 1.0f) * p.inv_temp;
          bc.lse2 = row_ok ? p.lse[row] * LOG2E_F : 0.f;
          bc.a = row_ok ? p.coef_lp[row] * gs : 0.f;
          bc.e = row_ok ? p.coef_ent[row] * gs : 0.f;
          bc.ent = row_ok ? p.entropy[row] : 0.f;
          const long long t = row_ok ? p.targets[row] - p.v_offset : -1;
          bc.tcol = (t >= 0 && t < p.N) ? (int)t : -1000000;
        }
        if constexpr (kEpi == EPI_STORE) {
          if (row_ok && (p.row_scale || p.row_coef)) re = load_row_epi(p, row);
        }
        ptx::mbar_wait(&tmem_full[as], aphase, 6);
        ptx::tc_fence_after();
        const uint32_t t_addr = tmem_base + lane_addr + as * ACC_COLS + rep * BLOCK_N;
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 32; ++c) {
          const int col0 = n0 + c * 32;
          if (col0 >= p.N) break;                      // warp-uniform
          uint32_t v[32];
          __syncwarp();
          ptx::tmem_ld_32x32(t_addr + c * 32, v);
          ptx::tmem_ld_wait();
          const int n_valid = min(32, p.N - col0);
          if constexpr (kEpi == EPI_FWD_STATS) {
            if (!row_ok) continue;
            fwd_chunk<kSave>(fs, v, col0, n_valid, scale2, kSave ? p.zsave + ((long long)it.nb * p.M + row) * PANEL_W + c * 32 : nullptr);
          }
          if constexpr (kEpi == EPI_BWD_DLOGITS) {
            if (row_ok) bwd_chunk(bc, v, col0, n_valid, scale2, p.dlogits + ((long long)it.nb * p.M + row) * PANEL_W + c * 32);
          }
          if constexpr (kEpi == EPI_STORE) {
            if (row_ok) store_chunk(p, v, row, col0, n_valid, it.ks, it.slot, (int)cta_rank * CTA_M + rep * GEMM_BLOCK_M + row_in_tile, c * 32, re);
          }
        }
        if constexpr (kEpi == EPI_FWD_STATS) {
          if (row_ok) {
            const float4 o = make_float4(0.f, fs.s, fs.u * p.inv_temp, fs.tgt * p.inv_temp);
            reinterpret_cast<float4*>(p.part)[(long long)it.nb * p.M + row] = o;
          }
        }
      }
      ptx::tc_fence_before();
      if (kCtas == 1) ptx::mbar_arrive(&tmem_empty[as]); else ptx::mbar_arrive_cluster(&tmem_empty[as], 0);
    }
  }

  // ===================== teardown =====================
  __syncwarp();
  ptx::tc_fence_before();
  if (kCtas > 1) ptx::cluster_sync(); else __syncthreads();
  if (warp == 2) { ptx::tc_fence_after(); ptx::tmem_dealloc<kCtas>(tmem_base, TMEM_COLS); }
}

}  // namespace b200grpo
