// K4: group-normalised advantages, replacing the pandas groupby/merge/apply path of
// /root/reference/pipelinerl/finetune/rl/__init__.py:394-499.
//   stats kernel : one warp per group key, mean reward / sample std (ddof=1) / mean row length in
//                  f64, rows folded in dataset order (:408-416)
//   expand kernel: one warp per rollout row writes its per-token f32 columns with
//                  16-byte stores (:461-469,:481-487); 16 B/token written.
#include "common.cuh"
#include <math.h>

namespace b200grpo {

struct GroupStat { double mean, std, gtok; };

// One warp per group key, eight keys per block.  The block stages the key column through shared memory in
// chunks of GS_CHUNK rows (coalesced, every thread keeps GS_CHUNK/256 loads in flight); each warp then ballots
// its key against the chunk 32 rows at a time and folds the matches IN DATASET ORDER (lowest lane first), so
// the f64 sums are the sequential sums a host loop over the rows would produce, whatever the group sizes.
// Two sweeps: mean, then squared deviations.
constexpr int GS_WARPS = 8;
constexpr int GS_CHUNK = 2048;
__global__ void __launch_bounds__(GS_WARPS * 32) group_stats_kernel(const double* __restrict__ row_reward,
                                                                    const int32_t* __restrict__ group_key,
                                                                    const int64_t* __restrict__ row_offsets, int n_rows,
                                                                    int n_groups, GroupStat* __restrict__ out) {
  __shared__ int32_t keys[GS_CHUNK];
  const int g = blockIdx.x * GS_WARPS + (threadIdx.x >> 5);   // may be >= n_groups in the last block: matches nothing
  const int lane = threadIdx.x & 31;
  double sum = 0.0, tok = 0.0, mean = 0.0, ss = 0.0;
  int cnt = 0;
  for (int pass = 0; pass < 2; ++pass) {
    for (int c0 = 0; c0 < n_rows; c0 += GS_CHUNK) {
      const int len = min(GS_CHUNK, n_rows - c0);
      __syncthreads();
#pragma unroll
      for (int k = 0; k < GS_CHUNK / (GS_WARPS * 32); ++k) {
        const int i = k * GS_WARPS * 32 + threadIdx.x;
        keys[i] = i < len ? __ldg(group_key + c0 + i) : -1;
      }
      __syncthreads();
      for (int base = 0; base < len; base += 32) {
        unsigned m = __ballot_sync(0xffffffffu, keys[base + lane] == g);
        if (pass == 0) cnt += __popc(m);
        while (m) {                                 // warp-uniform loop: every lane keeps the same running sums
          const int j = c0 + base + __ffs(m) - 1;
          m &= m - 1;
          if (pass == 0) {
            sum += __ldg(row_reward + j);
            tok += (double)(__ldg(row_offsets + j + 1) - __ldg(row_offsets + j));
          } else {
            const double d = __ldg(row_reward + j) - mean;
            ss += d * d;
          }
        }
      }
    }
    if (pass == 0) mean = cnt ? sum / cnt : 0.0;
  }
  if (g < n_groups && lane == 0) {
    GroupStat s;
    s.mean = mean;
    s.gtok = cnt ? tok / cnt : 0.0;
    s.std = cnt > 1 ? sqrt(ss / (cnt - 1)) : nan("");  // pandas std of one sample is NaN
    out[g] = s;
  }
}

// One warp per rollout row; 16-byte stores when the four output columns are 16-byte aligned (the Python side
// pads the column pitch), scalar head / tail around them.
__global__ void __launch_bounds__(256) group_expand_kernel(const double* __restrict__ row_reward, const double* __restrict__ token_rewards,
                                                           const int32_t* __restrict__ group_key, const int64_t* __restrict__ row_offsets,
                                                           const int32_t* __restrict__ n_labels, const uint8_t* __restrict__ has_eos,
                                                           const GroupStat* __restrict__ stats, int n_rows, int divide_by_std,
                                                           float* __restrict__ adv, float* __restrict__ gtok, float* __restrict__ nlab,
                                                           float* __restrict__ ovf, int vec_ok) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < n_rows; row += gridDim.x * warps_per_block) {
    const GroupStat s = stats[group_key[row]];
    const double denom = (isnan(s.std) ? 0.0 : s.std) + 1e-4;  // np.nan_to_num(std) + 1e-4 (:466)
    const double r0 = row_reward[row];
    const float a0 = (float)(divide_by_std ? (r0 - s.mean) / denom : (r0 - s.mean));
    const float g = (float)s.gtok, nl = (float)n_labels[row], ov = has_eos[row] ? 0.f : 1.f;
    const int64_t b = row_offsets[row], e = row_offsets[row + 1];
    auto scalar = [&](int64_t t) {
      float a = a0;
      if (token_rewards) {
        const double r = token_rewards[t];
        a = (float)(divide_by_std ? (r - s.mean) / denom : (r - s.mean));
      }
      adv[t] = a; gtok[t] = g; nlab[t] = nl; ovf[t] = ov;
    };
    if (!vec_ok || token_rewards) {
      for (int64_t t = b + lane; t < e; t += 32) scalar(t);
      continue;
    }
    const int64_t vb = min(e, (b + 3) & ~(int64_t)3), ve = max(vb, e & ~(int64_t)3);   // [vb, ve) is whole quads
    for (int64_t t = b + lane; t < vb; t += 32) scalar(t);
    const float4 a4 = make_float4(a0, a0, a0, a0), g4 = make_float4(g, g, g, g), n4 = make_float4(nl, nl, nl, nl),
                 o4 = make_float4(ov, ov, ov, ov);
    for (int64_t t = vb + 4 * lane; t < ve; t += 128) {
      *reinterpret_cast<float4*>(adv + t) = a4;
      *reinterpret_cast<float4*>(gtok + t) = g4;
      *reinterpret_cast<float4*>(nlab + t) = n4;
      *reinterpret_cast<float4*>(ovf + t) = o4;
    }
    for (int64_t t = ve + lane; t < e; t += 32) scalar(t);
  }
}

// Per-ROLLOUT form of the expansion: one value per row and column (the list-of-dict contract of populate_rl_data
// only needs these; the host repeats them per token with list multiplication).
__global__ void __launch_bounds__(256) group_rows_kernel(const double* __restrict__ row_reward, const int32_t* __restrict__ group_key,
                                                         const int32_t* __restrict__ n_labels, const uint8_t* __restrict__ has_eos,
                                                         const GroupStat* __restrict__ stats, int n_rows, int divide_by_std,
                                                         float* __restrict__ adv, float* __restrict__ gtok, float* __restrict__ nlab,
                                                         float* __restrict__ ovf) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n_rows) return;
  const GroupStat s = stats[group_key[row]];
  const double denom = (isnan(s.std) ? 0.0 : s.std) + 1e-4;  // np.nan_to_num(std) + 1e-4 (:466)
  const double r0 = row_reward[row];
  adv[row] = (float)(divide_by_std ? (r0 - s.mean) / denom : (r0 - s.mean));
  gtok[row] = (float)s.gtok;
  nlab[row] = (float)n_labels[row];
  ovf[row] = has_eos[row] ? 0.f : 1.f;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" size_t b200grpo_group_advantage_workspace(int n_rows, int n_groups) {
  (void)n_rows;
  return align_up((size_t)(n_groups > 0 ? n_groups : 1) * sizeof(GroupStat), 256);
}

extern "C" int b200grpo_group_advantage(const double* row_reward, const int32_t* group_key, const int64_t* row_offsets,
                                        const int32_t* n_labels, const uint8_t* has_eos, const double* token_rewards,
                                        int n_rows, int n_groups, int divide_by_std, float* advantages,
                                        float* group_tokens, float* num_labels, float* overflow, void* workspace,
                                        size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(n_rows >= 0 && n_groups >= 0, "group_advantage: negative size");
  if (n_rows == 0) return B200GRPO_OK;
  B200_REQUIRE(row_reward && group_key && row_offsets && n_labels && has_eos && advantages && group_tokens &&
               num_labels && overflow && workspace, "group_advantage: null pointer");
  if (workspace_bytes < b200grpo_group_advantage_workspace(n_rows, n_groups)) {
    set_error("group_advantage: workspace too small");
    return B200GRPO_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  GroupStat* stats = (GroupStat*)workspace;
  ProfScope ps(PROF_ADVANTAGE, st);
  if (n_groups > 0) count_launch();
  if (n_groups > 0)
    group_stats_kernel<<<ceil_div(n_groups, GS_WARPS), GS_WARPS * 32, 0, st>>>(row_reward, group_key, row_offsets, n_rows, n_groups, stats);
  const int vec_ok = (((uintptr_t)advantages | (uintptr_t)group_tokens | (uintptr_t)num_labels | (uintptr_t)overflow) & 15) == 0;
  int grid = ceil_div(n_rows, 8);
  const int cap = sm_count() * 8;
  if (grid > cap) grid = cap;
  group_expand_kernel<<<grid, 256, 0, st>>>(row_reward, token_rewards, group_key, row_offsets, n_labels, has_eos, stats,
                                            n_rows, divide_by_std, advantages, group_tokens, num_labels, overflow, vec_ok); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_group_advantage_rows(const double* row_reward, const int32_t* group_key, const int64_t* row_offsets,
                                             const int32_t* n_labels, const uint8_t* has_eos, int n_rows, int n_groups,
                                             int divide_by_std, float* advantages, float* group_tokens, float* num_labels,
                                             float* overflow, void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(n_rows >= 0 && n_groups >= 0, "group_advantage_rows: negative size");
  if (n_rows == 0) return B200GRPO_OK;
  B200_REQUIRE(row_reward && group_key && row_offsets && n_labels && has_eos && advantages && group_tokens && num_labels &&
               overflow && workspace, "group_advantage_rows: null pointer");
  if (workspace_bytes < b200grpo_group_advantage_workspace(n_rows, n_groups)) {
    set_error("group_advantage_rows: workspace too small");
    return B200GRPO_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  GroupStat* stats = (GroupStat*)workspace;
  ProfScope ps(PROF_ADVANTAGE, st);
  if (n_groups > 0) {
    group_stats_kernel<<<ceil_div(n_groups, GS_WARPS), GS_WARPS * 32, 0, st>>>(row_reward, group_key, row_offsets, n_rows, n_groups, stats);
    count_launch();
  }
  group_rows_kernel<<<ceil_div(n_rows, 256), 256, 0, st>>>(row_reward, group_key, n_labels, has_eos, stats, n_rows, divide_by_std,
                                                           advantages, group_tokens, num_labels, overflow); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
