// K4: group-normalised advantages, replacing the pandas groupby/merge/apply path of
// /root/reference/pipelinerl/finetune/rl/__init__.py:394-499.
//   stats kernel : per group key, mean reward / sample std (ddof=1) / mean row length in
//                  f64, rows visited in dataset order (:408-416)
//   expand kernel: one warp per rollout row writes its per-token f32 columns with
//                  coalesced stores (:461-469,:481-487); 16 B/token written.
#include "common.cuh"
#include <math.h>

namespace b200grpo {

struct GroupStat { double mean, std, gtok; };

__global__ void __launch_bounds__(128) group_stats_kernel(const double* row_reward, const int32_t* group_key,
                                                          const int64_t* row_offsets, int n_rows, int n_groups,
                                                          GroupStat* out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_groups) return;
  double sum = 0.0, tok = 0.0;
  int cnt = 0;
  for (int i = 0; i < n_rows; ++i)
    if (group_key[i] == g) { sum += row_reward[i]; tok += (double)(row_offsets[i + 1] - row_offsets[i]); ++cnt; }
  GroupStat s;
  s.mean = cnt ? sum / cnt : 0.0;
  s.gtok = cnt ? tok / cnt : 0.0;
  double ss = 0.0;
  for (int i = 0; i < n_rows; ++i)
    if (group_key[i] == g) { const double d = row_reward[i] - s.mean; ss += d * d; }
  s.std = cnt > 1 ? sqrt(ss / (cnt - 1)) : nan("");  // pandas std of one sample is NaN
  out[g] = s;
}

__global__ void __launch_bounds__(256) group_expand_kernel(const double* row_reward, const double* token_rewards,
                                                           const int32_t* group_key, const int64_t* row_offsets,
                                                           const int32_t* n_labels, const uint8_t* has_eos,
                                                           const GroupStat* stats, int n_rows, int divide_by_std,
                                                           float* adv, float* gtok, float* nlab, float* ovf) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < n_rows; row += gridDim.x * warps_per_block) {
    const GroupStat s = stats[group_key[row]];
    const double denom = (isnan(s.std) ? 0.0 : s.std) + 1e-4;  // np.nan_to_num(std) + 1e-4 (:466)
    const double r0 = row_reward[row];
    const float a0 = (float)(divide_by_std ? (r0 - s.mean) / denom : (r0 - s.mean));
    const float g = (float)s.gtok, nl = (float)n_labels[row], ov = has_eos[row] ? 0.f : 1.f;
    const int64_t b = row_offsets[row], e = row_offsets[row + 1];
    for (int64_t t = b + lane; t < e; t += 32) {
      float a = a0;
      if (token_rewards) {
        const double r = token_rewards[t];
        a = (float)(divide_by_std ? (r - s.mean) / denom : (r - s.mean));
      }
      adv[t] = a; gtok[t] = g; nlab[t] = nl; ovf[t] = ov;
    }
  }
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
  group_stats_kernel<<<ceil_div(n_groups, 128), 128, 0, st>>>(row_reward, group_key, row_offsets, n_rows, n_groups, stats);
  int grid = ceil_div(n_rows, 8);
  const int cap = sm_count() * 8;
  if (grid > cap) grid = cap;
  group_expand_kernel<<<grid, 256, 0, st>>>(row_reward, token_rewards, group_key, row_offsets, n_labels, has_eos, stats,
                                            n_rows, divide_by_std, advantages, group_tokens, num_labels, overflow);
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
