// K4: group-normalised advantages, replacing the pandas groupby/merge/apply path of
// /root/reference/pipelinerl/finetune/rl/__init__.py:394-499.
//   stats kernel : one thread per group walks ITS rows (a stable sort of the rows by group key, handed in by the
//                  host as row_order / group_starts) in dataset order -- O(rollouts) reads in total -- and folds them
//                  exactly the way pandas' groupby kernels do (libgroupby group_mean / group_var, pandas 3.0.2): the
//                  mean of the rewards and of the row lengths by Kahan-compensated summation, the ddof=1 variance by
//                  Welford's update, every operation an explicitly rounded f64 intrinsic (no FMA contraction), so
//                  mean / std / advantages are BIT-IDENTICAL to the reference's (:408-416, :466)
//   rows kernel  : one thread per rollout -> f64 advantage and f64 mean group length (what the list-of-dict contract
//                  carries: Python floats)
//   expand kernel: one warp per rollout row writes its per-token f32 columns with 16-byte stores
//                  (:461-469,:481-487); 16 B/token written (+ 8 B/token when the f64 advantages are wanted too).
#include "common.cuh"
#include <math.h>

namespace b200grpo {

struct GroupStat { double mean, std, gtok; };

// pandas group_mean: Kahan summation in row order, then sum / count
struct KahanSum {
  double sum = 0.0, comp = 0.0;
  __device__ __forceinline__ void add(double v) {
    const double y = __dsub_rn(v, comp);
    const double t = __dadd_rn(sum, y);
    comp = __dsub_rn(__dsub_rn(t, sum), y);
    if (comp != comp) comp = 0.0;   // pandas resets a NaN compensation (inf inputs)
    sum = t;
  }
};

__global__ void __launch_bounds__(128) group_stats_kernel(const double* __restrict__ row_reward,
                                                          const int32_t* __restrict__ row_order,
                                                          const int32_t* __restrict__ group_starts,
                                                          const int64_t* __restrict__ row_offsets, int n_groups,
                                                          GroupStat* __restrict__ out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_groups) return;
  const int b = group_starts[g], e = group_starts[g + 1];
  KahanSum rew, tok;
  double mean = 0.0, m2 = 0.0;   // Welford (pandas group_var): mean += (v - mean)/n; m2 += (v - mean_new) * (v - mean_old)
  for (int k = b; k < e; ++k) {
    const int j = __ldg(row_order + k);
    const double v = __ldg(row_reward + j);
    rew.add(v);
    tok.add((double)(__ldg(row_offsets + j + 1) - __ldg(row_offsets + j)));
    const double old = mean;
    mean = __dadd_rn(mean, __ddiv_rn(__dsub_rn(v, old), (double)(k - b + 1)));
    m2 = __dadd_rn(m2, __dmul_rn(__dsub_rn(v, mean), __dsub_rn(v, old)));
  }
  const int cnt = e - b;
  GroupStat s;
  s.mean = cnt ? __ddiv_rn(rew.sum, (double)cnt) : 0.0;
  s.gtok = cnt ? __ddiv_rn(tok.sum, (double)cnt) : 0.0;
  s.std = cnt > 1 ? __dsqrt_rn(__ddiv_rn(m2, (double)(cnt - 1))) : nan("");  // pandas std of one sample is NaN
  out[g] = s;
}

// (reward - mean) / (np.nan_to_num(std) + 1e-4)   (:466), each operation rounded once in f64 like the Python expression
__device__ __forceinline__ double advantage_f64(double r, const GroupStat& s, int divide_by_std) {
  const double d = __dsub_rn(r, s.mean);
  if (!divide_by_std) return d;
  return __ddiv_rn(d, __dadd_rn(isnan(s.std) ? 0.0 : s.std, 1e-4));
}

// One warp per rollout row; 16-byte stores when the four output columns are 16-byte aligned (the Python side
// pads the column pitch), scalar head / tail around them.
__global__ void __launch_bounds__(256) group_expand_kernel(const double* __restrict__ row_reward, const double* __restrict__ token_rewards,
                                                           const int32_t* __restrict__ group_key, const int64_t* __restrict__ row_offsets,
                                                           const int32_t* __restrict__ n_labels, const uint8_t* __restrict__ has_eos,
                                                           const GroupStat* __restrict__ stats, int n_rows, int divide_by_std,
                                                           float* __restrict__ adv, float* __restrict__ gtok, float* __restrict__ nlab,
                                                           float* __restrict__ ovf, double* __restrict__ adv64, int vec_ok) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < n_rows; row += gridDim.x * warps_per_block) {
    const GroupStat s = stats[group_key[row]];
    const double a0d = advantage_f64(row_reward[row], s, divide_by_std);
    const float a0 = (float)a0d;
    const float g = (float)s.gtok, nl = (float)n_labels[row], ov = has_eos[row] ? 0.f : 1.f;
    const int64_t b = row_offsets[row], e = row_offsets[row + 1];
    auto scalar = [&](int64_t t) {
      double ad = a0d;
      if (token_rewards) ad = advantage_f64(token_rewards[t], s, divide_by_std);
      adv[t] = (float)ad; gtok[t] = g; nlab[t] = nl; ovf[t] = ov;
      if (adv64) adv64[t] = ad;
    };
    if (!vec_ok || token_rewards || adv64) {
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

// Per-ROLLOUT form: the f64 advantage and f64 mean group length of every row -- the values the reference's lists hold as
// Python floats (the host repeats them per token with list multiplication; num_labels / overflow are host-side facts).
__global__ void __launch_bounds__(256) group_rows_kernel(const double* __restrict__ row_reward, const int32_t* __restrict__ group_key,
                                                         const GroupStat* __restrict__ stats, int n_rows, int divide_by_std,
                                                         double* __restrict__ adv, double* __restrict__ gtok) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= n_rows) return;
  const GroupStat s = stats[group_key[row]];
  adv[row] = advantage_f64(row_reward[row], s, divide_by_std);
  gtok[row] = s.gtok;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" size_t b200grpo_group_advantage_workspace(int n_rows, int n_groups) {
  (void)n_rows;
  return align_up((size_t)(n_groups > 0 ? n_groups : 1) * sizeof(GroupStat), 256);
}

extern "C" int b200grpo_group_advantage(const double* row_reward, const int32_t* group_key, const int32_t* row_order,
                                        const int32_t* group_starts, const int64_t* row_offsets, const int32_t* n_labels,
                                        const uint8_t* has_eos, const double* token_rewards, int n_rows, int n_groups,
                                        int divide_by_std, float* advantages, float* group_tokens, float* num_labels,
                                        float* overflow, double* advantages_f64, void* workspace, size_t workspace_bytes,
                                        void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(n_rows >= 0 && n_groups >= 0, "group_advantage: negative size");
  if (n_rows == 0) return B200GRPO_OK;
  B200_REQUIRE(row_reward && group_key && row_order && group_starts && row_offsets && n_labels && has_eos && advantages &&
               group_tokens && num_labels && overflow && workspace, "group_advantage: null pointer");
  if (workspace_bytes < b200grpo_group_advantage_workspace(n_rows, n_groups)) {
    set_error("group_advantage: workspace too small");
    return B200GRPO_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  GroupStat* stats = (GroupStat*)workspace;
  ProfScope ps(PROF_ADVANTAGE, st);
  if (n_groups > 0) {
    group_stats_kernel<<<ceil_div(n_groups, 128), 128, 0, st>>>(row_reward, row_order, group_starts, row_offsets, n_groups, stats);
    count_launch();
  }
  const int vec_ok = (((uintptr_t)advantages | (uintptr_t)group_tokens | (uintptr_t)num_labels | (uintptr_t)overflow) & 15) == 0;
  int grid = ceil_div(n_rows, 8);
  const int cap = sm_count() * 8;
  if (grid > cap) grid = cap;
  group_expand_kernel<<<grid, 256, 0, st>>>(row_reward, token_rewards, group_key, row_offsets, n_labels, has_eos, stats,
                                            n_rows, divide_by_std, advantages, group_tokens, num_labels, overflow, advantages_f64,
                                            vec_ok); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_group_advantage_rows(const double* row_reward, const int32_t* group_key, const int32_t* row_order,
                                             const int32_t* group_starts, const int64_t* row_offsets, int n_rows, int n_groups,
                                             int divide_by_std, double* advantages, double* group_tokens, void* workspace,
                                             size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(n_rows >= 0 && n_groups >= 0, "group_advantage_rows: negative size");
  if (n_rows == 0) return B200GRPO_OK;
  B200_REQUIRE(row_reward && group_key && row_order && group_starts && row_offsets && advantages && group_tokens && workspace,
               "group_advantage_rows: null pointer");
  if (workspace_bytes < b200grpo_group_advantage_workspace(n_rows, n_groups)) {
    set_error("group_advantage_rows: workspace too small");
    return B200GRPO_EWORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  GroupStat* stats = (GroupStat*)workspace;
  ProfScope ps(PROF_ADVANTAGE, st);
  if (n_groups > 0) {
    group_stats_kernel<<<ceil_div(n_groups, 128), 128, 0, st>>>(row_reward, row_order, group_starts, row_offsets, n_groups, stats);
    count_launch();
  }
  group_rows_kernel<<<ceil_div(n_rows, 256), 256, 0, st>>>(row_reward, group_key, stats, n_rows, divide_by_std, advantages,
                                                           group_tokens); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
