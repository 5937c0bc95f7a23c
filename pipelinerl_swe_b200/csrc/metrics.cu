// Step-metrics accumulator: folds one micro-batch's K3 metrics vector into a fixed-layout device
// vector, replacing the per-micro-batch python dict appends of pipelinerl/finetune_loop.py:669-670
// and making the pickled all_gather_object of :62-87 one fixed-size collective.  No host sync.
//
// acc layout (f64): [0,K) value per stat key, [K,2K) number of finite values folded into it,
// 2K flags (bit-or, as a double), 2K+1 micro-batches with labels, 2K+2 micro-batches seen,
// 2K+3 sequences seen (the local sample counter of :606).
#include "common.cuh"

namespace b200grpo {

__global__ void metrics_accumulate_kernel(const double* __restrict__ m, const int* __restrict__ slots, const int* __restrict__ kinds,
                                          int K, double kl_coef, double ent_coef, double input_size, double n_samples,
                                          int record, double* __restrict__ acc) {
  const int k = threadIdx.x;
  const bool empty = m[B200GRPO_M_N_MASKED] == 0.0;  // rl_step returned {"input_size": ...} only (rl/__init__.py:195)
  if (k < K && record) {
    const int s = slots[k];
    double v;
    bool present = !empty;
    if (s >= 0) v = m[s];
    else if (s == B200GRPO_STAT_KL_COEF) v = m[B200GRPO_M_N_SEQ] * kl_coef;
    else if (s == B200GRPO_STAT_ENTROPY_COEF) v = m[B200GRPO_M_N_SEQ] * ent_coef;
    else { v = input_size; present = true; }
    if (present && isfinite(v)) {  // gather_rl_metrics keeps finite values only (finetune_loop.py:85)
      const int kind = kinds[k];
      const double a = acc[k];
      acc[k] = kind == B200GRPO_STAT_MIN ? fmin(a, v) : kind == B200GRPO_STAT_MAX ? fmax(a, v) : a + v;
      acc[K + k] += 1.0;
    }
  }
  if (k == 0) {
    if (record) {
      const unsigned long long f = (unsigned long long)acc[2 * K] | (unsigned long long)m[B200GRPO_M_FLAGS];
      acc[2 * K] = (double)f;
      if (!empty) acc[2 * K + 1] += 1.0;
    }
    acc[2 * K + 2] += 1.0;
    acc[2 * K + 3] += n_samples;
  }
}

__global__ void metrics_reset_kernel(const int* __restrict__ kinds, int K, double* __restrict__ acc) {
  const int k = threadIdx.x;
  if (k < K) {
    const int kind = kinds[k];
    acc[k] = kind == B200GRPO_STAT_MIN ? (double)INFINITY : kind == B200GRPO_STAT_MAX ? -(double)INFINITY : 0.0;
    acc[K + k] = 0.0;
  }
  if (k < 4) acc[2 * K + k] = 0.0;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" int b200grpo_metrics_reset(const int* kinds, int n_stats, double* acc, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(kinds && acc && n_stats >= 1 && n_stats <= B200GRPO_STATS_MAX, "metrics_reset: bad argument (n_stats=%d)", n_stats);
  metrics_reset_kernel<<<1, B200GRPO_STATS_MAX, 0, (cudaStream_t)stream>>>(kinds, n_stats, acc); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}

extern "C" int b200grpo_metrics_accumulate(const double* metrics, const int* slots, const int* kinds, int n_stats, double kl_coef,
                                           double entropy_coef, double input_size, double n_samples, int record, double* acc,
                                           void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(metrics && slots && kinds && acc && n_stats >= 1 && n_stats <= B200GRPO_STATS_MAX,
               "metrics_accumulate: bad argument (n_stats=%d)", n_stats);
  metrics_accumulate_kernel<<<1, B200GRPO_STATS_MAX, 0, (cudaStream_t)stream>>>(metrics, slots, kinds, n_stats, kl_coef, entropy_coef,
                                                                                input_size, n_samples, record, acc); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
