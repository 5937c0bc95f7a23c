// Packed-batch ingest: expand the wire form of a packed micro-batch -- 16 B per token
// (input id, label, old / ref log-prob) plus 5 floats per rollout -- into the eleven [1, T] tensors
// rl_step reads, on the device.  Replaces the host-side per-token materialisation of
// pipelinerl/finetune/data.py:215-279 (collate_packed) + the 11 tensor H2D copies of
// finetune_loop.py:109 (60 B per token) and the per-token advantage / group_tokens / num_labels /
// overflow lists of rl/__init__.py:481-499.
//
// One thread per token: binary search of the token's rollout in the (<= 64 K entry) boundary table,
// then eleven coalesced stores.  HBM-bound: 16 B read + 76 B written per token.
#include "common.cuh"

namespace b200grpo {

constexpr int ING_THREADS = 256;

__global__ void __launch_bounds__(ING_THREADS)
expand_packed_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ labels, const float* __restrict__ old_lp,
                     const float* __restrict__ ref_lp, const int32_t* __restrict__ bounds, const float* __restrict__ scalars,
                     int T, int n_seq, b200grpo_packed_out o) {
  const int t = blockIdx.x * ING_THREADS + threadIdx.x;
  if (t >= T) return;
  int lo = 0, hi = n_seq;                       // largest s with bounds[s] <= t; empty rollouts (bounds[s] == bounds[s+1]) are skipped
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(bounds + mid) <= t) lo = mid; else hi = mid;
  }
  const int s = lo;
  o.input_ids[t] = ids[t];
  o.labels[t] = labels[t];
  o.attention_mask[t] = 1;
  o.position_ids[t] = t - __ldg(bounds + s);
  o.old_logprobs[t] = old_lp[t];
  o.ref_logprobs[t] = ref_lp[t];
  o.rewards[t] = __ldg(scalars + 0 * n_seq + s);
  o.advantages[t] = __ldg(scalars + 1 * n_seq + s);
  o.group_tokens[t] = __ldg(scalars + 2 * n_seq + s);
  o.num_labels[t] = __ldg(scalars + 3 * n_seq + s);
  o.overflow[t] = __ldg(scalars + 4 * n_seq + s);
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" int b200grpo_expand_packed(const int32_t* input_ids, const int32_t* labels, const float* old_logprobs,
                                      const float* ref_logprobs, const int32_t* seq_boundaries, const float* seq_scalars, int T,
                                      int n_seq, const b200grpo_packed_out* out, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(input_ids && labels && old_logprobs && ref_logprobs && seq_boundaries && seq_scalars && out,
               "expand_packed: null pointer");
  B200_REQUIRE(T >= 0 && n_seq >= 1, "expand_packed: T=%d n_seq=%d", T, n_seq);
  B200_REQUIRE(out->input_ids && out->labels && out->attention_mask && out->position_ids && out->rewards && out->advantages &&
               out->ref_logprobs && out->old_logprobs && out->group_tokens && out->num_labels && out->overflow,
               "expand_packed: null output");
  if (T == 0) return B200GRPO_OK;
  ProfScope ps(PROF_INGEST, (cudaStream_t)stream);
  expand_packed_kernel<<<ceil_div(T, ING_THREADS), ING_THREADS, 0, (cudaStream_t)stream>>>(
      input_ids, labels, old_logprobs, ref_logprobs, seq_boundaries, seq_scalars, T, n_seq, *out); count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
