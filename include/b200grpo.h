/*
 * b200grpo.h -- C ABI of libb200grpo.so: the trainer-side GRPO hot path of
 * ServiceNow/PipelineRL-SWE re-built as hand-written sm_100a CUDA kernels.
 *
 * The reference has no FFI (it is pure Python over torch); the boundary it offers is the
 * Python call signatures rl_step / populate_rl_data.  Each entry point
 * below names the reference lines whose work it replaces, relative to
 * /root/reference/pipelinerl/.  The Python mirror of those signatures lives in
 * pipelinerl_swe_b200/ and reaches this library through ctypes only
 * (INTEGRATION.md shows the binding).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), allocation
 *     free (the caller passes workspace), re-entrant, and returns 0 on success or a
 *     negative B200GRPO_E* code; b200grpo_last_error() gives a thread-local message;
 *   - bf16 tensors are row-major with the given leading dimension in ELEMENTS;
 *   - there is no CPU fallback: without a compute-capability-10.x device calls fail.
 */
#ifndef B200GRPO_H_
#define B200GRPO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200GRPO_ABI_VERSION 3

enum {
  B200GRPO_OK = 0,
  B200GRPO_EINVAL = -1,    /* bad argument (shape, alignment, null pointer) */
  B200GRPO_ECUDA = -2,     /* CUDA runtime / driver error */
  B200GRPO_EWORKSPACE = -3,/* workspace too small */
  B200GRPO_EARCH = -4      /* device is not sm_100 */
};

const char* b200grpo_last_error(void);
int b200grpo_abi_version(void);

/* ---------------------------------------------------------------------------------------
 * K1  lm_head forward fused with online log-softmax / entropy / target gather.
 * Replaces finetune/rl/__init__.py:197-208 (logits = lm_head(hidden); /temperature;
 * log_softmax; softmax; entropy; gather) without materialising [tokens x vocab].
 *
 *   hidden  bf16 [M, H]   (ldh elements between rows)   rows = shifted token positions
 *   weight  bf16 [Vloc,H] (ldw)                         this rank's vocab rows
 *   targets i64  [M]      global vocab id scored at each row (input_ids[t+1])
 *   v_offset              global id of weight row 0 (0 when the vocab is not sharded)
 *   stats   f32  [M,4]    out: (r, sum exp(z-r), sum exp(z-r)*z, z[target] or 0) with z = logit /
 *                         temperature over this rank's vocab rows and r a per-row reference known
 *                         before the GEMM (max of the target logit and four sampled logits), which
 *                         makes the epilogue one pass; b200grpo_lmhead_finalize merges ranks
 *   saved   nullable, b200grpo_lmhead_saved_bytes(), 256-byte aligned.  When given, the
 *           epilogue also keeps E = exp(z - r) of this micro-batch as bf16 (f32's exponent
 *           range, 2^-9 relative precision at any magnitude; every quantity of the backward
 *           is a row-scaled multiple of it) in [Vloc/256][M][256] panels, plus r, for a
 *           recompute-free backward: the [M x Vloc] scratch lives only between this call
 *           and the matching b200grpo_lmhead_bwd.
 *           NULL: nothing of size [M x Vloc] is written by the forward.
 * ------------------------------------------------------------------------------------- */
size_t b200grpo_lmhead_fwd_workspace(int M, int H, int Vloc);
size_t b200grpo_lmhead_saved_bytes(int M, int H, int Vloc);
int b200grpo_lmhead_fwd(const void* hidden, int64_t ldh, const void* weight, int64_t ldw,
                        const int64_t* targets, int M, int H, int Vloc, int64_t v_offset,
                        float inv_temperature, float* stats, void* saved, void* workspace,
                        size_t workspace_bytes, void* stream);

/* Combine per-rank stats (all-gathered, [R, M, 4]) into per-token results.
 * logprob = z_target - lse ; entropy = lse - (sum e*z)/(sum e)   (rl/__init__.py:202-208).
 * R = 1 when the vocab is not sharded.  Outputs f32 [M]. */
int b200grpo_lmhead_finalize(const float* stats, int R, int M, float* logprob, float* entropy,
                             float* lse, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3  per-token RL loss + all reductions + gradient coefficients in one launch.
 * Replaces rl/__init__.py:152-181 (masks, segments), :213-292 (ratios, KL, surrogate,
 * weights, loss), :294-313 (value loss), :315-375 (stats) and rl/utils.py:25-30,66-87.
 * ------------------------------------------------------------------------------------- */
typedef struct b200grpo_loss_cfg {
  int32_t policy_loss;          /* 0 = "ppo", 1 = "reinforce"              (:269-283) */
  int32_t use_advantages;       /* log_p_weights = advantages else rewards (:250)     */
  int32_t relu_log_p_weights;   /* clamp(min=0)                            (:251-252) */
  int32_t group_normalization;  /* weight = 1/group_tokens else 1/batch    (:220-225) */
  int32_t overlong_filtering;   /* weight *= 1-overflow                    (:227-230) */
  int32_t is_packed;            /* segments from position_ids==0           (:158-178) */
  int32_t has_values;           /* value-head branch                       (:239-246) */
  int32_t batch_size;           /* config.batch_size                                  */
  float epsilon;                /* ratio clip                                         */
  float clamp_log_ratio_ref_new;/* C                                       (:254-260) */
  float kl_coef;                /* already linearly decayed                (:266)     */
  float entropy_coef;           /* already linearly decayed                (:265)     */
  float value_loss_coef;        /*                                         (:308)     */
  float reserved;
} b200grpo_loss_cfg;

/* slots of the f64 metrics vector written by b200grpo_token_loss */
enum {
  B200GRPO_M_LOSS = 0,        /* final loss (policy + value_loss_coef*value)          */
  B200GRPO_M_N_MASKED,        /* masks_shifted.sum()                                  */
  B200GRPO_M_N_SEQ,           /* number of segments (packed) / rows (unpacked)        */
  B200GRPO_M_NUM_NANS,        /* isnan(per-token loss).sum()                          */
  B200GRPO_M_REWARD, B200GRPO_M_ENTROPY, B200GRPO_M_OLD_LP, B200GRPO_M_NEW_LP,
  B200GRPO_M_REF_LP, B200GRPO_M_ADVANTAGE, B200GRPO_M_KL, B200GRPO_M_POLICY,
  B200GRPO_M_SURR1, B200GRPO_M_SURR2, B200GRPO_M_RATIO, B200GRPO_M_RATIO_SUM,
  B200GRPO_M_RATIO_SQ_SUM, B200GRPO_M_RATIO_REF_NEW, B200GRPO_M_RATIO_REF_OLD,
  B200GRPO_M_CLAMP_REF_NEW, B200GRPO_M_CLAMP_NEW_OLD, B200GRPO_M_TOKEN_WEIGHT,
  B200GRPO_M_VALUE_MEAN, B200GRPO_M_VALUE_LOSS, B200GRPO_M_VALUE_MSE,
  B200GRPO_M_MAX_REWARD, B200GRPO_M_MIN_REWARD, B200GRPO_M_MAX_ADV, B200GRPO_M_MIN_ADV,
  B200GRPO_M_MAX_KL, B200GRPO_M_MIN_KL, B200GRPO_M_MAX_TW, B200GRPO_M_MIN_TW,
  B200GRPO_M_MAX_VALUE, B200GRPO_M_MIN_VALUE,
  B200GRPO_M_FLAGS,           /* bit set of B200GRPO_F_* as a double                  */
  B200GRPO_M_COUNT = 40
};
enum {
  B200GRPO_F_NEW_LP_NONFINITE = 1,  /* assert :209 */
  B200GRPO_F_REF_NEW_NONFINITE = 2, /* assert :237 */
  B200GRPO_F_KL_NONFINITE = 4,      /* assert :264 */
  B200GRPO_F_LOSS_NONFINITE = 8,    /* assert :313 */
  B200GRPO_F_GROUP_TOKENS_LE0 = 16  /* assert :222 */
};

/* Inputs are the UN-shifted [rows, L] batch tensors (finetune/types.py:54-65); the
 * kernel applies the [:,1:] shift itself.  logprob/entropy are [rows, L-1].
 * values (nullable unless has_values) is [rows, L] f32.
 * Outputs: metrics f64[B200GRPO_M_COUNT]; loss f32[1] (nullable; = metrics[0]);
 * coef_lp/coef_ent f32[rows,L-1] = dLoss/dlogprob,
 * dLoss/dentropy for an upstream gradient of 1; coef_value f32[rows,L] (nullable). */
size_t b200grpo_token_loss_workspace(int rows, int L);
int b200grpo_token_loss(const float* logprob, const float* entropy, const int64_t* labels,
                        const int64_t* position_ids, const float* rewards,
                        const float* advantages, const float* ref_logprobs,
                        const float* old_logprobs, const float* group_tokens,
                        const float* num_labels, const float* overflow, const float* values,
                        int rows, int L, const b200grpo_loss_cfg* cfg_host, double* metrics,
                        float* loss, float* coef_lp, float* coef_ent, float* coef_value, void* workspace,
                        size_t workspace_bytes, void* stream);

/* Segment boundaries exactly as rl/__init__.py:159-170: starts where position_ids==0 plus
 * index 0, then T.  bounds i32[T+1] (capacity), n_bounds i32[1] = S+1.
 * position_ids is read once (8 B/token); workspace: b200grpo_segment_bounds_workspace(T) bytes. */
size_t b200grpo_segment_bounds_workspace(int T);
int b200grpo_segment_bounds(const int64_t* position_ids, int T, int32_t* bounds,
                            int32_t* n_bounds, void* workspace, size_t workspace_bytes,
                            void* stream);

/* Allocation-free variant of b200grpo_segment_bounds.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same kernels and the same outputs,
 * bit for bit; the scratch (per-block counts + the 1-bit start mask, ~520 KB) is a file-scope
 * __device__ array inside the library, which every GPU that loads the library holds until exit.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   T > B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T  ->  B200GRPO_EWORKSPACE with a message.
 * All calls on a device share the one array: calls that overlap on the GPU (different
 * streams or threads) race -- serialise them.  Nothing in it needs clearing between calls. */
#define B200GRPO_SEGMENT_BOUNDS_STATIC_MAX_T (1 << 22) /* 4,194,304 tokens */
int b200grpo_segment_bounds_static(const int64_t* position_ids, int T, int32_t* bounds,
                                   int32_t* n_bounds, void* stream);

/* ---------------------------------------------------------------------------------------
 * K2  lm_head backward.  Replaces autograd through rl/__init__.py:197-208:
 *   dlogit[t,j] = (s/temperature) * ( coef_lp[t]*(1[j==target_t] - p[t,j])
 *                                   - coef_ent[t]*p[t,j]*(logp[t,j] + entropy[t]) )
 * written as bf16 scratch, then dH = dlogits . W and dW = dlogits^T . hidden as tcgen05
 * GEMMs.  Two ways to get the logits back:
 *   saved == NULL : recomputed tile by tile from hidden/weight and lse (a third GEMM; the
 *                   forward stored nothing of size [M x Vloc]);
 *   saved != NULL : the buffer the forward filled; turned into dlogits IN PLACE by one
 *                   HBM-bound pass (4 B/element) -- the buffer is consumed by this call.
 *                   With coef_ent != NULL this pass works from bf16 exponentials: fine for the policy term,
 *                   but on near-deterministic rows (entropy << 1e-2 nats) the entropy term p*(log p + H) at
 *                   the peak token is a difference of nearly equal numbers and can be off by tens of per
 *                   cent of that (tiny) term.  b200grpo_rl_step_host and the Python drop-in therefore run an
 *                   entropy gradient in recompute mode (saved == NULL) and keep `saved` for coef_ent == NULL.
 * s = *grad_scale (device scalar, nullable = 1).
 *   grad_hidden bf16 [M,H] (ldgh)   -- partial sum over this rank's vocab rows
 *   grad_weight bf16 or f32 [Vloc,H] (ldgw); accumulate != 0 adds into it (f32 only)
 * ------------------------------------------------------------------------------------- */
size_t b200grpo_lmhead_bwd_workspace(int M, int H, int Vloc, int with_saved);
int b200grpo_lmhead_bwd(const void* hidden, int64_t ldh, const void* weight, int64_t ldw,
                        const int64_t* targets, const float* lse, const float* entropy,
                        const float* coef_lp, const float* coef_ent, const float* grad_scale,
                        int M, int H, int Vloc, int64_t v_offset, float inv_temperature,
                        void* saved, void* grad_hidden, int64_t ldgh, void* grad_weight, int64_t ldgw,
                        int grad_weight_is_f32, int accumulate, void* workspace,
                        size_t workspace_bytes, void* stream);

/* The same backward in stages, for callers that put something between the two gradient GEMMs -- the
 * vocab-parallel path all-reduces grad_hidden over NCCL while grad_weight is still being computed:
 *   call 1: stages = B200GRPO_BWD_DLOGITS | B200GRPO_BWD_DH   (forms dlogits, writes grad_hidden)
 *   call 2: stages = B200GRPO_BWD_DW                          (reuses the dlogits of call 1)
 * with the same saved / workspace buffers in both.  The per-token inputs may be NULL when
 * B200GRPO_BWD_DLOGITS is not requested. */
enum { B200GRPO_BWD_DLOGITS = 1, B200GRPO_BWD_DH = 2, B200GRPO_BWD_DW = 4 };
int b200grpo_lmhead_bwd_staged(const void* hidden, int64_t ldh, const void* weight, int64_t ldw,
                        const int64_t* targets, const float* lse, const float* entropy,
                        const float* coef_lp, const float* coef_ent, const float* grad_scale,
                        int M, int H, int Vloc, int64_t v_offset, float inv_temperature,
                        void* saved, void* grad_hidden, int64_t ldgh, void* grad_weight, int64_t ldgw,
                        int grad_weight_is_f32, int accumulate, int stages, void* workspace,
                        size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * K4  group-normalised advantages.  Replaces the pandas path rl/__init__.py:394-499.
 * Per rollout row i (n_rows): reward f64 (rewards[0] of the row, :402), group_key i32
 * (dense id of group_id+"_"+step_index, :395), row_offsets i64[n_rows+1] (token ranges),
 * n_labels i32 (:487), has_eos u8 (:481-484).  row_order i32[n_rows] = the row indices stably
 * sorted by group_key (dataset order inside a group) and group_starts i32[n_groups+1] = where
 * each group begins in it: one thread per group folds its rows in that order, O(n_rows) reads.
 * Group mean / ddof=1 std / mean length follow pandas' groupby kernels operation by operation
 * (Kahan-compensated mean, Welford variance, f64, no FMA contraction; :408-416), so statistics
 * and advantages are bit-identical to the reference's; a singleton group has std NaN -> 0 (:466).
 * token_rewards f64[total] (nullable: every token of a row then carries row_reward, the
 * documented assumption at :401) feeds the numerator (:466).
 * Outputs f32[total]: advantages, group_tokens, num_labels, overflow (the device batch columns);
 * advantages_f64 (nullable) f64[total]: the un-rounded values the reference's lists hold.
 * ------------------------------------------------------------------------------------- */
size_t b200grpo_group_advantage_workspace(int n_rows, int n_groups);
int b200grpo_group_advantage(const double* row_reward, const int32_t* group_key,
                             const int32_t* row_order, const int32_t* group_starts,
                             const int64_t* row_offsets, const int32_t* n_labels,
                             const uint8_t* has_eos, const double* token_rewards, int n_rows,
                             int n_groups, int divide_by_std, float* advantages,
                             float* group_tokens, float* num_labels, float* overflow,
                             double* advantages_f64, void* workspace, size_t workspace_bytes,
                             void* stream);
/* Same statistics, one f64 value per ROLLOUT: advantages[n_rows] and group_tokens[n_rows] (mean row length of the
 * row's group) -- what the list-of-dict contract of populate_rl_data holds as Python floats when rewards are constant
 * inside a rollout (the host repeats the value; num_labels and overflow are host-side facts). */
int b200grpo_group_advantage_rows(const double* row_reward, const int32_t* group_key,
                                  const int32_t* row_order, const int32_t* group_starts,
                                  const int64_t* row_offsets, int n_rows, int n_groups, int divide_by_std,
                                  double* advantages, double* group_tokens, void* workspace,
                                  size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * K6  weight-update buckets.  Replaces the per-parameter loop of
 * finetune_loop.py:236-248 (sender: param.data.bfloat16() + broadcast) and
 * vllm1.py:81-94 (receiver: torch.empty + broadcast + load_weights) with a gather of many
 * parameters into one contiguous bf16 bucket (and the scatter back).  The NCCL broadcast
 * of the bucket itself is issued by the host layer on the same stream.
 *   src_ptrs_host / dst_ptrs_host: n device pointers (host array); numel_host: n counts;
 *   src_is_f32_host: per-tensor flag (fp32 master weights are cast to bf16 while packing).
 *   table workspace: device scratch the call fills asynchronously from the host arrays
 *   (the host arrays must stay valid until the stream reaches the kernel).
 * ------------------------------------------------------------------------------------- */
size_t b200grpo_bucket_workspace(int n_tensors);
/* elements the bucket needs: every tensor is padded to a multiple of 8 elements (16 B) */
int64_t b200grpo_bucket_numel(const int64_t* numel_host, int n_tensors);
int b200grpo_bucket_pack(const void* const* src_ptrs_host, const int64_t* numel_host,
                         const uint8_t* src_is_f32_host, int n_tensors, void* bucket_bf16,
                         int64_t bucket_capacity, void* workspace, size_t workspace_bytes,
                         void* stream);
int b200grpo_bucket_unpack(const void* bucket_bf16, void* const* dst_ptrs_host,
                           const int64_t* numel_host, const uint8_t* dst_is_f32_host,
                           int n_tensors, void* workspace, size_t workspace_bytes, void* stream);
/* Pack + broadcast in one pass: bucket_multicast_bf16 is the NVSwitch multicast address of a symmetric
 * buffer every rank of the update group owns (cuMulticast* / torch symmetric memory); the kernel reads the
 * parameters once and stores with multimem.st, so the bytes land in every rank's copy of the bucket
 * (the sender's included).  The caller orders it against the receivers with a barrier on the group. */
int b200grpo_bucket_pack_multicast(const void* const* src_ptrs_host, const int64_t* numel_host,
                                   const uint8_t* src_is_f32_host, int n_tensors,
                                   void* bucket_multicast_bf16, int64_t bucket_capacity, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* Scatter + relay: a 1 -> N broadcast at the line rate of one copy.  The bucket is cut into stripes of stripe_elems bf16
 * elements (a multiple of 8).  b200grpo_bucket_pack_scatter (sender) packs like b200grpo_bucket_pack but stores stripe j
 * into peer_buckets_host[j % n_peers] + offset -- device pointers to the receivers' copies of a symmetric slot, P2P
 * mapped (plain st.global over NVLink; no staging buffer on the sender).  After a barrier every receiver calls
 * b200grpo_bucket_relay_multicast: it re-emits the stripes it owns (j % n_peers == peer_index) from its local copy to
 * the NVSwitch multicast address of the slot (multimem.st), which completes every other copy.  After a second barrier
 * each receiver holds the whole bucket.  Sender egress = receiver ingress = one payload, for any N.
 * Replaces the same per-parameter broadcast loop (finetune_loop.py:236-248, vllm1.py:81-94). */
size_t b200grpo_bucket_scatter_workspace(int n_tensors, int n_peers);
int b200grpo_bucket_pack_scatter(const void* const* src_ptrs_host, const int64_t* numel_host,
                                 const uint8_t* src_is_f32_host, int n_tensors,
                                 void* const* peer_buckets_host, int n_peers, int64_t stripe_elems,
                                 int64_t bucket_capacity, void* workspace, size_t workspace_bytes,
                                 void* stream);
int b200grpo_bucket_relay_multicast(const void* local_bucket, void* bucket_multicast, int64_t numel,
                                    int peer_index, int n_peers, int64_t stripe_elems, int ctas,
                                    void* stream);
/* The same relay over plain P2P stores (no multicast mapping needed): the owned stripes are written into every other
 * receiver's copy, peer_buckets_host[q] (q != peer_index), n_peers <= 16. */
int b200grpo_bucket_relay_p2p(const void* local_bucket, void* const* peer_buckets_host, int64_t numel,
                              int peer_index, int n_peers, int64_t stripe_elems, int ctas, void* stream);

/* ---------------------------------------------------------------------------------------
 * Host-buffer entry point (the end-to-end call): one GRPO logprob+loss forward/backward
 * over a packed row whose inputs live in PINNED HOST memory.  Copies inputs to the device
 * workspace, runs K1 -> finalize -> K3 -> K2 and copies metrics (f64[B200GRPO_M_COUNT])
 * back, all on `stream`.  weight / grad_weight stay device resident (model state).
 * grad_hidden is always computed (the trunk's backward consumes it on the device, at
 * b200grpo_rl_step_host_grad_hidden()); it is copied to the host only when
 * grad_hidden_host_bf16 is given.  Same math as calling the entry points above one by one.
 * ------------------------------------------------------------------------------------- */
/* save_logits != 0: recompute-free backward (forward keeps bf16 logits in the workspace). */
typedef struct b200grpo_host_batch {
  const void* hidden_bf16;       /* [L, H]  pinned host */
  const int64_t* input_ids;      /* [L] */
  const int64_t* labels;         /* [L] */
  const int64_t* position_ids;   /* [L] */
  const float* rewards;          /* [L] ... all f32 [L] below */
  const float* advantages;
  const float* ref_logprobs;
  const float* old_logprobs;
  const float* group_tokens;
  const float* num_labels;
  const float* overflow;
} b200grpo_host_batch;

size_t b200grpo_rl_step_host_workspace(int L, int H, int V);
/* device address of grad_hidden bf16 [L,H] inside `workspace` (row L-1 is zero) */
void* b200grpo_rl_step_host_grad_hidden(void* workspace, int L, int H, int V);
int b200grpo_rl_step_host(const b200grpo_host_batch* batch_host, int L, int H, int V,
                          const void* weight, int64_t ldw, float temperature,
                          const b200grpo_loss_cfg* cfg_host, void* grad_hidden_host_bf16,
                          void* grad_weight, int grad_weight_is_f32, int accumulate,
                          double* metrics_host, void* workspace, size_t workspace_bytes,
                          void* stream);

/* ---------------------------------------------------------------------------------------
 * Building block exposed for tests: C[M,N] = A . B^T-style GEMM on the same tcgen05
 * mainloop the lm_head kernels use.  a_mn_major / b_mn_major select the operand layout:
 *   0: operand is [rows = M or N, cols = K] (K contiguous)
 *   1: operand is [rows = K, cols = M or N] (M/N contiguous)
 * C is f32 (c_is_f32) or bf16, row-major [M, ldc]; accumulate adds into f32 C.
 * group_m: row blocks per raster group (0 = all); k_splits > 1 splits K into f32 partial slabs
 * (workspace >= k_splits*M*N*4 bytes, N % 4 == 0) reduced by a second kernel; k_splits == 0 cuts only the
 * tiles of the last partial wave along K (workspace >= one f32 tile per cluster, 40 MB covers every case;
 * a smaller workspace silently disables the cut).
 * ------------------------------------------------------------------------------------- */
int b200grpo_gemm_bf16(const void* A, int64_t lda, int a_mn_major, const void* B, int64_t ldb,
                       int b_mn_major, void* C, int64_t ldc, int c_is_f32, int accumulate,
                       int M, int N, int K, int group_m, int k_splits, void* workspace,
                       size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * Packed-batch ingest.  Replaces pipelinerl/finetune/data.py:215-279
 * (collate_packed materialising eleven [1,T] host tensors, 60 B per token) + their H2D copies
 * (finetune_loop.py:109) + the per-token broadcast of the per-rollout advantage columns
 * (rl/__init__.py:481-499).  Wire form, already on the device (one H2D copy of one pinned
 * buffer): per token input_ids i32, labels i32 (-100 = masked, already masked at every rollout
 * start after the first as data.py:260-261 does), old / ref log-probs f32; per rollout
 * seq_boundaries i32[n_seq + 1] (token offsets, non-decreasing, [0] = 0, [n_seq] = T) and
 * seq_scalars f32[5][n_seq] = rewards, advantages, group_tokens, num_labels, overflow.
 * Output: the eleven tensors of the batch (i64 ids / labels / attention_mask (= 1) /
 * position_ids (restart per rollout), f32 columns), each T long.
 * ------------------------------------------------------------------------------------- */
typedef struct b200grpo_packed_out {
  int64_t *input_ids, *labels, *attention_mask, *position_ids;
  float *rewards, *advantages, *ref_logprobs, *old_logprobs, *group_tokens, *num_labels, *overflow;
} b200grpo_packed_out;
int b200grpo_expand_packed(const int32_t* input_ids, const int32_t* labels, const float* old_logprobs,
                           const float* ref_logprobs, const int32_t* seq_boundaries,
                           const float* seq_scalars, int T, int n_seq, const b200grpo_packed_out* out,
                           void* stream);

/* ---------------------------------------------------------------------------------------
 * Step-metrics accumulator.  Replaces, per micro-batch, the python
 * `rl_metrics[k].append(v)` of pipelinerl/finetune_loop.py:669-670 and, per optimizer step, the
 * pickled all_gather_object of :62-87 + aggregate_rl_stats (rl/utils.py:8-22): every
 * micro-batch folds its K3 metrics vector into acc on the device (no host sync); at the step
 * boundary the caller exchanges acc (one fixed-size collective) and combines by kind.
 *
 * n_stats keys in the order of the reference's stats dict (rl/__init__.py:325-375).  Per key:
 * slots[k] >= 0 is a B200GRPO_M_* index, < 0 one of B200GRPO_STAT_KL_COEF (= n_seq * kl_coef),
 * B200GRPO_STAT_ENTROPY_COEF (= n_seq * entropy_coef), B200GRPO_STAT_INPUT_SIZE;
 * kinds[k] is B200GRPO_STAT_SUM / _MIN / _MAX (chosen from the key NAME by the caller, as
 * aggregate_rl_stats does).  Non-finite values are skipped (finetune_loop.py:85); a micro-batch
 * without labels contributes input_size only (rl/__init__.py:195).  record == 0 (sentinel
 * batch, finetune_loop.py:664-666) only advances the micro-batch / sample counters.
 * acc: f64[2*n_stats + 4] = values | counts | flags, labelled micro-batches, micro-batches, samples.
 * ------------------------------------------------------------------------------------- */
enum { B200GRPO_STAT_SUM = 0, B200GRPO_STAT_MIN = 1, B200GRPO_STAT_MAX = 2 };
enum { B200GRPO_STAT_KL_COEF = -1, B200GRPO_STAT_ENTROPY_COEF = -2, B200GRPO_STAT_INPUT_SIZE = -3 };
#define B200GRPO_STATS_MAX 64
int b200grpo_metrics_reset(const int* kinds, int n_stats, double* acc, void* stream);
int b200grpo_metrics_accumulate(const double* metrics, const int* slots, const int* kinds, int n_stats,
                                double kl_coef, double entropy_coef, double input_size,
                                double n_samples, int record, double* acc, void* stream);

/* ---------------------------------------------------------------------------------------
 * Final RMSNorm + value head in one sweep.  Replaces, inside `model(**model_inputs)`
 * of finetune/rl/__init__.py:183-195, the trunk's last op -- the final RMSNorm of the HF model
 * (third-party transformers Qwen2RMSNorm / LlamaRMSNorm: ~7 elementwise launches over [T, H]) --
 * and the value head `Linear(H, 1)` of finetune/value_model.py:40-52,100-106 that reads the same
 * normalised tensor.  Numerics of the transformers module on bf16: f32 variance and scale, the
 * normalised value rounded to bf16, then a bf16 multiply by gamma.
 *
 *   x bf16 [T,H] (ldx)  pre-norm hidden state        gamma bf16 [H]
 *   value_w bf16 [H], value_b bf16 [1]  nullable: no value head (value / dvalue* unused)
 *   y bf16 [T,H] (ldy)  out: gamma * norm(x) -- operand A of b200grpo_lmhead_fwd
 *   rstd f32 [T]        out: rsqrt(mean(x^2) + eps), consumed by the backward
 *   value f32 [T]       out: y . value_w + value_b
 * Backward: dy bf16 [T,H] (gradient of y from K2's dH), dvalue f32 [T] (K3's coef_value), ->
 *   dx bf16 [T,H], dgamma f32 [H], dvalue_w f32 [H], dvalue_b f32 [1]; the column sums are
 *   reduced in fixed order (deterministic).  H % 8 == 0, H <= 8192, 16-byte aligned rows.
 * ------------------------------------------------------------------------------------- */
size_t b200grpo_final_norm_bwd_workspace(int T, int H);
int b200grpo_final_norm_fwd(const void* x, int64_t ldx, const void* gamma, float eps,
                            const void* value_w, const void* value_b, int T, int H, void* y,
                            int64_t ldy, float* rstd, float* value, void* stream);
int b200grpo_final_norm_bwd(const void* x, int64_t ldx, const void* gamma, const float* rstd,
                            const void* dy, int64_t lddy, const float* dvalue, const void* value_w,
                            int T, int H, void* dx, int64_t lddx, float* dgamma, float* dvalue_w,
                            float* dvalue_b, void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * K5  vocab-parallel exchange over NVSwitch multicast (a new capability -- the reference
 * always holds the whole lm_head, finetune/rl/__init__.py:197).  Both calls take addresses of a
 * SYMMETRIC buffer mapped through a multicast address (the caller allocates / rendezvous-es it,
 * e.g. torch symmetric memory) and leave cross-rank ordering to the caller's barrier on `stream`.
 *
 * b200grpo_lmhead_fwd_multicast: b200grpo_lmhead_fwd whose merge kernel stores the per-token
 *   partials with multimem.st to stats_mc = multicast address of THIS rank's [M,4] slot of a
 *   symmetric [R][M][4] f32 buffer: the all-gather of the forward exchange is the kernel's own
 *   store (16 B/token/rank through the switch).  After a barrier every rank finalizes its local
 *   copy with b200grpo_lmhead_finalize(R, M).
 * b200grpo_mc_allreduce_bf16: in place sum over ranks of a symmetric bf16 tensor (dH = sum over
 *   vocab shards; the dH GEMM writes its partial straight into the local copy).  Two-shot: rank r
 *   owns 1/R of the 16-byte vectors, multimem.ld_reduce (f32 accumulate in the switch) then
 *   multimem.st to every copy.  `ctas` (<= 0: 16) blocks of 512 threads, no shared memory: they
 *   co-reside with the dW GEMM instead of taking SMs from it.  Barrier before and after.
 * ------------------------------------------------------------------------------------- */
int b200grpo_lmhead_fwd_multicast(const void* hidden, int64_t ldh, const void* weight, int64_t ldw,
                                  const int64_t* targets, int M, int H, int Vloc, int64_t v_offset,
                                  float inv_temperature, void* stats_mc, void* saved, void* workspace,
                                  size_t workspace_bytes, void* stream);
int b200grpo_mc_allreduce_bf16(void* mc_ptr, int64_t numel, int rank, int world, int ctas, void* stream);

/* Tunables (for experiments from the test/bench harness; defaults are the shipped ones).
 * name in {"two_cta", "dh_ksplit", "dw_ksplit", "tail_split", "fwd_group_m", "dh_group_m", "dw_group_m", "dw_group_n",
 * "save_logits", "hint_resident", "hint_stream", "dh_big_m", "dw_big_m", "gemm_big_m", "fwd_big_m", "dw_zigzag"} (meanings: csrc/abi.cu);
 * returns the previous value, or a negative B200GRPO_E* code for an unknown name. */
int b200grpo_set_tunable(const char* name, int value);

/* Kernels launched by this library since it was loaded (every <<<>>> / cudaLaunchKernelEx it issues; copies
 * and memsets are not counted).  bench.py differences it around the timed region for "gpu_launches". */
uint64_t b200grpo_launch_count(void);

/* Per-launch device timing (diagnostic, single-threaded): while enabled, every kernel group the
 * library launches is bracketed by CUDA events on its stream.  read() waits for the recorded
 * events, writes up to `capacity` (tag, milliseconds) pairs in launch order, clears the ring and
 * returns the count.  Tags: 1 fwd GEMM+stats, 2 split merge, 3 finalize, 4 token loss,
 * 5 bwd recompute->dlogits GEMM, 6 dH GEMM, 7 dW GEMM, 8 plain GEMM, 9 advantages, 10 bucket,
 * 11 segments, 12 fused logits GEMM, 13 logits->dlogits pass, 14 split-K reduce, 15 ingest expand,
 * 16 final norm (+ value head), 17 multicast all-reduce.
 * enable(0) switches it off. */
int b200grpo_profile_enable(int capacity);
int b200grpo_profile_read(int32_t* tags_host, float* ms_host, int capacity);

#ifdef __cplusplus
}
#endif
#endif /* B200GRPO_H_ */
