// K3: per-token GRPO loss, every reduction and the gradient coefficients in one launch.
// Follows /root/reference/pipelinerl/finetune/rl/__init__.py:152-181,213-375 and
// rl/utils.py:25-30,66-87 (see include/b200grpo.h for the slot layout).
//
// HBM-bound byte work: 60 B/token algorithmic (9 f32 + 2 i64 in, 2 f32 out).  Grid is a
// multiple of the SM count, coalesced loads all issued before the arithmetic, warp-shuffle block reduce,
// deterministic last-block finish (no float atomics), one f64 vector out.
#include "common.cuh"
#include <float.h>

namespace b200grpo {

constexpr int TL_THREADS = 256;
// Measured on B200 at 8 M tokens (tools/bench_small_kernels.py --only k3): one token per thread per trip with three
// resident CTAs per SM (80 registers) reaches 0.52 of the HBM peak; 2 or 4 tokens per trip cost registers, hence
// occupancy, and are slower (0.38 / 0.28)
constexpr int TL_UNROLL = 1;
constexpr int TL_MINB = 3;
constexpr int N_SUM = B200GRPO_M_MAX_REWARD;  // slots [0, N_SUM) reduce by +
constexpr int N_EXT = 10;                     // max/min pairs: reward, adv, kl, tw, value

struct TokenLossArgs {
  const float *logprob, *entropy;
  const int64_t *labels, *position_ids;
  const float *rewards, *advantages, *ref_lp, *old_lp, *group_tokens, *num_labels, *overflow, *values;
  int rows, L;
  b200grpo_loss_cfg cfg;
  double* metrics;
  float *loss, *coef_lp, *coef_ent, *coef_value;
  double* partial_sum;    // [grid][N_SUM]
  float* partial_ext;     // [grid][N_EXT]
  unsigned* partial_flag; // [grid]
  unsigned* ticket;
};

// (value * mask).nan_to_num(0) for one element, rl/utils.py:28-30
__device__ __forceinline__ float masked_term(float v, bool m) {
  if (!m) return 0.f;          // finite*0 = 0 ; inf*0 = nan*0 = NaN -> 0
  if (isnan(v)) return 0.f;
  return fminf(fmaxf(v, -FLT_MAX), FLT_MAX);
}

__global__ void __launch_bounds__(TL_THREADS, TL_MINB) token_loss_kernel(const TokenLossArgs a) {
  const b200grpo_loss_cfg& c = a.cfg;
  const int L = a.L, Ls = a.L - 1;
  const long long n = (long long)a.rows * Ls;
  // per-thread partial sums: f32 for the 24 statistics (a thread sees 1-2 tokens of a real micro-batch, <= ~110 at the
  // 8 M-token bandwidth test), f64 for the two loss terms; everything is f64 from the warp reduction on
  float s[N_SUM];
  double s_loss = 0.0, s_vloss = 0.0;
  float ext[N_EXT];
#pragma unroll
  for (int i = 0; i < N_SUM; ++i) s[i] = 0.f;
#pragma unroll
  for (int i = 0; i < N_EXT; i += 2) { ext[i] = -INFINITY; ext[i + 1] = INFINITY; }
  unsigned flags = 0;
  const float inv_b = 1.0f / (float)c.batch_size;
  const float C = c.clamp_log_ratio_ref_new, eps = c.epsilon;

  const long long stride = (long long)gridDim.x * TL_THREADS;
  for (long long base = (long long)blockIdx.x * TL_THREADS + threadIdx.x; base < n; base += stride * TL_UNROLL) {
    // ---- phase 1: every load of TL_UNROLL tokens is issued before any arithmetic (11-13 independent 4/8-byte
    // coalesced loads per token in flight per thread; the stores below cannot alias them: __restrict__) ----
    float v_lp[TL_UNROLL], v_ent[TL_UNROLL], v_rew[TL_UNROLL], v_ref[TL_UNROLL], v_old[TL_UNROLL], v_gt[TL_UNROLL],
        v_nl[TL_UNROLL], v_ovf[TL_UNROLL], v_adv[TL_UNROLL], v_val[TL_UNROLL];
    int v_r[TL_UNROLL], v_t[TL_UNROLL];
    bool v_m[TL_UNROLL];
    int v_nseq[TL_UNROLL];
#pragma unroll
    for (int u = 0; u < TL_UNROLL; ++u) {
      const long long i = base + u * stride;
      v_t[u] = -1;
      if (i >= n) continue;
      int r = 0, t = (int)i;                         // packed rows (rows == 1) need no division
      if (a.rows > 1) {
        if (n < (1ll << 31)) { r = (int)((unsigned)i / (unsigned)Ls); t = (int)i - r * Ls; }
        else { r = (int)(i / Ls); t = (int)(i - (long long)r * Ls); }
      }
      const long long j = (long long)r * L + t + 1;  // shifted view [:,1:]
      v_r[u] = r; v_t[u] = t;
      v_lp[u] = __ldg(a.logprob + i); v_ent[u] = __ldg(a.entropy + i);
      v_m[u] = __ldg(a.labels + j) != -100;
      v_rew[u] = __ldg(a.rewards + j); v_ref[u] = __ldg(a.ref_lp + j); v_old[u] = __ldg(a.old_lp + j);
      v_gt[u] = __ldg(a.group_tokens + j); v_nl[u] = __ldg(a.num_labels + j);
      v_ovf[u] = c.overlong_filtering ? __ldg(a.overflow + j) : 0.f;
      v_val[u] = c.has_values ? __ldg(a.values + (j - 1)) : 0.f;
      v_adv[u] = c.has_values ? 0.f : __ldg(a.advantages + j);
      int ns = 0;
      if (c.is_packed) {  // segment starts in un-shifted coordinates (:159-171)
        if (t == 0 || __ldg(a.position_ids + (j - 1)) == 0) ++ns;
        if (t == Ls - 1 && __ldg(a.position_ids + j) == 0) ++ns;
      }
      v_nseq[u] = ns;
    }
    // ---- phase 2: arithmetic, in the same per-thread token order as the loads ----
#pragma unroll
    for (int u = 0; u < TL_UNROLL; ++u) {
      if (v_t[u] < 0) continue;
      const long long i = base + u * stride;
      const int r = v_r[u], t = v_t[u];
      const float lp = v_lp[u], ent = v_ent[u];
      const bool m = v_m[u];
      const float rew = v_rew[u], ref = v_ref[u], old = v_old[u];
      const float gt = v_gt[u], nl = v_nl[u];
      s[B200GRPO_M_N_SEQ] += (float)v_nseq[u];
      float tw;
      if (c.group_normalization) {
        if (!(gt > 0.f)) flags |= B200GRPO_F_GROUP_TOKENS_LE0;
        tw = 1.0f / gt;
      } else {
        tw = inv_b;
      }
      if (c.overlong_filtering) tw = tw * (1.0f - v_ovf[u]);
      if (!isfinite(lp)) flags |= B200GRPO_F_NEW_LP_NONFINITE;
      const float d_no = lp - old;
      float ratio = expf(d_no);
      const float d_rn = ref - lp;
      if (!isfinite(d_rn)) flags |= B200GRPO_F_REF_NEW_NONFINITE;
      float v_pred = 0.f, adv;
      if (c.has_values) { v_pred = v_val[u]; adv = rew - v_pred; }
      else adv = v_adv[u];
      float w = c.use_advantages ? adv : rew;
      if (c.relu_log_p_weights) w = fmaxf(w, 0.f);
      const bool kl_clamped = fabsf(d_rn) > C;
      const float cc = fminf(fmaxf(d_rn, -C), C);
      const float e_cc = expf(cc);
      const float kl = e_cc - cc - 1.0f;
      if (!isfinite(kl)) flags |= B200GRPO_F_KL_NONFINITE;
      float surr1 = 0.f, surr2 = 0.f, policy, dpol;  // dpol = d policy / d logprob
      bool r_clamped;
      if (c.policy_loss == 0) {
        surr1 = ratio * w;
        const float cr = fminf(fmaxf(ratio, 1.0f - eps), 1.0f + eps);
        r_clamped = cr != ratio;
        surr2 = cr * w;
        policy = fminf(surr1, surr2);
        const float g1 = ratio * w;
        const float g2 = (ratio >= 1.0f - eps && ratio <= 1.0f + eps) ? g1 : 0.f;  // clamp passes grad on inclusive bounds
        dpol = surr1 < surr2 ? g1 : (surr1 > surr2 ? g2 : 0.5f * (g1 + g2));       // torch.min splits ties
      } else {
        r_clamped = ratio > 1.0f + eps;
        ratio = fminf(fmaxf(ratio, 0.f), 1.0f + eps);  // rebinding: later stats see the clamp
        policy = lp * w * ratio;
        dpol = w * ratio;
      }
      const float per_tok = (policy - c.kl_coef * kl + c.entropy_coef * ent) * tw;
      if (isnan(per_tok)) s[B200GRPO_M_NUM_NANS] += 1.f;
      s_loss += (double)masked_term(per_tok, m);
      // gradient coefficients for an upstream gradient of 1
      const float inside = (d_rn >= -C && d_rn <= C) ? 1.f : 0.f;
      const float mw = m ? tw : 0.f;
      a.coef_lp[i] = -mw * (dpol + c.kl_coef * (e_cc - 1.0f) * inside);
      a.coef_ent[i] = -mw * c.entropy_coef;
      if (c.has_values) {
        const float dv = v_pred - rew;
        s_vloss += (double)masked_term(0.5f * dv * dv * tw, m);
        s[B200GRPO_M_VALUE_MEAN] += masked_term(v_pred / nl, m);
        s[B200GRPO_M_VALUE_MSE] += masked_term(dv * dv / nl, m);
        if (a.coef_value) {
          a.coef_value[(long long)r * L + t] = c.value_loss_coef * dv * mw;
          if (t == Ls - 1) a.coef_value[(long long)r * L + t + 1] = 0.f;
        }
        if (m) { ext[8] = fmaxf(ext[8], v_pred); ext[9] = fminf(ext[9], v_pred); }
      }
      if (m) {  // masked-out tokens contribute exact zeros to every sum below (masked_term): skip them
        s[B200GRPO_M_N_MASKED] += 1.f;
        ext[0] = fmaxf(ext[0], rew); ext[1] = fminf(ext[1], rew);
        ext[2] = fmaxf(ext[2], adv); ext[3] = fminf(ext[3], adv);
        ext[4] = fmaxf(ext[4], kl);  ext[5] = fminf(ext[5], kl);
        ext[6] = fmaxf(ext[6], tw);  ext[7] = fminf(ext[7], tw);
        // x / num_labels as x * (1 / num_labels): one division per token instead of seventeen (<= 1 ulp apart)
        const float inv_nl = 1.0f / nl;
        s[B200GRPO_M_REWARD] += masked_term(rew * inv_nl, true);
        s[B200GRPO_M_ENTROPY] += masked_term(ent * inv_nl, true);
        s[B200GRPO_M_OLD_LP] += masked_term(old * inv_nl, true);
        s[B200GRPO_M_NEW_LP] += masked_term(lp * inv_nl, true);
        s[B200GRPO_M_REF_LP] += masked_term(ref * inv_nl, true);
        s[B200GRPO_M_ADVANTAGE] += masked_term(adv * inv_nl, true);
        s[B200GRPO_M_KL] += masked_term(kl * inv_nl, true);
        s[B200GRPO_M_POLICY] += masked_term(policy * inv_nl, true);
        s[B200GRPO_M_SURR1] += masked_term(surr1 * inv_nl, true);
        s[B200GRPO_M_SURR2] += masked_term(surr2 * inv_nl, true);
        s[B200GRPO_M_RATIO] += masked_term(ratio * inv_nl, true);
        s[B200GRPO_M_RATIO_SUM] += masked_term(ratio, true);
        s[B200GRPO_M_RATIO_SQ_SUM] += masked_term(ratio * ratio, true);
        s[B200GRPO_M_RATIO_REF_NEW] += masked_term(expf(d_rn) * inv_nl, true);
        s[B200GRPO_M_RATIO_REF_OLD] += masked_term(expf(ref - old) * inv_nl, true);
        s[B200GRPO_M_CLAMP_REF_NEW] += masked_term(kl_clamped ? inv_nl : 0.f, true);
        s[B200GRPO_M_CLAMP_NEW_OLD] += masked_term(r_clamped ? inv_nl : 0.f, true);
        s[B200GRPO_M_TOKEN_WEIGHT] += masked_term(tw * inv_nl, true);
      }
    }
  }

  // ---- block reduce: warp shuffles, then one warp over the 8 warp partials --------------
  __shared__ double sh_sum[TL_THREADS / 32][N_SUM];
  __shared__ float sh_ext[TL_THREADS / 32][N_EXT];
  __shared__ unsigned sh_flag[TL_THREADS / 32];
  __shared__ bool is_last;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < N_SUM; ++i) {
    const double v = warp_sum(i == B200GRPO_M_LOSS ? s_loss : i == B200GRPO_M_VALUE_LOSS ? s_vloss : (double)s[i]);
    if (lane == 0) sh_sum[warp][i] = v;
  }
#pragma unroll
  for (int i = 0; i < N_EXT; i += 2) {
    const float mx = warp_max(ext[i]), mn = warp_min(ext[i + 1]);
    if (lane == 0) { sh_ext[warp][i] = mx; sh_ext[warp][i + 1] = mn; }
  }
  flags = warp_or(flags);
  if (lane == 0) sh_flag[warp] = flags;
  __syncthreads();
  if (threadIdx.x < N_SUM) {
    double v = 0.0;
    for (int w = 0; w < TL_THREADS / 32; ++w) v += sh_sum[w][threadIdx.x];
    a.partial_sum[(size_t)blockIdx.x * N_SUM + threadIdx.x] = v;
  } else if (threadIdx.x >= 32 && threadIdx.x < 32 + N_EXT) {
    const int k = threadIdx.x - 32;
    float v = sh_ext[0][k];
    for (int w = 1; w < TL_THREADS / 32; ++w) v = (k & 1) ? fminf(v, sh_ext[w][k]) : fmaxf(v, sh_ext[w][k]);
    a.partial_ext[(size_t)blockIdx.x * N_EXT + k] = v;
  } else if (threadIdx.x == 64) {
    unsigned f = 0;
    for (int w = 0; w < TL_THREADS / 32; ++w) f |= sh_flag[w];
    a.partial_flag[blockIdx.x] = f;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(a.ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();

  // ---- last block: fixed-order reduction over block partials (deterministic) ------------
  const int G = gridDim.x;
  if (threadIdx.x < N_SUM) {
    double v = 0.0;
    for (int b = 0; b < G; ++b) v += a.partial_sum[(size_t)b * N_SUM + threadIdx.x];
    sh_sum[0][threadIdx.x] = v;
  } else if (threadIdx.x >= 32 && threadIdx.x < 32 + N_EXT) {
    const int k = threadIdx.x - 32;
    float v = a.partial_ext[k];
    for (int b = 1; b < G; ++b) {
      const float p = a.partial_ext[(size_t)b * N_EXT + k];
      v = (k & 1) ? fminf(v, p) : fmaxf(v, p);
    }
    sh_ext[0][k] = v;
  } else if (threadIdx.x == 64) {
    unsigned f = 0;
    for (int b = 0; b < G; ++b) f |= a.partial_flag[b];
    sh_flag[0] = f;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double* M = a.metrics;
    for (int i = 0; i < N_SUM; ++i) M[i] = sh_sum[0][i];
    // final_loss = -sum_sum(loss) + value_loss_coef * value_loss   (:292,:308)
    const float policy_total = -(float)sh_sum[0][B200GRPO_M_LOSS];
    float loss = policy_total;
    if (c.has_values) loss = policy_total + c.value_loss_coef * (float)sh_sum[0][B200GRPO_M_VALUE_LOSS];
    M[B200GRPO_M_LOSS] = (double)loss;
    if (a.loss) *a.loss = loss;
    unsigned f = sh_flag[0];
    if (!isfinite(loss)) f |= B200GRPO_F_LOSS_NONFINITE;
    if (!c.is_packed) M[B200GRPO_M_N_SEQ] = (double)a.rows;
    else if (n == 0) M[B200GRPO_M_N_SEQ] = 1.0;
    for (int k = 0; k < N_EXT; ++k) M[B200GRPO_M_MAX_REWARD + k] = (double)sh_ext[0][k];
    M[B200GRPO_M_FLAGS] = (double)f;
    for (int i = B200GRPO_M_FLAGS + 1; i < B200GRPO_M_COUNT; ++i) M[i] = 0.0;
    *a.ticket = 0u;  // self-cleaning for the next launch on this workspace
  }
}

static int token_loss_grid(long long n) {
  const int sms = sm_count();
  long long want = (n + (long long)TL_THREADS * TL_UNROLL - 1) / ((long long)TL_THREADS * TL_UNROLL);
  if (want < 1) want = 1;
  const long long cap = (long long)sms * TL_MINB;  // resident 256-thread CTAs per SM
  if (want > cap) want = cap;
  if (want > sms) want = want / sms * sms;   // whole waves
  return (int)want;
}

static size_t token_loss_ws_layout(int grid, size_t* off_ext, size_t* off_flag, size_t* off_ticket) {
  size_t o = 0;
  o += align_up((size_t)grid * N_SUM * sizeof(double), 256);
  *off_ext = o;
  o += align_up((size_t)grid * N_EXT * sizeof(float), 256);
  *off_flag = o;
  o += align_up((size_t)grid * sizeof(unsigned), 256);
  *off_ticket = o;
  o += 256;
  return o;
}

}  // namespace b200grpo

using namespace b200grpo;

extern "C" size_t b200grpo_token_loss_workspace(int rows, int L) {
  (void)rows; (void)L;
  size_t a, b, c;
  int sms = sm_count();
  if (sms <= 0) sms = 148;
  return token_loss_ws_layout(sms * 8, &a, &b, &c);
}

extern "C" int b200grpo_token_loss(const float* logprob, const float* entropy, const int64_t* labels,
                                   const int64_t* position_ids, const float* rewards, const float* advantages,
                                   const float* ref_logprobs, const float* old_logprobs, const float* group_tokens,
                                   const float* num_labels, const float* overflow, const float* values, int rows,
                                   int L, const b200grpo_loss_cfg* cfg, double* metrics, float* loss, float* coef_lp,
                                   float* coef_ent, float* coef_value, void* workspace, size_t workspace_bytes,
                                   void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(cfg && metrics && coef_lp && coef_ent && workspace, "token_loss: null pointer");
  B200_REQUIRE(rows >= 1 && L >= 1, "token_loss: rows=%d L=%d", rows, L);
  B200_REQUIRE(cfg->policy_loss == 0 || cfg->policy_loss == 1, "Unknown algorithm id %d", cfg->policy_loss);
  B200_REQUIRE(!cfg->is_packed || (rows == 1 && position_ids), "token_loss: packed batch needs rows==1 and position_ids");
  B200_REQUIRE(!cfg->has_values || values, "token_loss: has_values without values");
  const long long n = (long long)rows * (L - 1);
  B200_REQUIRE(n == 0 || (logprob && entropy && labels && rewards && advantages && ref_logprobs && old_logprobs &&
                          group_tokens && num_labels && overflow), "token_loss: null input");
  const int grid = token_loss_grid(n);
  size_t off_ext, off_flag, off_ticket;
  const size_t need = token_loss_ws_layout(grid, &off_ext, &off_flag, &off_ticket);
  if (workspace_bytes < need) { set_error("token_loss: workspace %zu < %zu", workspace_bytes, need); return B200GRPO_EWORKSPACE; }
  char* ws = (char*)workspace;
  TokenLossArgs a;
  a.logprob = logprob; a.entropy = entropy; a.labels = labels; a.position_ids = position_ids;
  a.rewards = rewards; a.advantages = advantages; a.ref_lp = ref_logprobs; a.old_lp = old_logprobs;
  a.group_tokens = group_tokens; a.num_labels = num_labels; a.overflow = overflow; a.values = values;
  a.rows = rows; a.L = L; a.cfg = *cfg; a.metrics = metrics; a.loss = loss;
  a.coef_lp = coef_lp; a.coef_ent = coef_ent; a.coef_value = coef_value;
  a.partial_sum = (double*)ws; a.partial_ext = (float*)(ws + off_ext);
  a.partial_flag = (unsigned*)(ws + off_flag); a.ticket = (unsigned*)(ws + off_ticket);
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA_OK(cudaMemsetAsync(a.ticket, 0, sizeof(unsigned), st));
  ProfScope ps(PROF_TOKEN_LOSS, st);
  token_loss_kernel<<<grid, TL_THREADS, 0, st>>>(a);
  count_launch();
  B200_CUDA_OK(cudaGetLastError());
  return B200GRPO_OK;
}
