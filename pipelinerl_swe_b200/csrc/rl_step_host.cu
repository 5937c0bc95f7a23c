// End-to-end entry with HOST buffers: H2D of one packed row, K1 -> finalize -> K3 -> K2,
// D2H of the metrics (and optionally of grad_hidden), all stream-ordered.  This is the
// call bench.py times for its "e2e" number; it composes the public entry points only.
#include "common.cuh"

using namespace b200grpo;

namespace {
struct Layout {
  size_t hidden, ids, labels, pos, f32[7], stats, logprob, entropy, lse, coef_lp, coef_ent, metrics, loss, ghid, ws_fwd,
      ws_loss, ws_bwd, saved, total;
  size_t n_fwd, n_loss, n_bwd, n_saved;
  int save;
};
Layout make_layout(int L, int H, int V) {
  Layout o;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off += align_up(bytes, 256); return r; };
  const int M = L > 1 ? L - 1 : 1;
  o.hidden = take((size_t)L * H * 2);
  o.ids = take((size_t)L * 8); o.labels = take((size_t)L * 8); o.pos = take((size_t)L * 8);
  for (int i = 0; i < 7; ++i) o.f32[i] = take((size_t)L * 4);
  o.stats = take((size_t)M * 16);
  o.logprob = take((size_t)M * 4); o.entropy = take((size_t)M * 4); o.lse = take((size_t)M * 4);
  o.coef_lp = take((size_t)M * 4); o.coef_ent = take((size_t)M * 4);
  o.metrics = take(B200GRPO_M_COUNT * 8); o.loss = take(4);
  o.ghid = take((size_t)L * H * 2);
  o.n_fwd = b200grpo_lmhead_fwd_workspace(M, H, V); o.ws_fwd = take(o.n_fwd);
  o.n_loss = b200grpo_token_loss_workspace(1, L); o.ws_loss = take(o.n_loss);
  o.save = get_tunable(TUNE_SAVE_LOGITS) != 0;
  o.n_bwd = b200grpo_lmhead_bwd_workspace(M, H, V, o.save); o.ws_bwd = take(o.n_bwd);
  o.n_saved = o.save ? b200grpo_lmhead_saved_bytes(M, H, V) : 0; o.saved = take(o.n_saved);
  o.total = off;
  return o;
}
}  // namespace

extern "C" size_t b200grpo_rl_step_host_workspace(int L, int H, int V) { return make_layout(L, H, V).total; }

extern "C" void* b200grpo_rl_step_host_grad_hidden(void* workspace, int L, int H, int V) {
  return workspace ? (char*)workspace + make_layout(L, H, V).ghid : nullptr;
}

extern "C" int b200grpo_rl_step_host(const b200grpo_host_batch* b, int L, int H, int V, const void* weight, int64_t ldw,
                                     float temperature, const b200grpo_loss_cfg* cfg, void* grad_hidden_host_bf16,
                                     void* grad_weight, int grad_weight_is_f32, int accumulate, double* metrics_host,
                                     void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_device_sm100()) return e;
  B200_REQUIRE(b && cfg && weight && metrics_host && workspace, "rl_step_host: null pointer");
  B200_REQUIRE(L >= 2 && temperature > 0.f, "rl_step_host: L=%d temperature=%f", L, temperature);
  B200_REQUIRE(cfg->is_packed && !cfg->has_values, "rl_step_host: packed GRPO rows only");
  const Layout o = make_layout(L, H, V);
  if (workspace_bytes < o.total) { set_error("rl_step_host: workspace %zu < %zu", workspace_bytes, o.total); return B200GRPO_EWORKSPACE; }
  char* w = (char*)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  const int M = L - 1;
  B200_CUDA_OK(cudaMemcpyAsync(w + o.hidden, b->hidden_bf16, (size_t)L * H * 2, cudaMemcpyHostToDevice, st));
  B200_CUDA_OK(cudaMemcpyAsync(w + o.ids, b->input_ids, (size_t)L * 8, cudaMemcpyHostToDevice, st));
  B200_CUDA_OK(cudaMemcpyAsync(w + o.labels, b->labels, (size_t)L * 8, cudaMemcpyHostToDevice, st));
  B200_CUDA_OK(cudaMemcpyAsync(w + o.pos, b->position_ids, (size_t)L * 8, cudaMemcpyHostToDevice, st));
  const float* f32_src[7] = {b->rewards, b->advantages, b->ref_logprobs, b->old_logprobs, b->group_tokens, b->num_labels, b->overflow};
  for (int i = 0; i < 7; ++i) B200_CUDA_OK(cudaMemcpyAsync(w + o.f32[i], f32_src[i], (size_t)L * 4, cudaMemcpyHostToDevice, st));
  const int64_t* targets = (const int64_t*)(w + o.ids) + 1;  // row t is scored against input_ids[t+1]
  // An entropy gradient is formed from recomputed f32 logits, never from the saved bf16 exponentials: on near-deterministic rows
  // (entropy << 1e-2 nats) p * (log p + H) is a difference of nearly equal numbers at the peak token and bf16's 2^-9 on E ~ 1
  // is as large as H itself.  With entropy_coef != 0 the step therefore runs in recompute mode; the backward's dlogits scratch
  // then takes the (unused) saved-exponentials region that follows its workspace.
  const bool save = o.save && cfg->entropy_coef == 0.f;
  size_t n_bwd = o.n_bwd;
  if (o.save && !save) {
    n_bwd = (size_t)((w + o.saved + o.n_saved) - (w + o.ws_bwd));
    B200_REQUIRE(n_bwd >= b200grpo_lmhead_bwd_workspace(M, H, V, 0), "rl_step_host: workspace too small for the recompute backward");
  }
  int e = b200grpo_lmhead_fwd(w + o.hidden, H, weight, ldw, targets, M, H, V, 0, 1.0f / temperature, (float*)(w + o.stats),
                              save ? (void*)(w + o.saved) : nullptr, w + o.ws_fwd, o.n_fwd, stream);
  if (e) return e;
  e = b200grpo_lmhead_finalize((const float*)(w + o.stats), 1, M, (float*)(w + o.logprob), (float*)(w + o.entropy),
                               (float*)(w + o.lse), stream);
  if (e) return e;
  e = b200grpo_token_loss((const float*)(w + o.logprob), (const float*)(w + o.entropy), (const int64_t*)(w + o.labels),
                          (const int64_t*)(w + o.pos), (const float*)(w + o.f32[0]), (const float*)(w + o.f32[1]),
                          (const float*)(w + o.f32[2]), (const float*)(w + o.f32[3]), (const float*)(w + o.f32[4]),
                          (const float*)(w + o.f32[5]), (const float*)(w + o.f32[6]), nullptr, 1, L, cfg,
                          (double*)(w + o.metrics), (float*)(w + o.loss), (float*)(w + o.coef_lp), (float*)(w + o.coef_ent),
                          nullptr, w + o.ws_loss, o.n_loss, stream);
  if (e) return e;
  if (grad_weight || grad_hidden_host_bf16) {
    e = b200grpo_lmhead_bwd(w + o.hidden, H, weight, ldw, targets, (const float*)(w + o.lse), (const float*)(w + o.entropy),
                            (const float*)(w + o.coef_lp), save ? nullptr : (const float*)(w + o.coef_ent),
                            nullptr, M, H, V, 0,
                            1.0f / temperature, save ? (void*)(w + o.saved) : nullptr, (void*)(w + o.ghid), H, grad_weight, H,
                            grad_weight_is_f32, accumulate, w + o.ws_bwd, n_bwd, stream);
    if (e) return e;
    B200_CUDA_OK(cudaMemsetAsync(w + o.ghid + (size_t)M * H * 2, 0, (size_t)H * 2, st));  // last position has no target
    if (grad_hidden_host_bf16)
      B200_CUDA_OK(cudaMemcpyAsync(grad_hidden_host_bf16, w + o.ghid, (size_t)L * H * 2, cudaMemcpyDeviceToHost, st));
  }
  B200_CUDA_OK(cudaMemcpyAsync(metrics_host, w + o.metrics, B200GRPO_M_COUNT * 8, cudaMemcpyDeviceToHost, st));
  return B200GRPO_OK;
}
