// H2-d — the actor step's loss function as the engine calls it, and its backward, with nothing
// but the two (resp. one) streaming/[T]-sized kernels on the stream.
//
// ref: areal/engine/ppo/actor.py:292-367 (grpo_loss_fn): labels = roll(input_ids, -1);
// gather_logprobs_entropy (+ two more [T, V] passes for the logged row min/max); ppo_actor_loss_fn;
// a dozen derived logging tensors (`.float()` of masks, two `torch.where`). Eager torch needs
// ~40 launches for this; the drop-in path of round 1 needed 2 kernels + ~10 small aten ones
// (roll, bool cast, where, float casts) per micro-batch, which sat between the two big kernels.
// Here the forward is exactly logprob_fwd_kernel + ppo_loss_fwd_kernel: the roll is an index
// computation inside the kernels, the loss kernel reads the engine's fp32 (or bool) mask as it
// is and writes every logged tensor itself. The backward is ONE launch: the scalar loss's
// backward (g * dlogp / N) is evaluated per row inside logprob_bwd_kernel.
#include "common.cuh"

extern "C" {

size_t b200rl_grpo_loss_workspace_bytes(int64_t T) { return b200rl_ppo_loss_workspace_bytes(T); }

int b200rl_grpo_loss_fwd(const void* logits, int dtype, int64_t ld, const int64_t* input_ids,
                         int label_roll, int64_t T, int64_t V, float temperature,
                         const float* prox_logp, const float* old_logp, const float* adv,
                         const void* mask, int mask_is_f32, float clip_lo, float clip_hi,
                         float c_clip, float cap, float* logp, float* entropy, float* lse,
                         float* rowmax, float* rowmin, float* stat_loss, float* stat_imp_weight,
                         float* stat_approx_kl, uint8_t* stat_clip_mask,
                         uint8_t* stat_dual_clip_mask, float* stat_behave_imp_weight,
                         float* stat_behave_approx_kl, uint8_t* stat_behave_mask,
                         uint8_t* stat_valid, float* stat_clip_ratio, float* stat_dual_clip_ratio,
                         float* stat_clipped_new_logp, float* stat_clipped_old_logp, float* dlogp,
                         float* loss, float* denom, void* workspace, size_t workspace_bytes,
                         void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T >= 0, "negative T");
  B200RL_REQUIRE(loss && denom && workspace, "loss/denom/workspace must not be null");
  B200RL_REQUIRE(T == 0 || (logp && lse && dlogp), "logp, lse and dlogp are needed by the backward");
  if (T > 0) {
    int rc = logprob_fwd_impl(logits, dtype, ld, input_ids, T, V, temperature, logp, entropy, lse,
                              rowmax, rowmin, nullptr, entropy != nullptr,
                              rowmax != nullptr || rowmin != nullptr, label_roll != 0, stream);
    if (rc) return rc;
  }
  PpoLossArgs p{};
  p.logp = logp; p.prox = prox_logp; p.old = old_logp; p.adv = adv;
  p.mask = mask_is_f32 ? nullptr : static_cast<const uint8_t*>(mask);
  p.mask_f32 = mask_is_f32 ? static_cast<const float*>(mask) : nullptr;
  p.T = T;
  p.clip_lo = clip_lo; p.clip_hi = clip_hi; p.c_clip = c_clip; p.cap = cap;
  p.s_loss = stat_loss; p.s_iw = stat_imp_weight; p.s_kl = stat_approx_kl;
  p.s_clip = stat_clip_mask; p.s_dual = stat_dual_clip_mask;
  p.s_bw = stat_behave_imp_weight; p.s_bkl = stat_behave_approx_kl; p.s_bm = stat_behave_mask;
  p.s_valid = stat_valid; p.s_clip_ratio = stat_clip_ratio; p.s_dual_ratio = stat_dual_clip_ratio;
  p.s_clip_new = stat_clipped_new_logp; p.s_clip_old = stat_clipped_old_logp;
  p.dlogp = dlogp; p.loss = loss; p.denom = denom;
  return ppo_loss_fwd_launch(p, workspace, workspace_bytes, stream);
}

int b200rl_grpo_loss_bwd(const void* logits, int dtype, int64_t ld, const int64_t* input_ids,
                         int label_roll, int64_t T, int64_t V, float temperature, const float* lse,
                         const float* dlogp, const float* grad_loss, const float* denom,
                         void* dlogits, int64_t ld_out, void* stream) {
  using namespace b200rl;
  B200RL_REQUIRE(T == 0 || (grad_loss && denom), "grad_loss and denom must not be null");
  return logprob_bwd_impl(logits, dtype, ld, input_ids, lse, nullptr, dlogp, nullptr, T, V,
                          temperature, dlogits, ld_out, label_roll != 0, grad_loss, denom, stream);
}

}  // extern "C"
