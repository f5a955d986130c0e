/* b200rl — C ABI of the B200-native per-token RL math library (libb200rl.so).
 *
 * Every entry point is `extern "C"`, takes plain device/host pointers, sizes and a
 * CUDA stream handle (`void*`, a cudaStream_t; NULL = legacy default stream), and
 * returns an int: 0 on success, a positive cudaError_t value if a CUDA call failed,
 * a negative B200RL_E* code for argument errors. `b200rl_last_error()` returns a
 * thread-local, human readable description of the last non-zero return.
 *
 * The library never allocates device memory: the caller owns every buffer,
 * including the scratch ("workspace") buffers whose size the *_workspace_bytes
 * helpers report. (One entry point, b200rl_gae_1d_nolp_misalign_noalloc, takes no
 * workspace and uses a fixed scratch array built into the library image instead.)
 * Nothing here synchronises the host with the device.
 *
 * Citations `ref:` are relative to the AReaL-vllm reference tree and name the
 * reference interface that each entry point replaces.
 */
#ifndef B200RL_H_
#define B200RL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200RL_OK 0
#define B200RL_EINVAL (-1)       /* bad argument (null pointer, negative size, misalignment) */
#define B200RL_EUNSUPPORTED (-2) /* dtype / mode not built */
#define B200RL_EWORKSPACE (-3)   /* workspace too small */

/* element types of a logits / parameter buffer */
#define B200RL_BF16 0
#define B200RL_FP16 1
#define B200RL_FP32 2

const char* b200rl_last_error(void);
int b200rl_version(void);

/* ------------------------------------------------------------------------------------------
 * H1  log-prob / entropy gather over [T, V] logits.
 * ref: areal/utils/functional.py:10-16 (_gather_logprobs), :19-26 (_gather_logprobs_entropy),
 *      :29-80 (chunking wrappers; chunking is unnecessary here, results do not depend on it).
 *
 * fwd:  x = logits/temperature (fp32);  lse[t] = logsumexp_j x[t,j];
 *       logp[t] = x[t,labels[t]] - lse[t];  entropy[t] = -sum_j p_j log p_j.
 *       One streaming pass over the row (online softmax). `entropy`, `rowmax`, `rowmin` may be
 *       NULL. `rowmax/rowmin` are the raw (un-tempered) per-row logits extrema that
 *       grpo_loss_fn logs (ref: areal/engine/ppo/actor.py:351-352).
 * bwd:  dlogits[t,j] = (1/temperature) * ( g_logp[t]*(1[j==y_t] - p_j) - g_ent[t]*p_j*(log p_j + H_t) )
 *       rounded to the logits dtype; the softmax is recomputed from `lse`, never stored.
 *       `g_ent` may be NULL (entropy detached, the GRPO call site); then `entropy` may be NULL too.
 * `ld` / `ld_out` are row strides in elements. `dtype` is B200RL_BF16 / FP16 / FP32.
 */
int b200rl_logprob_entropy_fwd(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               int64_t T, int64_t V, float temperature, float* logp,
                               float* entropy, float* lse, float* rowmax, float* rowmin,
                               void* stream);

int b200rl_logprob_entropy_bwd(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               const float* lse, const float* entropy, const float* g_logp,
                               const float* g_ent, int64_t T, int64_t V, float temperature,
                               void* dlogits, int64_t ld_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * H2-d  the actor step's loss function as the engine calls it, and its backward.
 * ref: areal/engine/ppo/actor.py:292-367 (grpo_loss_fn) and loss.backward()
 *      (areal/engine/fsdp_engine.py:542-550).
 *
 * fwd = b200rl_logprob_entropy_fwd followed by b200rl_ppo_loss_fwd on the fresh log-probs: two
 *       launches and nothing between them.
 *       - `input_ids` with label_roll != 0: the label of row t is input_ids[(t+1) % T], i.e.
 *         torch.roll(input_ids, -1) (actor.py:308) evaluated as an index; label_roll == 0: the
 *         array already holds the labels.
 *       - `mask`: bytes (torch.bool) when mask_is_f32 == 0, else fp32 with non-zero = valid (the
 *         engine's loss_mask is a float tensor that grpo_loss_fn turns into bool, actor.py:304).
 *       - besides everything the two separate entry points write, the derived tensors that
 *         grpo_loss_fn logs (actor.py:339-365): stat_valid (the mask as bytes), stat_clip_ratio /
 *         stat_dual_clip_ratio (the two masks as fp32), stat_clipped_new_logp / _old_logp
 *         (where(clip_mask, logp | old_logp, 0)). Every stat_* pointer, entropy, rowmax and rowmin
 *         may be NULL; logp, lse, dlogp, loss, denom are required (the backward reads them).
 * bwd = ONE launch: dlogits[t, :] = d(grad_loss * loss)/d logits with entropy detached
 *       (actor.py:312): b200rl_ppo_loss_bwd's g_logp[t] = dlogp[t] * (grad_loss[0] / denom[0]) is
 *       evaluated per row inside the log-prob backward kernel (same operation order, same bits as
 *       the two separate calls). grad_loss and denom are DEVICE scalars.
 * workspace: b200rl_grpo_loss_workspace_bytes(T) (= the PPO loss's), cleared by the call.
 */
size_t b200rl_grpo_loss_workspace_bytes(int64_t T);
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
                         void* stream);
int b200rl_grpo_loss_bwd(const void* logits, int dtype, int64_t ld, const int64_t* input_ids,
                         int label_roll, int64_t T, int64_t V, float temperature, const float* lse,
                         const float* dlogp, const float* grad_loss, const float* denom,
                         void* dlogits, int64_t ld_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * H1 over a vocabulary-parallel (tensor-parallel) shard of the logits.
 * ref: areal/experimental/megatron_actor.py:60-66, 323-331 (logits.float()/T, then
 *      megatron.core.tensor_parallel.vocab_parallel_cross_entropy and _VocabParallelEntropy) and
 *      areal/experimental/utils/mcore/functional.py:8-52 (three all-reduces + four passes).
 *
 * b200rl_logprob_shard_fwd: one streaming pass over this rank's [T, V_shard] slice writes, per
 * row, 8 floats to `partial`: (M, S, W, x_y, max, min, 0, 0) - the online-softmax state in the
 * log2 domain (M = max of x*log2e/T, S = sum 2^(.-M), W = sum 2^(.-M)*(.-M)), the raw label
 * logit (NaN when `local_labels[t]` = label - vocab_start is outside [0, V_shard)), and the raw
 * row max / min (need_minmax). After ONE all-gather of the [T, 8] records over the tensor-parallel
 * group, b200rl_logprob_shard_combine folds `partials` [n_shards, T, 8] into the same outputs as
 * b200rl_logprob_entropy_fwd (entropy, rowmax, rowmin nullable). The backward needs no exchange:
 * b200rl_logprob_entropy_bwd on the shard with the global lse / entropy and the local labels
 * (out-of-range labels add no one-hot term).
 */
int b200rl_logprob_shard_fwd(const void* logits, int dtype, int64_t ld, const int64_t* local_labels,
                             int64_t T, int64_t V_shard, float temperature, int need_entropy,
                             int need_minmax, float* partial, void* stream);
int b200rl_logprob_shard_combine(const float* partials, int32_t n_shards, int64_t T,
                                 float temperature, float* logp, float* entropy, float* lse,
                                 float* rowmax, float* rowmin, void* stream);

/* The exchange of the [T, 8] records WITHOUT a collective library call: peer stores over NVLink into
 * a buffer that every rank of the tensor-parallel group (one node) has mapped, e.g. with
 * b200rl_ipc_export / b200rl_ipc_open. No counterpart in the reference, which all-reduces.
 *   b200rl_tp_exchange_bytes(tp, max_tokens): size of one rank's buffer (flags + two record slots);
 *     zero-fill it once, before any rank's first push (the caller synchronises that).
 *   b200rl_tp_push: stores this rank's `partial` [T, 8] into slot (epoch & 1) of EVERY rank's
 *     buffer (`peer_bufs[r]` = rank r's buffer as mapped in THIS process, r = rank is the local
 *     one), then publishes `epoch` in its flags behind a system-scope fence.
 *   b200rl_tp_combine: waits (inside the kernel, bounded to ~2 s, then *status = 1) until every
 *     rank's records of `epoch` have arrived in `my_buf`, then computes the outputs of
 *     b200rl_logprob_shard_combine. `status` is a DEVICE int32 the caller zeroes once.
 * All ranks call push/combine in the same order with epoch = 1, 2, 3, ...; T <= max_tokens and
 * <= 262,144 rows per call, tp <= 16. Two slots suffice because a rank can be at most one call
 * ahead of a peer (its combine of call e needs the peer's push of call e). */
size_t b200rl_tp_exchange_bytes(int32_t tp, int64_t max_tokens);
int b200rl_tp_push(const float* partial, int64_t T, void* const* peer_bufs, int32_t tp, int32_t rank,
                   int64_t max_tokens, uint32_t epoch, void* stream);
int b200rl_tp_combine(void* my_buf, int32_t tp, int64_t T, int64_t max_tokens, uint32_t epoch,
                      float temperature, float* logp, float* entropy, float* lse, float* rowmax,
                      float* rowmin, int32_t* status, void* stream);

/* ------------------------------------------------------------------------------------------
 * H1+H2 fused: the whole GRPO actor step on one packed micro-batch in ONE pass over HBM.
 * ref: areal/engine/ppo/actor.py:292-367 (grpo_loss_fn) followed by loss.backward()
 *      (areal/engine/fsdp_engine.py:542-550).
 *
 * Computes everything b200rl_logprob_entropy_fwd (with entropy and row max/min) and
 * b200rl_ppo_loss_fwd compute, and also dlogits = d(upstream * loss)/d logits, with 4*V instead of
 * 6*V bytes of HBM traffic per token: each row is read once from HBM for the forward statistics
 * and a second time, immediately, out of L2 for the backward. Entropy is treated as detached (it
 * is in grpo_loss_fn, actor.py:312). `upstream` (DEVICE scalar, NULL = 1) is the factor the
 * caller will multiply the loss by before backward(); the engine's `loss_scale` does not depend on
 * the loss (fsdp_engine.py:543-547), so it is known here. `loss[0]` is NOT multiplied by it.
 * `denom[0]` receives N = max(count(mask), 1). Same dtypes/flags as the separate entry points;
 * returns B200RL_EUNSUPPORTED when rows are not 16-byte aligned (use the separate calls then).
 * b200rl_scale_inplace multiplies a gradient buffer by a DEVICE scalar and does nothing (no
 * memory traffic) when that scalar is exactly 1: the autograd backward of the fused step uses it
 * to stay correct for any incoming gradient.
 */
size_t b200rl_grpo_fused_workspace_bytes(int64_t T);
/* b200rl_grpo_fused_step with cap == -3 is the packed SFT step (ref areal/engine/sft/lm_engine.py:
 * 45-61 + backward): per-token loss -logp on the masked tokens, loss = sum / count(mask) with no
 * `or 1`, dlogits = upstream * (softmax - onehot) * mask / count; prox_logp / old_logp / adv must
 * still be valid pointers (their content is ignored), the PPO stat pointers may be NULL. */
int b200rl_grpo_fused_step(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                           int64_t T, int64_t V, float temperature, const float* prox_logp,
                           const float* old_logp, const float* adv, const uint8_t* mask,
                           float clip_lo, float clip_hi, float c_clip, float cap,
                           const float* upstream, float* logp, float* entropy, float* lse,
                           float* rowmax, float* rowmin, float* stat_loss, float* stat_imp_weight,
                           float* stat_approx_kl, uint8_t* stat_clip_mask,
                           uint8_t* stat_dual_clip_mask, float* stat_behave_imp_weight,
                           float* stat_behave_approx_kl, uint8_t* stat_behave_mask, void* dlogits,
                           int64_t ld_out, float* loss, float* denom, void* workspace,
                           size_t workspace_bytes, void* stream);
/* The same step with the conveniences of b200rl_grpo_loss_fwd: `labels` may be the packed input_ids
 * (label_roll != 0: label of row t = labels[(t+1) % T]), the mask may be fp32 (mask_is_f32), and
 * the derived logging tensors of grpo_loss_fn (stat_valid ... stat_clipped_old_logp, nullable) are
 * written by the kernel, so that the engine-facing wrapper launches nothing else. */
int b200rl_grpo_fused_step_ids(const void* logits, int dtype, int64_t ld, const int64_t* labels,
                               int label_roll, int64_t T, int64_t V, float temperature,
                               const float* prox_logp, const float* old_logp, const float* adv,
                               const void* mask, int mask_is_f32, float clip_lo, float clip_hi,
                               float c_clip, float cap, const float* upstream, float* logp,
                               float* entropy, float* lse, float* rowmax, float* rowmin,
                               float* stat_loss, float* stat_imp_weight, float* stat_approx_kl,
                               uint8_t* stat_clip_mask, uint8_t* stat_dual_clip_mask,
                               float* stat_behave_imp_weight, float* stat_behave_approx_kl,
                               uint8_t* stat_behave_mask, uint8_t* stat_valid, float* stat_clip_ratio,
                               float* stat_dual_clip_ratio, float* stat_clipped_new_logp,
                               float* stat_clipped_old_logp, void* dlogits, int64_t ld_out, float* loss,
                               float* denom, void* workspace, size_t workspace_bytes, void* stream);
int b200rl_scale_inplace(void* buf, int dtype, int64_t numel, const float* scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * H3  GAE over packed variable-length sequences ("1d, no-overlap, misaligned").
 * ref: csrc/cugae/gae.cu:10-60 (gae_kernel_1d_nolp_misalign + launcher),
 *      realhf/impl/model/utils/ppo_functional.py:292-366 (pygae1d_nolp_misalign, cugae1d_nolp_misalign_func).
 *
 * rewards/adv/ret: [total] fp32; values: [total + batch] fp32 (sequence i's values start at
 * cu_seqlens[i] + i); cu_seqlens: [batch+1] int32; bootstrap: [batch] bytes (0/1).
 *   delta_t = r_t + gamma * v_{t+1} * (t is last ? bootstrap : 1) - v_t
 *   adv_t   = delta_t + gamma*lam * adv_{t+1};   ret_t = adv_t + v_t.
 * mode: B200RL_GAE_AUTO picks SERIAL when gamma*lam is so close to 1 that re-associating the
 * fp32 additions would move results by more than the 1e-5 parity bar, else SCAN.
 *   SERIAL — the recurrence is evaluated in the reference's order (bit-identical to the reference
 *            CUDA kernel) with warp-coalesced loads/stores; by batch shape one of: a warp per
 *            sequence (chain in one lane, or relayed lane to lane in registers), a two-warp team
 *            per sequence (few long sequences), 8 ticket-fed chains per warp or 8/32 sequences per
 *            warp in lock step (very many short sequences).
 *   SCAN   — segmented reverse affine scan over fixed tiles of the packed stream with a
 *            decoupled look-back across tiles (long sequences spread over many warps).
 * SCAN needs `workspace` of b200rl_gae_workspace_bytes(); it is cleared by the call itself.
 */
#define B200RL_GAE_AUTO 0
#define B200RL_GAE_SERIAL 1
#define B200RL_GAE_SCAN 2
/* SERIAL picks among six kernels by batch shape (same bits from all of them); these force one: */
#define B200RL_GAE_SERIAL_WARP 3    /* one warp per sequence */
#define B200RL_GAE_SERIAL_BATCH8 4  /* 8 sequences per warp, one chain per lane */
#define B200RL_GAE_SERIAL_BATCH32 5 /* 32 sequences per warp */
#define B200RL_GAE_SERIAL_RELAY 7   /* one warp per sequence, chain relayed lane to lane in registers */
#define B200RL_GAE_SERIAL_TEAM 8    /* two warps per sequence: one feeds and drains a ring of blocks, one
                                       lane of the other runs the chain (few, long sequences) */
#define B200RL_GAE_SERIAL_SLOTS 6   /* 8 independent sequence slots per warp, refilled from a ticket
                                       counter in `workspace` (static round-robin without one) */
size_t b200rl_gae_workspace_bytes(int32_t batch, int64_t total);
int b200rl_gae_1d_nolp_misalign(const float* rewards, const float* values,
                                const int32_t* cu_seqlens, const uint8_t* bootstrap, float* adv,
                                float* ret, int32_t batch, int64_t total, float gamma, float lam,
                                int mode, void* workspace, size_t workspace_bytes, void* stream);

/* Same computation as b200rl_gae_1d_nolp_misalign in SCAN mode, for callers that cannot give the
 * library a workspace: e.g. a trainer colocated with an inference server that has already claimed
 * nearly all of HBM, or a bare C / FFI host with no device allocator of its own. This entry point
 * has no counterpart in the reference. It performs no device allocation of any kind: the tile
 * look-back state lives in a fixed array inside the library image (about 33 KB, resident on every
 * GPU that loads the library until the process exits). The caller still owns all inputs and outputs.
 *  - Limit: total <= 4,194,304 packed tokens; larger inputs are refused with B200RL_EINVAL and a
 *    message pointing at the workspace variant.
 *  - The first call on each device looks up the array's device address (one-time set-up with that
 *    device current). Make that first call before starting any stream capture.
 *  - All calls on a device share the one array and each call clears the part it uses, so two calls
 *    that may overlap on the GPU (different streams, or different threads) race with each other:
 *    serialise them on one stream.
 *  - Numerics are those of SCAN (see above): fp32 additions are re-associated, so for
 *    gamma*lam > 0.99 results can leave the 1e-5 band around the reference's serial order. */
int b200rl_gae_1d_nolp_misalign_noalloc(const float* rewards, const float* values,
                                        const int32_t* cu_seqlens, const uint8_t* bootstrap,
                                        float* adv, float* ret, int32_t batch, int64_t total,
                                        float gamma, float lam, void* stream);

/* Padded [B, H] GAE with episode boundaries inside a row.
 * ref: csrc/cugae/gae.cu:62-131 (gae_2d_olp); gae_2d_nolp (:133-206) launches the same kernel
 * (:199), so both bindings map here. rewards [B,H], values / done [B,H+1] contiguous,
 * done_y_indices [num_dones] int32 (time indices of the dones, row after row), cu_num_dones
 * [B+1] int32, is_truncate [num_dones] bool. Segment s of row b covers
 * [s ? y[cu[b]+s-1] : 0, s == last ? H : y[cu[b]+s]); the next value of a segment's final step
 * is 0 when done[b, t+1]; a non-final segment with is_truncate gets lastgae = 0 at its final
 * step. adv / ret must be zero-filled by the caller (as at::zeros_like, gae.cu:117-118) if the
 * segments may leave steps uncovered. Bit-identical to the reference kernel. */
int b200rl_gae_2d_olp(const float* rewards, const float* values, const uint8_t* done,
                      const int32_t* done_y_indices, const int32_t* cu_num_dones,
                      const uint8_t* is_truncate, float* adv, float* ret, int32_t batch,
                      int32_t horizon, int32_t max_num_dones, float gamma, float lam, void* stream);

/* ------------------------------------------------------------------------------------------
 * H2  clipped PPO / GRPO actor loss with decoupled-PPO behaviour weights.
 * ref: areal/utils/functional.py:124-188 (ppo_actor_loss_fn);
 *      twin realhf/impl/model/utils/ppo_functional.py:51-149 (actor_loss_fn).
 *
 * All per-token arrays have T elements. `mask` is bytes (bool). `clip_lo`/`clip_hi` are the
 * clamp bounds 1-eps_clip and 1+(eps_clip_higher or eps_clip), computed by the caller in double
 * and rounded to fp32 exactly as torch.clamp receives them (functional.py:146-150).
 * `c_clip` <= 0 means None, `cap` < 0 means None; `cap` == -2 switches the behaviour weight off
 * altogether (w = 1 on every token: the legacy actor_loss_fn called without proximal_logprobs).
 * Optional fused advantage normalisation (north_star kernel 3): when `adv_mean` or `adv_denom`
 * is non-NULL the kernel uses
 *     A_t = float( ((double)adv[t]*mask_f[t] - adv_mean[g]) / adv_denom[g] ),  g = adv_group[t]
 * (g = 0 when adv_group is NULL; a NULL mean is 0, a NULL denom is 1) — the arithmetic of
 * masked_normalization (ref: areal/utils/functional.py:83-121) — and, when `adv_out` is
 * non-NULL, also writes the normalised advantage there.
 * Outputs: the eight stat arrays of the reference (`stat_*`, any may be NULL), `loss[0]` (fp32
 * scalar = sum(mask*pg)/N), `denom[0]` (fp32, N = max(count(mask),1)), and, if non-NULL,
 * `dlogp[t]` = d(N*loss)/d logp[t] (NOT yet divided by N) for the backward.
 * `workspace`: b200rl_ppo_loss_workspace_bytes(T) bytes; cleared by the call itself.
 * The reduction order is fixed (deterministic loss bits for a given T).
 */
size_t b200rl_ppo_loss_workspace_bytes(int64_t T);
int b200rl_ppo_loss_fwd(const float* logp, const float* prox_logp, const float* old_logp,
                        const float* adv, const uint8_t* mask, int64_t T, float clip_lo,
                        float clip_hi, float c_clip, float cap, const double* adv_mean,
                        const double* adv_denom, const int32_t* adv_group, float* adv_out,
                        float* stat_loss, float* stat_imp_weight, float* stat_approx_kl,
                        uint8_t* stat_clip_mask, uint8_t* stat_dual_clip_mask,
                        float* stat_behave_imp_weight, float* stat_behave_approx_kl,
                        uint8_t* stat_behave_mask, float* dlogp, float* loss, float* denom,
                        void* workspace, size_t workspace_bytes, void* stream);

/* backward of the scalar loss: g_logp[t] = grad_loss[0] * dlogp[t] / denom[0]; grad_loss and
 * denom are DEVICE scalars, so there is no host sync (the reference syncs at
 * `count_nonzero() or 1`, functional.py:143). ref: autograd of functional.py:143-175. */
int b200rl_ppo_loss_bwd(const float* dlogp, const float* grad_loss, const float* denom,
                        float* g_logp, int64_t T, void* stream);

/* ------------------------------------------------------------------------------------------
 * Companions of the GAE path in the legacy PPO interface.
 * ref: realhf/impl/model/utils/ppo_functional.py:150-225 (critic_loss_fn, _mse_loss, _huber_loss),
 *      :228-253 (get_packed_rewards).
 *
 * critic loss: vl = max(f(value, target), f(old + clamp(value-old, -eps, eps), target)), f = mse
 * (loss_type 0) or huber(huber_delta) (loss_type 1); loss = sum(mask*vl)/count(mask) (no `or 1`,
 * as in the reference) or mean(vl) when mask is NULL; stat loss/clip_mask; dvalue = d(N*loss)/d value.
 * The backward of the scalar is b200rl_ppo_loss_bwd (g * dvalue / denom).
 * packed rewards: kl = tot = -kl_ctl*(log_probs - ref_log_probs); tot[cu[i+1]-1] += clip(score_i)
 * (0 instead when mask_no_eos_with_zero and seq_no_eos_mask[i]).
 */
size_t b200rl_critic_loss_workspace_bytes(int64_t T);
int b200rl_critic_loss_fwd(const float* value, const float* old_value, const float* target_value,
                           const uint8_t* mask, int64_t T, float value_eps_clip, int loss_type,
                           float huber_delta, float* stat_loss, uint8_t* stat_clip_mask,
                           float* dvalue, float* loss, float* denom, void* workspace,
                           size_t workspace_bytes, void* stream);
int b200rl_packed_rewards(const float* log_probs, const float* ref_log_probs,
                          const float* reward_score, const int32_t* short1cu_seqlens,
                          const uint8_t* seq_no_eos_mask, int32_t batch, int64_t T, float kl_ctl,
                          float clip_reward_value, int mask_no_eos_with_zero, float* kl_rewards,
                          float* tot_rewards, void* stream);

/* ------------------------------------------------------------------------------------------
 * Packed SFT loss tail (the SFT twin of grpo_loss_fn).
 * ref: areal/engine/sft/lm_engine.py:45-61 (compute_packed_sft_loss).
 * loss = -sum(mask ? logp : 0) / count(mask) (no `or 1`); masked_logp[t] = mask ? logp : 0;
 * dlogp[t] = mask ? -1 : 0 (= d(N*loss)/d logp; backward is b200rl_ppo_loss_bwd); per sequence
 * i over [cu[i], cu[i+1]): seqlogp[i] = (double)(float mean of the valid log-probs), nan when the
 * sequence has none; ppl[i] = exp(-seqlogp[i]). cu_seqlens is int32, or int64 when cu_is_int64.
 * masked_logp, dlogp, seqlogp, ppl are nullable.
 */
size_t b200rl_sft_loss_workspace_bytes(int64_t T, int64_t batch);
int b200rl_sft_loss_fwd(const float* logp, const uint8_t* mask, const void* cu_seqlens,
                        int cu_is_int64, int64_t T, int32_t batch, float* masked_logp, float* dlogp,
                        double* seqlogp, float* ppl, float* loss, float* denom, void* workspace,
                        size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * H2  masked moments + normalisation (advantage normalisation, batch / group level).
 * ref: areal/utils/functional.py:83-121 (masked_normalization),
 *      areal/engine/ppo/actor.py:412-604 (AdvNorm.__call__, _compute_mean, _compute_std).
 *
 * x: [n_elems] fp32; moment group g owns the contiguous slice [g*group_elems, (g+1)*group_elems)
 * (n_elems must be a multiple of group_elems). mask: same shape, fp32 weights, or NULL (ones).
 *   moments[g] = { sum(mask), sum(x*mask), sum(((x - center[i / center_group_elems])*mask)^2) }
 * in fp64 (center == NULL -> 0: the raw second moment). Fixed summation order.
 * `workspace`: b200rl_moments_workspace_bytes(n_elems, group_elems); cleared by the call.
 * finalize: mean = m1/m0; var = centered ? m2/m0 : m2/m0 - mean^2; unbiased: var *= m0/(m0-1);
 *           denom = sqrt(var) + eps.   (functional.py:116-121, actor.py:600-604)
 * normalize: out[i] = float( ((double)x[i]*(mask_x ? mask[i] : 1) - mean[i/mean_group_elems])
 *                            / denom[i/denom_group_elems] );  NULL mean = 0, NULL denom = 1.
 */
size_t b200rl_moments_workspace_bytes(int64_t n_elems, int64_t group_elems);
int b200rl_masked_moments(const float* x, const float* mask, int64_t n_elems,
                          int64_t group_elems, const double* center, int64_t center_group_elems,
                          double* moments, void* workspace, size_t workspace_bytes, void* stream);
int b200rl_moments_finalize(const double* moments, int64_t n_groups, int centered, int unbiased,
                            double eps, double* mean, double* denom, void* stream);
int b200rl_masked_normalize(const float* x, const float* mask, int64_t n_elems,
                            const double* mean, int64_t mean_group_elems, const double* denom,
                            int64_t denom_group_elems, int mask_x, float* out, void* stream);
/* The three calls above back to back for the common case of one process and the same level for
 * mean and std (masked_normalization with all_reduce off, AdvNorm "group"/"group"):
 * moments (raw second moment) -> finalize -> out = (x*mask - mean[g]) / denom[g], g = i / group_elems.
 * `moments` [G,3], `mean` [G], `denom` [G] are caller-owned fp64 outputs. One host call, three
 * launches; same bits as the separate calls. */
int b200rl_masked_normalize_groups(const float* x, const float* mask, int64_t n_elems,
                                   int64_t group_elems, int unbiased, double eps, float* out,
                                   double* moments, double* mean, double* denom, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * H4  interval gather / scatter on a flat parameter buffer.
 * ref: csrc/interval_op/interval_op.cu:12-114 (copyDataKernel, slice_intervals, set_intervals),
 *      realhf/impl/model/nn/flatten_param.py:65-186 (python wrappers and fallbacks),
 *      csrc/interval_op/interval_op.cpp:4-25 (merge_intervals).
 *
 * intervals: DEVICE int64 [n, 2] of [start, end) element offsets into the big buffer.
 * slice: packed[off_i : off_i + len_i] = big[start_i : end_i]    (off = exclusive prefix sum of len)
 * set:   big[start_i : end_i] = packed[off_i : off_i + len_i]
 * Pure byte moves, bit-exact for any payload. elem_size in {1,2,4,8}.
 * `packed_numel` is the packed side's element count (slice: output size; set: src size).
 * workspace: b200rl_interval_workspace_bytes(n, packed_numel, elem_size) bytes. The plan
 * (prefix sums + tile->interval map) is rebuilt on every call unless `reuse_plan` != 0, in which
 * case the workspace must hold the plan of a previous call with identical intervals.
 * After the call workspace's first int64 holds the sum of interval lengths and the second a
 * non-zero flag if an interval was malformed or out of bounds (nothing is copied for those).
 */
size_t b200rl_interval_workspace_bytes(int64_t n_intervals, int64_t packed_numel, int elem_size);
int b200rl_slice_intervals(const void* big, int64_t big_numel, void* packed, int64_t packed_numel,
                           const int64_t* intervals, int64_t n_intervals, int elem_size,
                           void* workspace, size_t workspace_bytes, int reuse_plan, void* stream);
int b200rl_set_intervals(const void* packed, int64_t packed_numel, void* big, int64_t big_numel,
                         const int64_t* intervals, int64_t n_intervals, int elem_size,
                         void* workspace, size_t workspace_bytes, int reuse_plan, void* stream);

/* Fused slice -> set without the packed staging buffer (the kernel asked for
 * by the TODO at realhf/impl/model/nn/real_llm_api.py:719-720, and with `src` pointing at a peer
 * GPU's flat parameter buffer the replacement of slice_intervals -> chunked_bcast ->
 * set_intervals, real_llm_api.py:721-805). Equivalent to
 *   set_intervals(slice_intervals(src, src_intervals), dst, dst_intervals)
 * where both interval lists sum to total_numel elements but may cut the stream differently.
 * Intervals, workspace and dst live on the launching device; src may be peer memory that the
 * launching device can access (b200rl_enable_peer_access). The kernel never waits on another
 * GPU: ordering against the owner of `src` is the caller's (events / barrier). */
size_t b200rl_move_workspace_bytes(int64_t n_src, int64_t n_dst, int64_t total_numel, int elem_size);
int b200rl_move_intervals(const void* src, int64_t src_numel, void* dst, int64_t dst_numel,
                          const int64_t* src_intervals, int64_t n_src, const int64_t* dst_intervals,
                          int64_t n_dst, int64_t total_numel, int elem_size, void* workspace,
                          size_t workspace_bytes, int reuse_plan, void* stream);
/* cudaDeviceEnablePeerAccess(device -> peer), idempotent; B200RL_EUNSUPPORTED when impossible. */
int b200rl_enable_peer_access(int device, int peer);
/* CUDA IPC plumbing for the peer-memory use of b200rl_move_intervals (no reference counterpart:
 * the reference stages through NCCL). export: 64-byte handle of the allocation containing `ptr`
 * and ptr's byte offset inside it. open: map that allocation with `device` (the CONSUMER's GPU)
 * current, peer access enabled lazily; returns the mapped base. close: unmap. A handle can be
 * opened once per process and never by the exporting process. */
int b200rl_ipc_export(const void* ptr, unsigned char* handle64, int64_t* offset_bytes);
int b200rl_ipc_open(const unsigned char* handle64, int device, void** base_out);
int b200rl_ipc_close(void* base, int device);

/* host: merge touching neighbours [a,b),[b,c) -> [a,c), order preserving, no sort.
 * in/out: n (resp. return value) pairs of uint64. out may alias in. */
size_t b200rl_merge_intervals(const uint64_t* in, size_t n, uint64_t* out);

#ifdef __cplusplus
}
#endif
#endif /* B200RL_H_ */
