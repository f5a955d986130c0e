// Per-token arithmetic of the clipped PPO / GRPO actor loss with decoupled-PPO behaviour weights,
// shared by ppo_loss.cu and grpo_fused.cu so both produce the same bits.
// ref: areal/utils/functional.py:143-187; gradient semantics of torch.max / min / clamp
// (ties split 1/2 + 1/2, clamp passes the gradient on its inclusive bounds).
#pragma once
#include "common.cuh"

namespace b200rl {

// torch.maximum / minimum / clamp propagate NaN; fmaxf/fminf do not.
__device__ __forceinline__ float tmax(float a, float b) {
  return (a != a || b != b) ? __int_as_float(0x7fc00000) : fmaxf(a, b);
}
__device__ __forceinline__ float tmin(float a, float b) {
  return (a != a || b != b) ? __int_as_float(0x7fc00000) : fminf(a, b);
}

struct PpoToken {
  float pgw;    // per-token loss (stat "loss"): pg * behaviour weight
  float ratio;  // importance_weight
  float kl;     // approx_kl = logp - prox (not masked)
  float w;      // behave_imp_weight
  float bkl;    // behave_approx_kl
  float dlogp;  // d(N * loss) / d logp
  bool clip, dual, bmask;
};

// c_clip <= 0: no dual clip. cap < 0: no cap; cap < -1.5: no behaviour weighting at all;
// cap < -2.5: not PPO at all but the SFT negative log-likelihood.
__device__ __forceinline__ PpoToken ppo_token(float lp, float px, float ol, float A, bool m,
                                              float clip_lo, float clip_hi, float c_clip,
                                              float cap) {
  PpoToken o;
  if (cap < -2.5f) {
    // negative log-likelihood (packed SFT loss, areal/engine/sft/lm_engine.py:51-53):
    // per-token loss -logp on the masked tokens; prox / old / advantages are ignored
    o.pgw = m ? -lp : 0.f;
    o.ratio = 0.f;
    o.kl = 0.f;
    o.w = 0.f;
    o.bkl = 0.f;
    o.dlogp = m ? -1.f : 0.f;
    o.clip = o.dual = false;
    o.bmask = m;
    return o;
  }
  const bool has_c = c_clip > 0.f;
  const bool has_cap = cap >= 0.f;
  const bool no_bw = cap < -1.5f;
  // functional.py:144
  const float d = lp - px;
  const float r = m ? expf(d) : 0.f;
  // functional.py:146-150  clamp(r, lo, hi) = min(max(r, lo), hi)
  const float rc = tmin(tmax(r, clip_lo), clip_hi);
  // functional.py:152-155
  const float nA = -A;
  const float l1 = nA * r, l2 = nA * rc;
  o.clip = (l1 < l2) && m;
  float pg = tmax(l1, l2);
  const float dl1 = m ? l1 : 0.f;  // d l1 / d lp = -A * r
  const float dl2 = (r >= clip_lo && r <= clip_hi) ? dl1 : 0.f;
  float gp = (l1 > l2) ? dl1 : ((l1 < l2) ? dl2 : 0.5f * (dl1 + dl2));
  o.dual = false;
  if (has_c) {  // functional.py:156-160
    const float sg = (A > 0.f) ? 1.f : ((A < 0.f) ? -1.f : ((A == 0.f) ? 0.f : A));
    const float l3 = (sg * c_clip) * A;
    o.dual = (l3 < pg) && m;
    gp = (pg < l3) ? gp : ((pg == l3) ? 0.5f * gp : 0.f);
    pg = tmin(pg, l3);
  }
  // functional.py:163-172
  float bk = px - ol;
  float w = expf(bk);
  const bool bm = has_cap ? ((w <= cap) && m) : m;
  bk = bm ? bk : 0.f;
  w = bm ? w : 0.f;
  if (no_bw) w = 1.f;  // legacy actor_loss_fn without proximal_logprobs: no weighting at all
  o.pgw = pg * w;
  o.ratio = r;
  o.kl = d;
  o.w = w;
  o.bkl = bk;
  o.bmask = bm;
  o.dlogp = m ? gp * w : 0.f;
  return o;
}

}  // namespace b200rl
