"""fp32 torch (CPU or GPU) restatement of the reference's H1/H2 math. TEST INFRASTRUCTURE.

The reference for these functions *is* a sequence of torch ops, so the oracle is the same
sequence of torch ops written out again, op by op, with the reference line beside each step.
Pinned by tests/golden/h1_*.npz, h2_*.npz (generated from the real reference by gen_golden.py).
"""

from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------
# H1  ref: areal/utils/functional.py:10-26
# ------------------------------------------------------------------------------------------
def logprobs_entropy(
    logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0, rows_per_step: int = 1024
) -> Tuple[torch.Tensor, torch.Tensor]:
    """(logp[T], entropy[T]) through the eager fp32 path; differentiable wrt ``logits``.

    functional.py:14  log_softmax(logits.float() / temperature, dim=-1)
    functional.py:15  gather at labels
    functional.py:24  entropy = -sum(exp(lp) * lp)
    Processed ``rows_per_step`` rows at a time like the reference's chunk loop (:42-51, :69-80).
    """
    lps, ents = [], []
    for lo in range(0, logits.shape[0], rows_per_step):
        blk = logits[lo:lo + rows_per_step]
        lp_all = F.log_softmax(blk.float() / temperature, dim=-1)
        ents.append(-(lp_all.exp() * lp_all).sum(dim=-1))
        lps.append(lp_all.gather(-1, labels[lo:lo + rows_per_step].unsqueeze(-1)).squeeze(-1))
    return torch.cat(lps), torch.cat(ents)


def logprobs_entropy_f64(logits, labels, temperature=1.0):
    """Same quantities in float64 (ground truth for error budgeting; not the parity reference)."""
    x = logits.double() / temperature
    lp_all = F.log_softmax(x, dim=-1)
    ent = -(lp_all.exp() * lp_all).sum(-1)
    return lp_all.gather(-1, labels.unsqueeze(-1)).squeeze(-1), ent


# ------------------------------------------------------------------------------------------
# H2  ref: areal/utils/functional.py:83-121
# ------------------------------------------------------------------------------------------
@torch.no_grad()
def masked_normalization(
    x: torch.Tensor,
    mask: Optional[torch.Tensor] = None,
    dim=None,
    unbiased: bool = False,
    eps: float = 1e-5,
    high_precision: bool = True,
) -> torch.Tensor:
    """Single-process masked_normalization (no all-reduce)."""
    wide = torch.float64 if high_precision else torch.float32          # :94
    xw = x.to(wide)
    dims = tuple(range(xw.dim())) if dim is None else dim                # :96-97
    if mask is None:
        count = torch.tensor(float(math.prod(xw.shape[d] for d in dims)), dtype=wide)  # :98-101
    else:
        mw = mask.to(wide)                                               # :103
        xw = xw * mw                                                     # :104
        count = mw.sum(dims, keepdim=True)                               # :105
    s1 = xw.sum(dim=dims, keepdim=True)                                  # :106
    s2 = (xw * xw).sum(dim=dims, keepdim=True)                           # :107
    mu = s1 / count                                                      # :116
    var = s2 / count - mu * mu                                           # :117-118
    if unbiased:
        var = var * (count / (count - 1))                                # :119-120
    return ((xw - mu) / (var.sqrt() + eps)).float()                      # :121


# ------------------------------------------------------------------------------------------
# H2  ref: areal/utils/functional.py:124-188
# ------------------------------------------------------------------------------------------
def ppo_actor_loss(
    logprobs: torch.Tensor,
    proximal_logprobs: torch.Tensor,
    old_logprobs: torch.Tensor,
    advantages: torch.Tensor,
    eps_clip: float,
    loss_mask: torch.Tensor,
    eps_clip_higher: Optional[float] = None,
    c_clip: Optional[float] = None,
    behav_imp_weight_cap: Optional[float] = None,
) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
    n_valid = loss_mask.count_nonzero() or 1                                       # :143
    log_ratio = logprobs - proximal_logprobs
    ratio = torch.where(loss_mask, log_ratio.exp(), 0)                             # :144
    upper = 1.0 + (eps_clip_higher if eps_clip_higher is not None else eps_clip)
    ratio_c = ratio.clamp(1.0 - eps_clip, upper)                                   # :146-150
    surr1 = -advantages * ratio                                                    # :152
    surr2 = -advantages * ratio_c                                                  # :153
    clipped = surr1.detach() < surr2.detach()                                      # :154
    pg = torch.max(surr1, surr2)                                                   # :155
    if c_clip is not None:
        assert c_clip > 1.0, c_clip                                                # :157
        surr3 = torch.sign(advantages) * c_clip * advantages                       # :158
        dual = surr3.detach() < pg.detach()                                        # :159
        pg = torch.min(pg, surr3)                                                  # :160
    else:
        dual = torch.zeros_like(clipped)                                           # :162
    b_logr = proximal_logprobs - old_logprobs                                      # :163
    b_w = b_logr.exp()                                                             # :164
    if behav_imp_weight_cap is not None:
        b_mask = (b_w <= behav_imp_weight_cap) & loss_mask                         # :165-167
    else:
        b_mask = loss_mask                                                         # :168-169
    b_logr = torch.where(b_mask, b_logr, 0.0)                                      # :170
    b_w = torch.where(b_mask, b_w, 0.0)                                            # :171
    pg = pg * b_w                                                                  # :172
    per_token = pg.detach()                                                        # :173
    loss = torch.where(loss_mask, pg, 0).sum() / n_valid                           # :174
    stat = dict(
        loss=per_token,
        importance_weight=ratio.detach(),
        approx_kl=log_ratio.detach(),
        clip_mask=clipped & loss_mask,                                             # :175
        dual_clip_mask=dual & loss_mask,                                           # :176
        behave_imp_weight=b_w,
        behave_approx_kl=b_logr,
        behave_mask=b_mask,
    )
    return loss, stat


# ------------------------------------------------------------------------------------------
# AdvNorm  ref: areal/engine/ppo/actor.py:412-604
# ------------------------------------------------------------------------------------------
def _level_mean(x, mask):
    """actor.py:538-566 (_compute_mean, fp64, no all-reduce)."""
    xw = x.double()
    if mask is None:
        return (xw.sum() / xw.numel()).reshape([1] * xw.dim())
    mw = mask.double()
    return ((xw * mw).sum() / mw.sum()).reshape([1] * xw.dim())


def _level_std(x, mask, mean, unbiased):
    """actor.py:568-604 (_compute_std: centred sum of squares, fp64)."""
    xw = x.double()
    if mask is None:
        cnt = torch.tensor(float(xw.numel()), dtype=torch.float64)
        ssq = ((xw - mean) ** 2).sum()
    else:
        mw = mask.double()
        cnt = mw.sum()
        ssq = ((xw * mw - mean * mw) ** 2).sum()
    var = ssq / cnt
    if unbiased:
        var = var * (cnt / (cnt - 1))
    return var.sqrt().reshape([1] * xw.dim())


@torch.no_grad()
def adv_norm(
    advantages: torch.Tensor,
    loss_mask: Optional[torch.Tensor],
    mean_level: str,
    std_level: str,
    group_size: Optional[int] = None,
    eps: float = 1e-5,
    unbiased: bool = False,
) -> torch.Tensor:
    bs = advantages.size(0)
    groups = (
        [slice(g * group_size, (g + 1) * group_size) for g in range(bs // group_size)]
        if group_size
        else []
    )
    if mean_level == std_level:                                           # actor.py:440-470
        if mean_level == "batch":
            return masked_normalization(advantages, loss_mask, unbiased=unbiased, eps=eps)
        if mean_level == "none":
            return advantages.float()
        return torch.cat(
            [
                masked_normalization(
                    advantages[s], None if loss_mask is None else loss_mask[s],
                    unbiased=unbiased, eps=eps,
                )
                for s in groups
            ],
            0,
        )
    if mean_level == "batch":                                             # actor.py:478-494
        mean = _level_mean(advantages, loss_mask).expand_as(advantages)
    elif mean_level == "group":
        mean = torch.zeros_like(advantages)
        for s in groups:
            mean[s] = _level_mean(advantages[s], None if loss_mask is None else loss_mask[s]).expand_as(advantages[s])
    else:
        mean = torch.zeros_like(advantages)
    centred = advantages - mean                                           # :497
    if std_level == "none":
        return centred.float()                                            # :500-501
    if std_level == "batch":                                              # :503-514
        std = _level_std(advantages, loss_mask, mean, unbiased).expand_as(advantages)
    else:                                                                 # :515-533
        std = torch.zeros_like(advantages)
        for s in groups:
            std[s] = _level_std(
                advantages[s], None if loss_mask is None else loss_mask[s], mean[s], unbiased
            ).expand_as(advantages[s])
    return (centred / (std + eps)).float()                                # :536


@torch.no_grad()
def group_reward_norm(reward_score: torch.Tensor, group_size: int) -> torch.Tensor:
    """actor.py:94-98."""
    out = reward_score.clone()
    for g in range(reward_score.shape[0] // group_size):
        s = slice(g * group_size, (g + 1) * group_size)
        r = reward_score[s]
        out[s] = (r - r.mean()) / (r.std() + 1e-9)
    return out


@torch.no_grad()
def padded_gae(rewards, values, seq_no_eos_mask, gamma, lam):
    """The padded GAE loop of PPOActor.compute_advantages (areal/engine/ppo/actor.py:131-148),
    on a copy of ``values`` (the reference's ``nextvalues *= mask`` writes through a view)."""
    values = values.clone()
    bs, max_seqlen = rewards.shape
    rev = [torch.zeros(bs, dtype=torch.float32, device=values.device)]          # :136-138
    last = 0
    for t in reversed(range(max_seqlen - 1)):                                   # :140
        nxt = values[:, t + 1]                                                  # :141
        if t == max_seqlen - 2:
            nxt *= seq_no_eos_mask                                              # :142-143
        delta = rewards[:, t] + gamma * nxt - values[:, t]                      # :144
        last = delta + gamma * lam * last                                       # :145
        rev.append(last)                                                        # :146
    return torch.stack(rev[::-1], dim=1)                                        # :147


def critic_loss(value, old_value, target_value, value_eps_clip, loss_mask=None, loss_fn_type="mse"):
    """Restates realhf/impl/model/utils/ppo_functional.py:150-225 (critic_loss_fn with _mse_loss /
    _huber_loss, delta = 10). Returns (loss, {clip_mask, loss})."""
    delta = 10.0

    def f(x, y):
        if loss_fn_type == "mse":
            return 0.5 * (x - y) ** 2
        if loss_fn_type == "huber":
            diff = torch.abs(x - y)
            return torch.where(diff < delta, 0.5 * diff**2, delta * (diff - 0.5 * delta))
        raise NotImplementedError(loss_fn_type)

    orig = f(value, target_value)
    clipped_value = old_value + (value - old_value).clamp(-value_eps_clip, value_eps_clip)
    clipped = f(clipped_value, target_value)
    vl = torch.max(orig, clipped)
    with torch.no_grad():
        clip_mask = clipped.detach() > orig.detach()
        if loss_mask is not None:
            clip_mask = clip_mask.logical_and(loss_mask)
        stat = dict(clip_mask=clip_mask, loss=vl.detach())
    if loss_mask is not None:
        loss = torch.where(loss_mask, vl, 0).sum() / loss_mask.count_nonzero()
    else:
        loss = vl.mean()
    return loss, stat


def packed_rewards(kl_ctl, clip_reward_value, log_probs, ref_log_probs, reward_score,
                   short1cu_seqlens, seq_no_eos_mask, mask_no_eos_with_zero=False):
    """Restates realhf/impl/model/utils/ppo_functional.py:228-253 (get_packed_rewards)."""
    tot = -kl_ctl * (log_probs - ref_log_probs)
    kl = tot.clone()
    score = reward_score.clip(-clip_reward_value, clip_reward_value)
    idx = (short1cu_seqlens[1:] - 1).long()
    if mask_no_eos_with_zero:
        tot[idx] += torch.where(seq_no_eos_mask, 0, score)
    else:
        tot[idx] += score
    return kl, tot


def packed_sft_loss(logits, input_ids, cu_seqlens, loss_mask):
    """Restates areal/engine/sft/lm_engine.py:45-83 (compute_packed_sft_loss) without the stats
    tracker: returns (loss, dict(seqlogp f64, ppl f32, loss per token, vocab_min/max_logits))."""
    labels = torch.roll(input_ids, shifts=-1, dims=-1)
    lp = torch.nn.functional.log_softmax(logits.float(), dim=-1).gather(-1, labels.unsqueeze(-1)).squeeze(-1)
    m = torch.roll(loss_mask.bool(), shifts=-1, dims=-1)
    lp = torch.where(m, lp, 0)
    loss = -lp.sum() / m.count_nonzero()
    with torch.no_grad():
        B = cu_seqlens.shape[0] - 1
        seqlogp = torch.zeros(B, dtype=torch.float64)
        for i in range(B):
            mi = m[cu_seqlens[i]: cu_seqlens[i + 1]]
            li = lp[cu_seqlens[i]: cu_seqlens[i + 1]]
            seqlogp[i] = torch.where(mi, li.detach(), 0.0).sum() / mi.count_nonzero()
        stat = dict(seqlogp=seqlogp, ppl=(-seqlogp).exp().float(), loss=-lp.detach(),
                    vocab_min_logits=logits.detach().min(-1).values.float(),
                    vocab_max_logits=logits.detach().max(-1).values.float())
    return loss, stat
