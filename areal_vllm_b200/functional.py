"""Drop-in for ``areal.utils.functional`` (reference: areal/utils/functional.py).

Same names, positional order and defaults as the reference, so the call sites
(``areal/engine/ppo/actor.py:14-21``, ``areal/engine/sft/lm_engine.py:9``) only change their
import. The math runs in the sm_100a kernels behind ``areal_vllm_b200.ops``; tensors must be on
CUDA — there is no CPU path.
"""

from __future__ import annotations

import warnings
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from . import ops

__all__ = [
    "gather_logprobs",
    "gather_logprobs_entropy",
    "gather_logprobs_entropy_minmax",
    "masked_normalization",
    "ppo_actor_loss_fn",
    "dynamic_sampling",
    "reward_overlong_penalty",
]


def _as_2d(logits: torch.Tensor, labels: torch.Tensor):
    """The reference works on any leading shape (log_softmax over the last dim); the engines
    hand it packed [T, V]. Flatten leading dims (a view for contiguous inputs)."""
    lead = logits.shape[:-1]
    if labels.shape != lead:
        raise RuntimeError(f"labels shape {tuple(labels.shape)} != logits leading shape {tuple(lead)}")
    l2 = logits if logits.dim() == 2 else logits.reshape(-1, logits.shape[-1])
    if l2.stride(-1) != 1:
        l2 = l2.contiguous()
    y = labels.reshape(-1)
    if y.dtype != torch.int64:
        y = y.long()
    return l2, y.contiguous(), lead


def gather_logprobs(
    logits: torch.Tensor,
    labels: torch.Tensor,
    temperature: float = 1.0,
    chunk_size: int = 1024,
):
    """log_softmax(logits.float() / temperature)[labels]  (ref functional.py:10-16, 29-51).

    ``chunk_size`` is accepted for signature compatibility and ignored: the kernel streams every
    row once and never materialises [chunk, V] fp32 temporaries, so there is nothing to chunk.
    """
    l2, y, lead = _as_2d(logits, labels)
    logp, _, _, _, _ = ops.logprob_entropy_fwd(l2, y, float(temperature), False, False)
    return logp.view(lead)


def gather_logprobs_entropy(
    logits: torch.Tensor,
    labels: torch.Tensor,
    temperature: float = 1.0,
    chunk_size: int = 1024,
):
    """(log-prob of the label, entropy of the row), both differentiable (ref functional.py:19-26, 54-80)."""
    l2, y, lead = _as_2d(logits, labels)
    logp, ent, _, _, _ = ops.logprob_entropy_fwd(l2, y, float(temperature), True, False)
    return logp.view(lead), ent.view(lead)


def gather_logprobs_entropy_minmax(
    logits: torch.Tensor, labels: torch.Tensor, temperature: float = 1.0
):
    """Extension: also returns the raw per-row max / min logits from the same single pass
    (``grpo_loss_fn`` otherwise re-reads [T, V] twice for them, ref areal/engine/ppo/actor.py:351-352)."""
    l2, y, lead = _as_2d(logits, labels)
    logp, ent, _, rmax, rmin = ops.logprob_entropy_fwd(l2, y, float(temperature), True, True)
    return logp.view(lead), ent.view(lead), rmax.view(lead), rmin.view(lead)


@torch.no_grad()
def masked_normalization(
    x: torch.Tensor,
    mask: Optional[torch.Tensor] = None,
    dim=None,
    unbiased=False,
    eps=1e-5,
    high_precision=True,
    all_reduce=True,
    reduce_group=None,
):
    """((x*mask - mean) / (sqrt(var) + eps)).float() with masked moments over ``dim``
    (ref functional.py:83-121).

    Moments are always accumulated in fp64 (the reference's default ``high_precision=True``;
    with ``False`` the reference uses fp32 sums, which this differs from by rounding only).
    When torch.distributed is initialised and ``all_reduce`` is set, the three moments are
    summed over ``reduce_group`` with ONE all-reduce of a [G, 3] fp64 tensor instead of the
    reference's three.
    """
    if not x.is_cuda:
        raise RuntimeError("x must be on CUDA (b200rl has no CPU path)")
    nd = x.dim()
    if dim is None:
        dims = tuple(range(nd))
    elif isinstance(dim, int):
        dims = (dim % nd,)
    else:
        dims = tuple(sorted(d % nd for d in dim))
    keep = tuple(d for d in range(nd) if d not in dims)
    perm = keep + dims
    xf = x.float()
    mf = None
    if mask is not None:
        mf = mask.to(torch.float32).expand_as(xf)
    needs_perm = perm != tuple(range(nd))
    if needs_perm:
        xf = xf.permute(perm)
        mf = mf.permute(perm) if mf is not None else None
    xf = xf.contiguous()
    mf = mf.contiguous() if mf is not None else None
    group_elems = 1
    for d in dims:
        group_elems *= x.shape[d]
    if xf.numel() == 0 or group_elems == 0:
        # an empty shard still has to join the collective (the reference all-reduces zero moments,
        # functional.py:108-115): the other ranks of reduce_group are waiting in it
        if dist.is_initialized() and all_reduce:
            G = 1
            for d in keep:
                G *= x.shape[d]
            if G > 0:
                dist.all_reduce(torch.zeros(G, 3, dtype=torch.float64, device=x.device),
                                op=dist.ReduceOp.SUM, group=reduce_group)
        return x.float()
    reduce = dist.is_initialized() and all_reduce
    if not reduce and not torch.compiler.is_compiling():
        # no collective in the middle: moments -> finalize -> normalise in one host call
        out = ops._raw_masked_normalize_groups(
            xf.view(-1), None if mf is None else mf.view(-1), group_elems, bool(unbiased), float(eps)
        ).view(xf.shape)
    else:
        mom = ops.masked_moments(xf.view(-1), None if mf is None else mf.view(-1), group_elems, None, 1)
        if reduce:
            dist.all_reduce(mom, op=dist.ReduceOp.SUM, group=reduce_group)
        mean, denom = ops.moments_finalize(mom, False, bool(unbiased), float(eps))
        out = ops.masked_normalize(
            xf.view(-1), None if mf is None else mf.view(-1), mean, group_elems, denom, group_elems, True
        ).view(xf.shape)
    if needs_perm:
        inv = [0] * nd
        for i, p in enumerate(perm):
            inv[p] = i
        out = out.permute(inv).contiguous()
    return out


def ppo_actor_loss_fn(
    logprobs: torch.Tensor,
    proximal_logprobs: torch.Tensor,
    old_logprobs: torch.Tensor,
    advantages: torch.Tensor,
    eps_clip: float,
    loss_mask: torch.Tensor,
    eps_clip_higher: Optional[float] = None,
    c_clip: Optional[float] = None,
    behav_imp_weight_cap: Optional[float] = None,
) -> Tuple[torch.Tensor, Dict]:
    """Clipped PPO actor loss with decoupled-PPO behaviour weights (ref functional.py:124-188).

    One fused kernel instead of ~25 eager ops and a host sync. Returns ``(loss, stat)`` with the
    reference's stat keys, dtypes (fp32 / bool) and shapes. Gradient flows to ``logprobs`` only
    (in the reference the other inputs never require grad at the call sites).
    """
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    out = ops.ppo_loss_fwd(
        logprobs,
        proximal_logprobs,
        old_logprobs,
        advantages,
        loss_mask,
        1.0 - eps_clip,
        1.0 + (eps_clip if eps_clip_higher is None else eps_clip_higher),
        -1.0 if c_clip is None else float(c_clip),
        -1.0 if behav_imp_weight_cap is None else float(behav_imp_weight_cap),
        None,
        None,
        None,
    )
    loss = out[0]
    # the reference's stats are detached; ours come back non-differentiable already (eager path),
    # so only tensors that still carry a graph (compiled custom-op path) are detached
    det = [t.detach() if t.requires_grad else t for t in out]
    stat = dict(
        loss=det[4],
        importance_weight=det[5],
        approx_kl=det[6],
        clip_mask=det[7].view(torch.bool),
        dual_clip_mask=det[8].view(torch.bool),
    )
    if proximal_logprobs is not None:
        stat["behave_imp_weight"] = det[9]
        stat["behave_approx_kl"] = det[10]
        stat["behave_mask"] = det[11].view(torch.bool)
    return loss, stat


def dynamic_sampling(data, group_size: int):
    """DAPO-style filter: drop groups whose rewards are all equal (ref functional.py:191-233).

    TensorDict plumbing, not a kernel; kept because the actor imports it from this module.
    Reference quirk preserved: the two early exits return ``data`` alone, the normal path
    returns ``(data, stats)``.
    """
    rewards = data["rewards"]
    n = rewards.shape[0]
    if n % group_size != 0:
        warnings.warn(
            "The group size is not divisible by the batch size. Return the original data"
        )
        return data
    grouped = rewards.view(n // group_size, group_size)
    keep_group = (grouped != grouped[:, :1]).any(dim=1)
    keep = keep_group.repeat_interleave(group_size)
    if not keep.any():
        return data
    kept = int(keep_group.sum().item())
    return data[keep], dict(n_group_keeped=kept, n_group_filtered=n // group_size - kept)


def reward_overlong_penalty(
    data,
    overlong_tokens: int,
    overlong_penalty_factor: float,
    max_response_length: int,
):
    """DAPO overlong penalty on sequence rewards (ref functional.py:237-259), vectorised:
    reward += min(-(resp_len - (max_len - overlong_tokens)) / overlong_tokens * factor, 0)."""
    rewards = data["rewards"]
    resp_len = data["loss_mask"].sum(dim=-1).long()
    exceed = resp_len - (max_response_length - overlong_tokens)
    penalty = torch.clamp(-exceed / overlong_tokens * overlong_penalty_factor, max=0)
    data["rewards"] = rewards + penalty.to(rewards.dtype)
    return data
