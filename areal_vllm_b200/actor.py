"""Drop-in for the hot-path pieces of ``areal/engine/ppo/actor.py``: ``AdvNorm`` (:370-604),
``grpo_loss_fn`` (:292-367), and the group reward normalisation of ``compute_advantages``
(:94-98), plus the fused "advantage normalisation inside the PPO loss" entry point.

Everything numeric goes through the CUDA kernels of ``areal_vllm_b200.ops``.
"""

from __future__ import annotations

import sys
from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from . import functional, ops

__all__ = [
    "grpo_loss_fn_fused",
    "padded_gae",
    "AdvNorm",
    "grpo_loss_fn",
    "group_advantage_stats",
    "ppo_actor_loss_fn_fused_advnorm",
    "group_reward_norm",
]


# grpo_loss_fn: take the two-launch composite (csrc/grpo_loss.cu) in eager mode. Tests set this to
# False to compare it against the custom-op path, which is what a compiled graph uses.
FAST_PATH = True


@torch.no_grad()
def padded_gae(
    rewards: torch.Tensor,
    values: torch.Tensor,
    seq_no_eos_mask: torch.Tensor,
    gamma: float,
    lam: float,
) -> torch.Tensor:
    """The padded ``[B, L]`` GAE of ``PPOActor.compute_advantages`` (ref actor.py:131-148), which
    the reference evaluates with a python loop over ``L-1`` time steps (~5 launches per step):

        for t in reversed(range(L - 1)):
            nextvalues = values[:, t + 1] * (seq_no_eos_mask if t == L - 2 else 1)
            delta = rewards[:, t] + gamma * nextvalues - values[:, t]
            lastgaelam = delta + gamma * lam * lastgaelam

    and returns ``[B, L]`` advantages whose last column is zero. A row of the padded tensors IS a
    packed "misaligned" sequence: ``L-1`` rewards against ``L`` values, so the whole batch is one
    call of the packed GAE kernel with ``cu_seqlens = arange(B+1) * (L-1)`` and
    ``bootstrap = seq_no_eos_mask`` — no new kernel, one launch instead of ``5 * (L-1)``.
    Unlike the reference loop this does not zero ``values[:, -1]`` in place where the mask is 0.
    """
    if rewards.dim() != 2 or rewards.shape != values.shape:
        raise RuntimeError("rewards and values must both be [B, L]")
    B, L = rewards.shape
    out = torch.zeros((B, L), dtype=torch.float32, device=rewards.device)
    if L < 2 or B == 0:
        return out
    cu = torch.arange(B + 1, dtype=torch.int32, device=rewards.device) * (L - 1)
    adv, _ = ops.gae_1d_nolp_misalign(
        rewards[:, : L - 1].float().contiguous().view(-1),
        values.float().contiguous().view(-1),
        cu,
        seq_no_eos_mask.bool().contiguous(),
        float(gamma),
        float(lam),
        0,
    )
    out[:, : L - 1] = adv.view(B, L - 1)
    return out


class AdvNorm:
    """Advantage normalisation with independent levels for the mean and the std
    ("batch" | "group" | "none"), reference areal/engine/ppo/actor.py:370-604.

    * same level  -> masked_normalization semantics (x*mask centred, raw second moment).
      "group": groups are ``group_size`` consecutive ROWS; rows beyond the last complete group
      are dropped from the output, as the reference's torch.cat of complete groups does (:455-470).
    * mixed       -> mean and centred std computed separately (:538-604); x is NOT masked in the
      numerator; rows beyond the last complete group see mean 0 / std 0 (-> x/eps) for a
      group-level statistic, as in the reference (:486-494, :518-533).

    One moments launch per level instead of a python loop over groups; moments in fp64.
    The constructor's argument validation (the three ValueError messages) is the reference's own
    (actor.py:388-412), kept word for word because callers and the reference's tests
    (areal/tests/test_adv_norm_config.py) match on it; ``__call__`` is a different algorithm.
    """

    def __init__(self, advNorm_cfg):
        if advNorm_cfg is None:
            return None
        if advNorm_cfg.mean_level not in {"batch", "group", "none"}:
            raise ValueError(
                f"mean_level must be 'batch', 'group' or 'none', got {advNorm_cfg.mean_level}"
            )
        if advNorm_cfg.std_level not in {"batch", "group", "none"}:
            raise ValueError(
                f"std_level must be 'batch', 'group', or 'none', got {advNorm_cfg.std_level}"
            )
        if (
            advNorm_cfg.mean_level == "group" or advNorm_cfg.std_level == "group"
        ) and advNorm_cfg.group_size is None:
            raise ValueError("group_size must be provided if using group normalization")
        self.mean_level = advNorm_cfg.mean_level
        self.std_level = advNorm_cfg.std_level
        self.group_size = advNorm_cfg.group_size

    @torch.no_grad()
    def __call__(
        self,
        advantages: torch.Tensor,
        loss_mask: Optional[torch.Tensor] = None,
        eps: float = 1e-5,
        unbiased: bool = False,
        high_precision: bool = True,
        reduce_group=None,
    ) -> torch.Tensor:
        bs = advantages.size(0)
        if self.mean_level == self.std_level:
            if self.mean_level == "batch":
                return functional.masked_normalization(
                    advantages, mask=loss_mask, unbiased=unbiased, eps=eps,
                    high_precision=high_precision, all_reduce=True, reduce_group=reduce_group,
                )
            if self.mean_level == "none":
                return advantages.float()
            # group / group
            G = bs // self.group_size
            rows = G * self.group_size
            x = advantages[:rows].float().contiguous()
            m = None if loss_mask is None else loss_mask[:rows].to(torch.float32).contiguous()
            if x.numel() == 0:
                return x
            ge = x.numel() // G
            if not torch.compiler.is_compiling():  # moments -> finalize -> normalise in one host call
                return ops._raw_masked_normalize_groups(
                    x.view(-1), None if m is None else m.view(-1), ge, bool(unbiased), float(eps)).view(x.shape)
            mom = ops.masked_moments(x.view(-1), None if m is None else m.view(-1), ge, None, 1)
            mean, denom = ops.moments_finalize(mom, False, bool(unbiased), float(eps))
            return ops.masked_normalize(
                x.view(-1), None if m is None else m.view(-1), mean, ge, denom, ge, True
            ).view(x.shape)

        # ---- mixed levels ---------------------------------------------------------------
        x = advantages.float().contiguous()
        m = None if loss_mask is None else loss_mask.to(torch.float32).contiguous()
        n = x.numel()
        if n == 0:
            # keep an empty data-parallel shard in step with the others: one zero-moment
            # all-reduce per batch-level statistic, in the order the non-empty ranks issue them
            if dist.is_initialized():
                for level in (self.mean_level, self.std_level):
                    if level == "batch":
                        dist.all_reduce(torch.zeros(1, 3, dtype=torch.float64, device=x.device),
                                        op=dist.ReduceOp.SUM, group=reduce_group)
            return x
        xf = x.view(-1)
        mf = None if m is None else m.view(-1)
        row_elems = n // bs
        need_groups = "group" in (self.mean_level, self.std_level)
        if need_groups:
            gs = self.group_size
            G = bs // gs
            ge = gs * row_elems
            Gp = (bs + gs - 1) // gs  # padded group count (last one partial)
            n_full = G * ge

        def allreduce(mom):
            if dist.is_initialized():
                dist.all_reduce(mom, op=dist.ReduceOp.SUM, group=reduce_group)
            return mom

        # step 1: mean (actor.py:478-494)
        mean = None
        mean_ge = n
        if self.mean_level == "batch":
            mom = allreduce(ops.masked_moments(xf, mf, n, None, 1))
            mean, _ = ops.moments_finalize(mom, False, False, 0.0)
        elif self.mean_level == "group":
            mean = torch.zeros(Gp, dtype=torch.float64, device=x.device)
            if G > 0:
                mom = ops.masked_moments(xf[:n_full], None if mf is None else mf[:n_full], ge, None, 1)
                mg, _ = ops.moments_finalize(mom, False, False, 0.0)
                mean[:G] = mg
            mean_ge = ge
        # step 2: std around that mean (actor.py:503-533, 568-604)
        denom = None
        denom_ge = n
        if self.std_level == "batch":
            mom = allreduce(ops.masked_moments(xf, mf, n, mean, mean_ge))
            _, denom = ops.moments_finalize(mom, True, bool(unbiased), float(eps))
        elif self.std_level == "group":
            denom = torch.full((Gp,), float(eps), dtype=torch.float64, device=x.device)
            if G > 0:
                mom = ops.masked_moments(
                    xf[:n_full], None if mf is None else mf[:n_full], ge, mean, mean_ge
                )
                _, dg = ops.moments_finalize(mom, True, bool(unbiased), float(eps))
                denom[:G] = dg
            denom_ge = ge
        return ops.masked_normalize(xf, None, mean, mean_ge, denom, denom_ge, False).view(x.shape)


@torch.no_grad()
def group_advantage_stats(
    advantages: torch.Tensor,
    loss_mask: Optional[torch.Tensor],
    group_size: Optional[int],
    eps: float = 1e-5,
    unbiased: bool = False,
    reduce_group=None,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """(mean[G], denom[G] = std + eps) of the masked advantages per group of ``group_size``
    consecutive rows (``None`` = the whole batch, all-reduced when torch.distributed runs) —
    the inputs of the fused normalise-inside-the-loss kernel."""
    x = advantages.float().contiguous()
    m = None if loss_mask is None else loss_mask.to(torch.float32).contiguous()
    if group_size is None:
        ge = x.numel()
    else:
        if x.size(0) % group_size != 0:
            raise ValueError("batch size must be a multiple of group_size")
        ge = group_size * (x.numel() // x.size(0))
    mom = ops.masked_moments(x.view(-1), None if m is None else m.view(-1), ge, None, 1)
    if group_size is None and dist.is_initialized():
        dist.all_reduce(mom, op=dist.ReduceOp.SUM, group=reduce_group)
    return ops.moments_finalize(mom, False, bool(unbiased), float(eps))


def ppo_actor_loss_fn_fused_advnorm(
    logprobs: torch.Tensor,
    proximal_logprobs: torch.Tensor,
    old_logprobs: torch.Tensor,
    raw_advantages: torch.Tensor,
    eps_clip: float,
    loss_mask: torch.Tensor,
    adv_mean: torch.Tensor,
    adv_denom: torch.Tensor,
    adv_group: Optional[torch.Tensor] = None,
    eps_clip_higher: Optional[float] = None,
    c_clip: Optional[float] = None,
    behav_imp_weight_cap: Optional[float] = None,
) -> Tuple[torch.Tensor, Dict, torch.Tensor]:
    """``ppo_actor_loss_fn`` on UN-normalised advantages: the kernel applies
    ``(adv*mask - mean[g]) / denom[g]`` (g = ``adv_group[t]``, the token's GRPO group) on the fly,
    forward and backward, and also returns the normalised advantages it used (they are logged by
    the reference, areal/engine/ppo/actor.py:200-205). Equivalent to AdvNorm(group, group) followed
    by ppo_actor_loss_fn, minus one [B, L] round trip through HBM and the per-group python loop."""
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    out = ops.ppo_loss_fwd(
        logprobs, proximal_logprobs, old_logprobs, raw_advantages, loss_mask,
        1.0 - eps_clip, 1.0 + (eps_clip if eps_clip_higher is None else eps_clip_higher),
        -1.0 if c_clip is None else float(c_clip),
        -1.0 if behav_imp_weight_cap is None else float(behav_imp_weight_cap),
        adv_mean, adv_denom, adv_group,
    )
    det = [t.detach() for t in out]
    stat = dict(
        loss=det[4], importance_weight=det[5], approx_kl=det[6],
        clip_mask=det[7].view(torch.bool), dual_clip_mask=det[8].view(torch.bool),
        behave_imp_weight=det[9], behave_approx_kl=det[10], behave_mask=det[11].view(torch.bool),
    )
    return out[0], stat, det[3]


@torch.no_grad()
def group_reward_norm(reward_score: torch.Tensor, group_size: int) -> torch.Tensor:
    """Per-group reward normalisation of ``compute_advantages`` (ref actor.py:94-98):
    ``(r - mean) / (std_unbiased + 1e-9)`` over each complete group of ``group_size``
    sequences; rewards beyond the last complete group are left untouched."""
    bs = reward_score.shape[0]
    G = bs // group_size
    out = reward_score.clone()
    if G == 0:
        return out
    rows = G * group_size
    x = reward_score[:rows].float().contiguous()
    mom = ops.masked_moments(x, None, group_size, None, 1)
    mean, denom = ops.moments_finalize(mom, False, True, 1e-9)
    out[:rows] = ops.masked_normalize(x, None, mean, group_size, denom, group_size, False).to(out.dtype)
    return out


def grpo_loss_fn_fused(
    logits: torch.Tensor,
    input_data: Dict,
    temperature: float,
    eps_clip: float,
    eps_clip_higher: float | None,
    c_clip: float | None,
    behav_imp_weight_cap: float | None,
    loss_scale: Optional[torch.Tensor] = None,
    stats: Optional[Dict] = None,
):
    """``grpo_loss_fn`` and its ``backward()`` in one pass over the logits (4*V instead of 6*V bytes
    of HBM traffic per token; see csrc/grpo_fused.cu).

    Same inputs as ``grpo_loss_fn`` plus ``loss_scale``: the factor the engine multiplies the loss
    by before ``backward()`` (``loss_weight_fn(mb) / total_loss_weight * dp_world_size``,
    fsdp_engine.py:543-547 — it does not depend on the loss, so compute it first and pass it in).
    Returns ``loss * loss_scale``; do NOT multiply again. ``logits.grad`` is produced during this
    call and handed out by ``backward()``; if the gradient that arrives there is not 1 the stored
    gradient is rescaled in place, so the result is right (just slower) for any caller.
    Falls back to ``grpo_loss_fn`` for layouts the fused kernel does not take."""
    from . import ops as _ops

    input_ids = input_data["input_ids"]
    old_logp = input_data["logprobs"]
    advantages = input_data["advantages"]
    prox_logp = input_data["prox_logp"]
    st = _stats_tracker()
    want_stats = stats is not None or st is not None
    out = None
    if (logits.is_cuda and logits.dim() == 2 and logits.requires_grad and torch.is_grad_enabled()
            and input_ids.dim() == 1 and input_ids.dtype == torch.int64 and input_ids.is_contiguous()
            and logits.stride(1) == 1 and not torch.compiler.is_compiling()):
        if c_clip is not None:
            assert c_clip > 1.0, c_clip
        up = None
        if loss_scale is not None:
            up = loss_scale if isinstance(loss_scale, torch.Tensor) else torch.tensor(
                float(loss_scale), dtype=torch.float32, device=logits.device)
        # one host call: mask count + the fused kernel; the label roll, the mask cast and the logged
        # tensors all happen inside it (csrc/grpo_fused.cu, b200rl_grpo_fused_step_ids)
        out = _ops._raw_grpo_fused_step_ids(
            logits.detach(), input_ids, True, prox_logp, old_logp, advantages, input_data["loss_mask"],
            float(temperature), 1.0 - eps_clip, 1.0 + (eps_clip if eps_clip_higher is None else eps_clip_higher),
            -1.0 if c_clip is None else float(c_clip),
            -1.0 if behav_imp_weight_cap is None else float(behav_imp_weight_cap), up, want_stats)
    if out is None:
        loss = grpo_loss_fn(logits, input_data, temperature, eps_clip, eps_clip_higher, c_clip,
                            behav_imp_weight_cap, stats=stats)
        return loss * loss_scale if loss_scale is not None else loss
    loss_unscaled, dlogits, r, up = out
    loss = _ops._GrpoFusedFn.apply(logits, loss_unscaled, dlogits, up)
    if want_stats:
        collected = dict(
            n_valid_tokens=r["valid"], clipped_tokens=r["s_clip"], dual_clipped_tokens=r["s_dual"],
            importance_weight=r["s_iw"], approx_kl=r["s_kl"], new_logp=r["logp"], old_logp=old_logp,
            entropy=r["entropy"], actor_loss=r["s_loss"], clip_ratio=r["clip_ratio"],
            dual_clip_ratio=r["dual_clip_ratio"], unclipped_behave_tokens=r["s_bm"],
            behave_imp_weight=r["s_bw"], behave_approx_kl=r["s_bkl"], vocab_min_logits=r["rowmin"],
            vocab_max_logits=r["rowmax"], clipped_new_logp=r["clipped_new_logp"],
            clipped_old_logp=r["clipped_old_logp"],
        )
        if stats is not None:
            stats.update(collected)
        _register_stats(st, logits, collected)
    return loss


def _register_stats(st, logits, c):
    if st is None:
        return
    st.denominator(
        n_tokens=torch.ones(logits.shape[0], dtype=torch.bool, device=logits.device),
        n_valid_tokens=c["n_valid_tokens"], clipped_tokens=c["clipped_tokens"],
        dual_clipped_tokens=c["dual_clipped_tokens"],
    )
    st.stat(
        importance_weight=c["importance_weight"], approx_kl=c["approx_kl"], new_logp=c["new_logp"],
        old_logp=c["old_logp"], entropy=c["entropy"], actor_loss=c["actor_loss"],
        clip_ratio=c["clip_ratio"], dual_clip_ratio=c["dual_clip_ratio"], denominator="n_valid_tokens",
    )
    st.denominator(unclipped_behave_tokens=c["unclipped_behave_tokens"])
    st.stat(behave_imp_weight=c["behave_imp_weight"], behave_approx_kl=c["behave_approx_kl"],
            denominator="unclipped_behave_tokens")
    st.stat(vocab_min_logits=c["vocab_min_logits"], vocab_max_logits=c["vocab_max_logits"],
            denominator="n_tokens")
    st.stat(clipped_new_logp=c["clipped_new_logp"], clipped_old_logp=c["clipped_old_logp"],
            denominator="clipped_tokens")


def _stats_tracker():
    """The reference's ``areal.utils.stats_tracker`` when this module runs inside AReaL, else None.
    Looked up in ``sys.modules`` (the engine imports it long before the first loss call): a failed
    ``import areal...`` on every call costs ~100 us of path walking, more than the kernels of a
    small micro-batch."""
    if torch.compiler.is_compiling():  # the tracker is python state with locks: eager only
        return None
    return sys.modules.get("areal.utils.stats_tracker")


def grpo_loss_fn(
    logits: torch.Tensor,
    input_data: Dict,
    temperature: float,
    eps_clip: float,
    eps_clip_higher: float | None,
    c_clip: float | None,
    behav_imp_weight_cap: float | None,
    stats: Optional[Dict] = None,
):
    """Loss function of the actor step on one packed micro-batch (ref actor.py:292-367).

    Same inputs and return value. The per-row max/min logits that the reference obtains with two
    more passes over [T, V] (:351-352) come out of the single forward pass here. Statistics are
    registered with ``areal.utils.stats_tracker`` when that module is importable; pass a dict as
    ``stats`` to receive the same tensors directly."""
    input_ids = input_data["input_ids"]
    old_logp = input_data["logprobs"]
    advantages = input_data["advantages"]
    prox_logp = input_data["prox_logp"]
    st = _stats_tracker()
    want_stats = stats is not None or st is not None

    if (FAST_PATH and logits.is_cuda and logits.dim() == 2 and input_ids.dim() == 1 and input_ids.dtype == torch.int64
            and input_ids.is_contiguous() and logits.stride(1) == 1 and logits.dtype in _lib_dtypes()
            and not torch.compiler.is_compiling()):
        # eager fast path: the whole function is TWO launches (log-prob forward with the label roll
        # as an index computation; loss kernel that reads the engine's float mask as it is and
        # writes every logged tensor), its backward ONE (csrc/grpo_loss.cu). Nobody consumes the
        # logging tensors -> entropy and row min/max are not even computed.
        if c_clip is not None:
            assert c_clip > 1.0, c_clip
        loss, r = ops.grpo_loss(
            logits, input_ids, True, prox_logp, old_logp, advantages, input_data["loss_mask"],
            float(temperature), 1.0 - eps_clip, 1.0 + (eps_clip if eps_clip_higher is None else eps_clip_higher),
            -1.0 if c_clip is None else float(c_clip),
            -1.0 if behav_imp_weight_cap is None else float(behav_imp_weight_cap), want_stats)
        if not want_stats:
            return loss
        collected = dict(
            n_valid_tokens=r["valid"], clipped_tokens=r["s_clip"], dual_clipped_tokens=r["s_dual"],
            importance_weight=r["s_iw"], approx_kl=r["s_kl"], new_logp=r["logp"], old_logp=old_logp,
            entropy=r["entropy"], actor_loss=r["s_loss"], clip_ratio=r["clip_ratio"],
            dual_clip_ratio=r["dual_clip_ratio"], unclipped_behave_tokens=r["s_bm"],
            behave_imp_weight=r["s_bw"], behave_approx_kl=r["s_bkl"], vocab_min_logits=r["rowmin"],
            vocab_max_logits=r["rowmax"], clipped_new_logp=r["clipped_new_logp"],
            clipped_old_logp=r["clipped_old_logp"],
        )
        if stats is not None:
            stats.update(collected)
        _register_stats(st, logits, collected)
        return loss

    # custom-op path (torch.compile, leading batch dims, int32 ids, ...)
    loss_mask = input_data["loss_mask"].bool()
    labels = torch.roll(input_ids, shifts=-1, dims=-1)
    logprobs, entropy, vmax, vmin = functional.gather_logprobs_entropy_minmax(
        logits, labels, temperature
    )
    entropy = entropy.detach()
    loss, stat = functional.ppo_actor_loss_fn(
        logprobs=logprobs,
        old_logprobs=old_logp,
        advantages=advantages,
        eps_clip=eps_clip,
        eps_clip_higher=eps_clip_higher,
        loss_mask=loss_mask,
        c_clip=c_clip,
        proximal_logprobs=prox_logp,
        behav_imp_weight_cap=behav_imp_weight_cap,
    )
    if not want_stats:
        return loss  # nobody consumes the logging tensors: do not launch the derived ones
    clip_mask = stat["clip_mask"]
    new_logp = logprobs.detach()
    collected = dict(
        n_valid_tokens=loss_mask,
        clipped_tokens=clip_mask,
        dual_clipped_tokens=stat["dual_clip_mask"],
        importance_weight=stat["importance_weight"],
        approx_kl=stat["approx_kl"],
        new_logp=new_logp,
        old_logp=old_logp,
        entropy=entropy.float(),
        actor_loss=stat["loss"],
        clip_ratio=clip_mask.float(),
        dual_clip_ratio=stat["dual_clip_mask"].float(),
        unclipped_behave_tokens=stat["behave_mask"],
        behave_imp_weight=stat["behave_imp_weight"],
        behave_approx_kl=stat["behave_approx_kl"],
        vocab_min_logits=vmin.detach(),
        vocab_max_logits=vmax.detach(),
        clipped_new_logp=torch.where(clip_mask, new_logp, 0.0),
        clipped_old_logp=torch.where(clip_mask, old_logp, 0.0),
    )
    if stats is not None:
        stats.update(collected)
    _register_stats(st, logits, collected)
    return loss


def _lib_dtypes():
    from . import _lib

    return _lib.DTYPE_CODE
