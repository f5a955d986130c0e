"""Drop-in for the loss function of ``areal/engine/sft/lm_engine.py`` (the SFT twin of
``grpo_loss_fn``): ``compute_packed_sft_loss(logits, input_)``.

Reference (lm_engine.py:45-83): gather_logprobs over the rolled input ids, masked mean of -logp,
a python loop over ``cu_seqlens`` for the per-sequence mean log-prob (one host sync per slice
bound), and two more full passes over ``[T, V]`` for ``logits.min(-1)`` / ``max(-1)``.
Here: one streaming pass over the logits yields logp AND the row min/max; one small launch yields
the loss, per-sequence log-prob / perplexity and the gradient seed. No host sync.
"""

from __future__ import annotations

import sys
from typing import Dict, Optional

import torch

from . import ops
from .functional import _as_2d


def _stats_tracker():
    """The reference's ``areal.utils.stats_tracker`` when this module runs inside AReaL, else None.
    Looked up in ``sys.modules`` (the engine imports it long before the first loss call): a failed
    ``import areal...`` on every call costs ~100 us of path walking, more than the kernels of a
    small micro-batch."""
    if torch.compiler.is_compiling():  # the tracker is python state with locks: eager only
        return None
    return sys.modules.get("areal.utils.stats_tracker")


def compute_packed_sft_loss(logits: torch.Tensor, input_: Dict, stats: Optional[Dict] = None) -> torch.Tensor:
    """ref areal/engine/sft/lm_engine.py:45-83. ``stats`` (extension): when a dict is passed the
    logged tensors are also stored there (ppl, loss, vocab_min_logits, vocab_max_logits, seqlogp,
    n_valid_tokens); inside AReaL they are registered with ``stats_tracker`` under the
    reference's keys and denominators."""
    packed_input_ids: torch.Tensor = input_["input_ids"]
    cu_seqlens: torch.Tensor = input_["cu_seqlens"]
    loss_mask = input_["loss_mask"].bool()

    labels = torch.roll(packed_input_ids, shifts=-1, dims=-1)
    loss_mask = torch.roll(loss_mask, shifts=-1, dims=-1)
    l2, y, lead = _as_2d(logits, labels)
    logp, _, _, vmax, vmin = ops.logprob_entropy_fwd(l2, y, 1.0, False, True)
    out = ops.sft_loss_fwd(logp.view(-1), loss_mask.reshape(-1), cu_seqlens)
    loss = out[0]
    with torch.no_grad():
        collected = dict(
            ppl=out[5].detach(),
            loss=-out[3].detach().view(lead),
            vocab_min_logits=vmin.detach().view(lead),
            vocab_max_logits=vmax.detach().view(lead),
            seqlogp=out[4].detach(),
            n_valid_tokens=loss_mask,
        )
        if stats is not None:
            stats.update(collected)
        st = _stats_tracker()
        if st is not None:
            st.denominator(
                n_seqs=torch.ones(cu_seqlens.shape[0] - 1, dtype=torch.bool, device=logits.device),
                n_tokens=torch.ones(logits.shape[0], dtype=torch.bool, device=logits.device),
                n_valid_tokens=loss_mask,
                prompt_tokens=loss_mask.logical_not(),
            )
            st.stat(ppl=collected["ppl"], denominator="n_seqs")
            st.stat(loss=collected["loss"], denominator="n_valid_tokens")
            st.stat(vocab_min_logits=collected["vocab_min_logits"],
                    vocab_max_logits=collected["vocab_max_logits"], denominator="n_tokens")
    return loss


def compute_packed_sft_loss_fused(logits: torch.Tensor, input_: Dict, loss_scale=None,
                                  stats: Optional[Dict] = None) -> torch.Tensor:
    """``compute_packed_sft_loss`` and its ``backward()`` in one pass over the logits: each row is
    read from HBM once (forward statistics) and again from L2 to write ``logits.grad`` (4*V instead
    of 6*V bytes per token, the SFT twin of ``actor.grpo_loss_fn_fused``). ``loss_scale`` is the
    factor the engine multiplies the loss by before ``backward()`` (fsdp_engine.py:543-547); the
    return value is already multiplied by it. Falls back to ``compute_packed_sft_loss`` for layouts
    the fused kernel does not take (then the scale is applied on top)."""
    out = None
    if logits.is_cuda and logits.dim() == 2 and logits.requires_grad and torch.is_grad_enabled() \
            and not torch.compiler.is_compiling():
        labels = torch.roll(input_["input_ids"], shifts=-1, dims=-1)
        loss_mask = torch.roll(input_["loss_mask"].bool(), shifts=-1, dims=-1)
        up = None
        if loss_scale is not None:
            up = loss_scale if isinstance(loss_scale, torch.Tensor) else torch.tensor(
                float(loss_scale), dtype=torch.float32, device=logits.device)
        dummy = torch.zeros(logits.shape[0], dtype=torch.float32, device=logits.device)
        out = ops._raw_grpo_fused_step(logits.detach(), labels, dummy, dummy, dummy, loss_mask, 1.0,
                                       1.0, 1.0, -1.0, -3.0, up, True)
    if out is None:
        loss = compute_packed_sft_loss(logits, input_, stats=stats)
        return loss * loss_scale if loss_scale is not None else loss
    loss = ops._GrpoFusedFn.apply(logits, out["loss"], out["dlogits"], up)
    st = _stats_tracker()
    if stats is not None or st is not None:
        cu_seqlens = input_["cu_seqlens"]
        with torch.no_grad():
            seq = ops._raw_sft_loss_fwd(out["logp"], loss_mask, cu_seqlens)  # per-sequence ppl only
            collected = dict(ppl=seq[5], loss=out["s_loss"], vocab_min_logits=out["rowmin"],
                             vocab_max_logits=out["rowmax"], seqlogp=seq[4], n_valid_tokens=loss_mask)
            if stats is not None:
                stats.update(collected)
            if st is not None:
                st.denominator(
                    n_seqs=torch.ones(cu_seqlens.shape[0] - 1, dtype=torch.bool, device=logits.device),
                    n_tokens=torch.ones(logits.shape[0], dtype=torch.bool, device=logits.device),
                    n_valid_tokens=loss_mask, prompt_tokens=loss_mask.logical_not())
                st.stat(ppl=collected["ppl"], denominator="n_seqs")
                st.stat(loss=collected["loss"], denominator="n_valid_tokens")
                st.stat(vocab_min_logits=collected["vocab_min_logits"],
                        vocab_max_logits=collected["vocab_max_logits"], denominator="n_tokens")
    return loss
