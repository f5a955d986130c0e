"""Drop-in for the GAE entry points and the legacy actor loss of
``realhf/impl/model/utils/ppo_functional.py`` and for the ``cugae`` / ``realhf._C.cugae`` binding.

Reference: ppo_functional.py:51-149 (actor_loss_fn), :292-323 (pygae1d_nolp_misalign),
:326-366 (cugae1d_nolp_misalign_func), :369-394 (get_packed_advantages_and_returns);
csrc/cugae/gae.cu:30-60 (gae_1d_nolp_misalign).
"""

from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib, ops

__all__ = [
    "gae_1d_nolp_misalign",
    "cugae1d_nolp_misalign_func",
    "pygae1d_nolp_misalign",
    "get_packed_advantages_and_returns",
    "actor_loss_fn",
]

_VALIDATE_DEFAULT = os.environ.get("B200RL_VALIDATE", "0") not in ("", "0")

_MODES = {"auto": _lib.GAE_AUTO, "serial": _lib.GAE_SERIAL, "scan": _lib.GAE_SCAN,
          # SERIAL chooses among these by batch shape (identical bits); forcing is for tests
          "serial_warp": 3, "serial_batch8": 4, "serial_batch32": 5, "serial_slots": 6, "serial_relay": 7,
          "serial_team": 8}


def gae_1d_nolp_misalign(
    rewards: torch.Tensor,
    values: torch.Tensor,
    cu_seqlens: torch.Tensor,
    bootstrap: torch.Tensor,
    gamma: float,
    lmbda: float,
    mode: str = "auto",
    validate: Optional[bool] = None,
) -> List[torch.Tensor]:
    """The pybind entry ``realhf._C.cugae.gae_1d_nolp_misalign`` (gae.cu:30-60): same argument
    checks (device, contiguity, shapes, dtypes) raising RuntimeError, returns ``[adv, ret]``.
    ``mode``: "auto" | "serial" | "scan" (see include/b200rl.h).

    The reference also checks ``cu_seqlens[0] == 0`` and ``cu_seqlens[-1] == total`` (gae.cu:48-49)
    with two ``.item()`` calls, i.e. two device->host syncs per call, and zero-fills its outputs.
    By default neither happens here (the step must not sync; the kernels overwrite every token of a
    well-formed ``cu_seqlens``), so a MALFORMED ``cu_seqlens`` is undefined behaviour: tokens
    outside ``[cu[0], cu[-1])`` stay uninitialised and ``cu[-1] > total`` reads/writes out of
    bounds. Pass ``validate=True`` (or set ``B200RL_VALIDATE=1``) to get the reference's check and
    its RuntimeError at the price of one sync - e.g. while bringing up a new data pipeline."""
    if validate is None:
        validate = _VALIDATE_DEFAULT
    if validate and cu_seqlens.numel() >= 1 and cu_seqlens.is_cuda:
        first, last = (int(x) for x in cu_seqlens[[0, -1]].tolist())  # one sync
        if first != 0 or last != rewards.shape[0]:
            raise RuntimeError(
                f"cu_seqlens must start at 0 and end at the total sequence length "
                f"({rewards.shape[0]}), got [{first}, ..., {last}]")
    adv, ret = ops.gae_1d_nolp_misalign(
        rewards, values, cu_seqlens, bootstrap, float(gamma), float(lmbda), _MODES[mode]
    )
    return [adv, ret]


def cugae1d_nolp_misalign_func(
    rewards: torch.Tensor,
    values: torch.Tensor,
    cu_seqlens: torch.Tensor,
    truncate: torch.Tensor,
    gamma: float,
    lam: float,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """GAE over packed sequences; ``values`` is longer than ``rewards`` by one bootstrap slot
    per sequence, ``truncate[i]`` says whether sequence i's last next-value is bootstrapped
    (ref ppo_functional.py:326-366). The reference's two asserts on ``cu_seqlens[0]`` /
    ``cu_seqlens[-1]`` force device->host syncs (:365); they are dropped."""
    assert len(rewards.shape) == len(values.shape) == len(cu_seqlens.shape) == 1
    adv, ret = gae_1d_nolp_misalign(rewards, values, cu_seqlens, truncate, gamma, lam)
    return adv, ret


def pygae1d_nolp_misalign(
    rewards: torch.Tensor,
    values: torch.Tensor,
    cu_seqlens_: torch.Tensor,
    bootstrap: torch.Tensor,
    gamma: float,
    lam: float,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Same contract as the reference's python loop (ppo_functional.py:292-323), computed by the
    CUDA kernel in its reference-order (serial) mode. Unlike the reference it does not zero
    ``values`` in place when ``bootstrap[i] == 0`` (:313-315 mutates through a view)."""
    if not rewards.is_cuda:
        raise RuntimeError(
            "pygae1d_nolp_misalign: CPU tensors are not supported (b200rl has no CPU path; "
            "the CPU oracle lives under oracle/ and is test infrastructure)"
        )
    adv, ret = gae_1d_nolp_misalign(
        rewards.float(), values.float(), cu_seqlens_.int(), bootstrap.bool(), gamma, lam, mode="serial"
    )
    return adv, ret


@torch.no_grad()
def get_packed_advantages_and_returns(
    gamma: float,
    lam: float,
    values: torch.Tensor,
    rewards: torch.Tensor,
    short1cu_seqlens: torch.Tensor,
    seq_no_eos_mask: torch.Tensor,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """ref ppo_functional.py:369-394. The reference silently falls back to the python loop on
    CPU tensors or when the extension is missing; here CUDA is the only path."""
    return cugae1d_nolp_misalign_func(
        rewards, values, short1cu_seqlens.int(), seq_no_eos_mask.bool(), gamma, lam
    )


def actor_loss_fn(
    logprobs: torch.Tensor,
    old_logprobs: torch.Tensor,
    advantages: torch.Tensor,
    eps_clip: float,
    loss_mask: Optional[torch.Tensor] = None,
    c_clip: Optional[float] = None,
    proximal_logprobs: Optional[torch.Tensor] = None,
    behav_imp_weight_cap: Optional[float] = None,
) -> Tuple[torch.Tensor, Dict]:
    """Legacy twin of ppo_actor_loss_fn (ref ppo_functional.py:51-149): no ``eps_clip_higher``;
    ``loss_mask`` optional (all ones); without ``proximal_logprobs`` the ratio is taken against
    ``old_logprobs`` and there is no behaviour weight (and no behave_* stats)."""
    assert logprobs.dtype == torch.float32
    assert old_logprobs.dtype == torch.float32
    assert advantages.dtype == torch.float32
    if c_clip is not None:
        assert c_clip > 1.0, c_clip
    if loss_mask is None:
        loss_mask = torch.ones_like(logprobs, dtype=torch.bool)
    denorm = proximal_logprobs if proximal_logprobs is not None else old_logprobs
    out = ops.ppo_loss_fwd(
        logprobs,
        denorm,
        old_logprobs if proximal_logprobs is not None else denorm,  # prox == old -> weight 1
        advantages,
        loss_mask,
        1.0 - eps_clip,
        1.0 + eps_clip,
        -1.0 if c_clip is None else float(c_clip),
        -2.0 if proximal_logprobs is None
        else (-1.0 if behav_imp_weight_cap is None else float(behav_imp_weight_cap)),
        None,
        None,
        None,
    )
    det = [t.detach() for t in out]
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
    return out[0], stat


def critic_loss_fn(
    value: torch.Tensor,
    old_value: torch.Tensor,
    target_value: torch.Tensor,
    value_eps_clip: float,
    loss_mask: Optional[torch.Tensor] = None,
    loss_fn_type: str = "mse",
) -> Tuple[torch.Tensor, Dict]:
    """Clipped value loss (ref ppo_functional.py:171-225; huber delta 10.0 as at :197). One launch
    yields the scalar, both stats and d loss / d value. Like the reference the masked mean divides
    by ``count_nonzero(mask)`` with no guard, so an all-false mask gives nan."""
    assert value.dtype == torch.float32
    assert old_value.dtype == torch.float32
    assert target_value.dtype == torch.float32
    if loss_fn_type == "huber":
        kind = 1
    elif loss_fn_type == "mse":
        kind = 0
    else:
        raise NotImplementedError(f"Unknown loss fn type: {loss_fn_type}")
    out = ops.critic_loss_fwd(value, old_value, target_value, loss_mask, float(value_eps_clip), kind, 10.0)
    stat = dict(clip_mask=out[4].detach().view(torch.bool), loss=out[3].detach())
    return out[0], stat


@torch.no_grad()
def get_packed_rewards(
    kl_ctl: float,
    clip_reward_value: float,
    log_probs: torch.Tensor,
    ref_log_probs: torch.Tensor,
    reward_score: torch.Tensor,
    short1cu_seqlens: torch.Tensor,
    seq_no_eos_mask: torch.Tensor,
    mask_no_eos_with_zero: bool = False,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """KL-shaped token rewards plus the clipped sequence score on each last token
    (ref ppo_functional.py:228-253). Returns ``(kl_rewards, tot_rewards)``."""
    kl, tot = ops.packed_rewards(log_probs, ref_log_probs, reward_score, short1cu_seqlens,
                                 seq_no_eos_mask, float(kl_ctl), float(clip_reward_value),
                                 bool(mask_no_eos_with_zero))
    return kl, tot


def gae_2d_olp(
    rewards: torch.Tensor,
    values: torch.Tensor,
    done: torch.Tensor,
    done_y_indices: torch.Tensor,
    cu_num_dones: torch.Tensor,
    max_num_dones: int,
    is_truncate: torch.Tensor,
    gamma: float,
    lmbda: float,
):
    """Binding twin of ``cugae.gae_2d_olp`` (ref csrc/cugae/gae.cu:94-131; same argument order,
    checks and error class). Returns ``[adv, ret]``, both ``zeros_like(rewards)``-initialised."""
    for name, t in (("rewards", rewards), ("values", values), ("done", done),
                    ("done_y_indices", done_y_indices), ("cu_num_dones", cu_num_dones),
                    ("is_truncate", is_truncate)):
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be on CUDA")
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")
    if rewards.dim() != 2:
        raise RuntimeError("rewards must be [batch, horizon]")
    B, H = rewards.shape
    n_dones = done_y_indices.numel()
    if tuple(values.shape) != (B, H + 1) or tuple(done.shape) != (B, H + 1):
        raise RuntimeError(f"values and done must have shape ({B}, {H + 1})")
    if tuple(cu_num_dones.shape) != (B + 1,) or tuple(is_truncate.shape) != (n_dones,):
        raise RuntimeError("cu_num_dones must be [batch + 1] and is_truncate [num_dones]")
    if done_y_indices.dtype != torch.int32 or cu_num_dones.dtype != torch.int32:
        raise RuntimeError("done_y_indices and cu_num_dones must be int32")
    if is_truncate.dtype != torch.bool or done.dtype != torch.bool:
        raise RuntimeError("is_truncate and done must be bool")
    if rewards.dtype != torch.float32 or values.dtype != torch.float32:
        raise RuntimeError("rewards and values must be float32")
    adv, ret = torch.zeros_like(rewards), torch.zeros_like(rewards)
    dev = rewards.device
    with torch.cuda.device(dev):
        rc = _lib.lib.b200rl_gae_2d_olp(
            _lib.ptr(rewards), _lib.ptr(values), _lib.ptr(done.view(torch.uint8)), _lib.ptr(done_y_indices),
            _lib.ptr(cu_num_dones), _lib.ptr(is_truncate.view(torch.uint8)), _lib.ptr(adv), _lib.ptr(ret),
            B, H, int(max_num_dones), float(gamma), float(lmbda), _lib.stream_handle(dev))
    _lib.check(rc, "b200rl_gae_2d_olp")
    return [adv, ret]


def gae_2d_nolp(rewards, values, on_reset, on_reset_y_indices, cu_num_resets, max_num_resets,
                bootstrap, gamma, lmbda):
    """Binding twin of ``cugae.gae_2d_nolp``. The reference's host function launches the
    *olp* kernel with these arguments (csrc/cugae/gae.cu:199), so that is what runs here too."""
    return gae_2d_olp(rewards, values, on_reset, on_reset_y_indices, cu_num_resets, max_num_resets,
                      bootstrap, gamma, lmbda)
