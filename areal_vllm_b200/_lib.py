"""ctypes binding of libb200rl.so (the C ABI declared in include/b200rl.h).

There is no fallback: if the shared library is missing, importing this module raises, and every
wrapper refuses non-CUDA tensors. Build the library with ``python -c "import __graft_entry__ as
g; g.build()"`` (or ``make -C areal_vllm_b200/csrc``).
"""

from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200RL_LIB") or os.path.join(_HERE, "libb200rl.so")  # override: tuning builds

BF16, FP16, FP32 = 0, 1, 2
DTYPE_CODE = {torch.bfloat16: BF16, torch.float16: FP16, torch.float32: FP32}

GAE_AUTO, GAE_SERIAL, GAE_SCAN = 0, 1, 2

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: the CUDA library has not been built. Run "
        "`python -c 'import __graft_entry__ as g; g.build()'` at the repo root. "
        "This package has no CPU or PyTorch fallback."
    )

lib = ctypes.CDLL(LIB_PATH)

# name -> (restype, argtypes); must list every symbol of include/b200rl.h
SIGNATURES = {
    "b200rl_last_error": (ctypes.c_char_p, []),
    "b200rl_version": (c_int, []),
    "b200rl_logprob_entropy_fwd": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_int64, c_int64, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "b200rl_logprob_entropy_bwd": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_int64, c_int64, c_float, c_void_p, c_int64, c_void_p],
    ),
    "b200rl_grpo_loss_workspace_bytes": (c_size_t, [c_int64]),
    "b200rl_grpo_loss_fwd": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_int, c_int64, c_int64, c_float,   # logits..temperature
         c_void_p, c_void_p, c_void_p, c_void_p, c_int,                          # prox, old, adv, mask, mask_is_f32
         c_float, c_float, c_float, c_float,                                     # clip_lo/hi, c_clip, cap
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,                       # logp, entropy, lse, rowmax, rowmin
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,  # 8 stats
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,                       # valid, 2 ratios, 2 clipped logp
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],            # dlogp, loss, denom, ws, stream
    ),
    "b200rl_grpo_loss_bwd": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_int, c_int64, c_int64, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "b200rl_grpo_fused_workspace_bytes": (c_size_t, [c_int64]),
    "b200rl_grpo_fused_step": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_int64, c_int64, c_float,        # logits..temperature
         c_void_p, c_void_p, c_void_p, c_void_p,                               # prox, old, adv, mask
         c_float, c_float, c_float, c_float, c_void_p,                         # clip_lo/hi, c_clip, cap, upstream
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,                     # logp, entropy, lse, rowmax, rowmin
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,  # 8 stats
         c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],  # dlogits, ld_out, loss, denom, ws
    ),
    "b200rl_grpo_fused_step_ids": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_int, c_int64, c_int64, c_float,   # logits..temperature
         c_void_p, c_void_p, c_void_p, c_void_p, c_int,                          # prox, old, adv, mask, mask_is_f32
         c_float, c_float, c_float, c_float, c_void_p,                           # clip_lo/hi, c_clip, cap, upstream
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,                       # logp, entropy, lse, rowmax, rowmin
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,  # 8 stats
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,                       # valid, 2 ratios, 2 clipped logp
         c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],   # dlogits, ld_out, loss, denom, ws
    ),
    "b200rl_scale_inplace": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p]),
    "b200rl_gae_workspace_bytes": (c_size_t, [c_int32, c_int64]),
    "b200rl_gae_1d_nolp_misalign": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int64,
         c_float, c_float, c_int, c_void_p, c_size_t, c_void_p],
    ),
    "b200rl_gae_1d_nolp_misalign_noalloc": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int64,
         c_float, c_float, c_void_p],
    ),
    "b200rl_ppo_loss_workspace_bytes": (c_size_t, [c_int64]),
    "b200rl_ppo_loss_fwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float,
         c_float, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "b200rl_ppo_loss_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "b200rl_critic_loss_workspace_bytes": (c_size_t, [c_int64]),
    "b200rl_critic_loss_fwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_int, c_float,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "b200rl_packed_rewards": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_float, c_float, c_int,
         c_void_p, c_void_p, c_void_p],
    ),
    "b200rl_logprob_shard_fwd": (
        c_int,
        [c_void_p, c_int, c_int64, c_void_p, c_int64, c_int64, c_float, c_int, c_int, c_void_p, c_void_p],
    ),
    "b200rl_logprob_shard_combine": (
        c_int,
        [c_void_p, c_int32, c_int64, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "b200rl_tp_exchange_bytes": (c_size_t, [c_int32, c_int64]),
    "b200rl_tp_push": (c_int, [c_void_p, c_int64, c_void_p, c_int32, c_int32, c_int64, ctypes.c_uint32, c_void_p]),
    "b200rl_tp_combine": (
        c_int,
        [c_void_p, c_int32, c_int64, c_int64, ctypes.c_uint32, c_float, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p],
    ),
    "b200rl_gae_2d_olp": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32,
         c_int32, c_float, c_float, c_void_p],
    ),
    "b200rl_move_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int]),
    "b200rl_move_intervals": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int,
         c_void_p, c_size_t, c_int, c_void_p],
    ),
    "b200rl_enable_peer_access": (c_int, [c_int, c_int]),
    "b200rl_ipc_export": (c_int, [c_void_p, c_void_p, c_void_p]),
    "b200rl_ipc_open": (c_int, [c_void_p, c_int, c_void_p]),
    "b200rl_ipc_close": (c_int, [c_void_p, c_int]),
    "b200rl_sft_loss_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "b200rl_sft_loss_fwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int32, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "b200rl_moments_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "b200rl_masked_moments": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_size_t,
         c_void_p],
    ),
    "b200rl_moments_finalize": (
        c_int, [c_void_p, c_int64, c_int, c_int, c_double, c_void_p, c_void_p, c_void_p]
    ),
    "b200rl_masked_normalize": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p,
         c_void_p],
    ),
    "b200rl_masked_normalize_groups": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_size_t, c_void_p],
    ),
    "b200rl_interval_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int]),
    "b200rl_slice_intervals": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_size_t,
         c_int, c_void_p],
    ),
    "b200rl_set_intervals": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p, c_size_t,
         c_int, c_void_p],
    ),
    "b200rl_merge_intervals": (c_size_t, [c_void_p, c_size_t, c_void_p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here = header and library disagree
    _fn.restype = _res
    _fn.argtypes = _args


def last_error() -> str:
    msg = lib.b200rl_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, what: str) -> None:
    """C-ABI return code -> exception (RuntimeError, as TORCH_CHECK raises in the reference)."""
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {last_error()}")


def ptr(t: torch.Tensor | None) -> c_void_p | None:
    return None if t is None else c_void_p(t.data_ptr())


def stream_handle(device: torch.device) -> c_void_p:
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(**tensors: torch.Tensor | None) -> None:
    for name, t in tensors.items():
        if t is None:
            continue
        if not isinstance(t, torch.Tensor):
            raise TypeError(f"{name} must be a torch.Tensor, got {type(t)}")
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be on CUDA (b200rl has no CPU path)")


def require_contiguous(**tensors: torch.Tensor | None) -> None:
    for name, t in tensors.items():
        if t is not None and not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")


def workspace(nbytes: int, device: torch.device) -> torch.Tensor:
    """Caller-owned scratch for a C-ABI call (torch's caching allocator hands out 512-B aligned blocks)."""
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
