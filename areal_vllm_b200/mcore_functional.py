"""Vocabulary-parallel (tensor-parallel) log-prob / entropy: drop-in for the TP branch of the
Megatron actor.

Reference: ``areal/experimental/megatron_actor.py:60-66, 323-331`` upcasts the ``[T, V/tp]`` shard
(``logits.float() / temperature``), then calls ``megatron.core.tensor_parallel.
vocab_parallel_cross_entropy`` (max, gathered target logit and sum-exp: three all-reduces, several
passes, an fp32 softmax saved for backward) and ``_VocabParallelEntropy``
(``areal/experimental/utils/mcore/functional.py:8-52``: three more all-reduces, an fp32 softmax and
the fp32 logits saved). megatron.core is a third-party dependency that is not under
/root/reference; both functions compute exactly ``log_softmax(cat(shards) / T)`` gathered at the
label, and its entropy, so parity is anchored on the whole-vocabulary oracle.

Here: ONE streaming pass over the shard in its own dtype emits an 8-float record per row, ONE
all-gather of ``[T, 8]`` fp32 over the TP group, one tiny combine launch; nothing of size
``[T, V]`` is saved except the shard itself. Backward is one pass over the shard with the global
log-sum-exp and no communication (the upstream gradients are replicated over the TP group).
"""

from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.distributed as dist

from . import ops
from .functional import ppo_actor_loss_fn


def tp_group_from_megatron():
    """The Megatron tensor-parallel group when megatron.core is importable and initialised."""
    try:
        from megatron.core import parallel_state as mpu  # type: ignore
    except Exception:
        return None
    if mpu.is_initialized() and mpu.get_tensor_model_parallel_world_size() > 1:
        return mpu.get_tensor_model_parallel_group()
    return None


def vocab_range(per_partition_vocab_size: int, rank: int, world_size: int) -> Tuple[int, int]:
    """First / one-past-last global column of ``rank``'s shard (Megatron's
    ``VocabUtility.vocab_range_from_per_partition_vocab_size``: equal contiguous shards)."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside [0, {world_size})")
    start = rank * per_partition_vocab_size
    return start, start + per_partition_vocab_size


def gather_partials(partial: torch.Tensor, group=None) -> torch.Tensor:
    """[T, 8] -> [tp, T, 8] in rank order: the one exchange of the forward."""
    if group is None or dist.get_world_size(group) == 1:
        return partial.unsqueeze(0)
    n = dist.get_world_size(group)
    if partial.is_cuda and dist.get_backend(group) == "gloo":
        # a gloo group (several ranks sharing one GPU, or a CPU-side control group): the 32-byte
        # records travel through host memory; everything else stays on the device
        host = torch.empty((n * partial.shape[0],) + tuple(partial.shape[1:]), dtype=partial.dtype)
        dist.all_gather_into_tensor(host, partial.cpu().contiguous(), group=group)
        return host.to(partial.device).view((n,) + tuple(partial.shape))
    out = torch.empty((n * partial.shape[0],) + tuple(partial.shape[1:]), dtype=partial.dtype,
                      device=partial.device)
    dist.all_gather_into_tensor(out, partial.contiguous(), group=group)
    return out.view((n,) + tuple(partial.shape))


class PeerExchange:
    """The forward's one exchange as peer stores instead of an NCCL all-gather (one node).

    Every rank of ``group`` owns a small buffer (flags + two slots of ``[tp, max_tokens, 8]`` fp32
    records) that all other ranks map through CUDA IPC. A call then is: shard kernel ->
    ``b200rl_tp_push`` (32-byte records stored straight into every rank's buffer over NVLink, flag
    behind a system-scope fence) -> ``b200rl_tp_combine`` (waits in-kernel for the tp flags of its
    rows, then folds the records). No collective launch, no host round trip; at tp = 8 and a few
    thousand tokens per micro-batch the all-gather's launch was most of the step.

    Collective constructor (every rank of the group, same ``max_tokens``). Calls must be issued in
    the same order on all ranks, like any collective. ``check()`` syncs and raises if a combine
    ever timed out waiting for a peer (the kernel gives up after ~2 s instead of hanging)."""

    def __init__(self, group=None, max_tokens: int = 65536, device=None):
        import ctypes

        from . import _lib
        from .dist import PeerBuffer, export_buffer

        self.group = group
        self.tp = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.max_tokens = int(max_tokens)
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        nbytes = _lib.lib.b200rl_tp_exchange_bytes(self.tp, self.max_tokens)
        if nbytes == 0:
            raise RuntimeError(f"unsupported exchange shape tp={self.tp}, max_tokens={max_tokens}")
        self.buf = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)  # zero-filled before any peer can push into it
        handles = [None] * self.tp
        dist.all_gather_object(handles, export_buffer(self.buf), group=group)
        self._peers, ptrs = [], []
        for r in range(self.tp):
            if r == self.rank:
                ptrs.append(self.buf.data_ptr())
            else:
                pb = PeerBuffer(*handles[r], dev)
                self._peers.append(pb)
                ptrs.append(pb.data_ptr())
        self.ptrs = (ctypes.c_void_p * self.tp)(*ptrs)
        self.epoch = 0
        dist.barrier(group=group)  # every rank has mapped every buffer

    def exchange_and_combine(self, partial, temperature, need_entropy, need_minmax):
        """partial [T, 8] of this rank -> (logp, entropy or empty, lse, rowmax or empty, rowmin or empty)."""
        from . import _lib
        from ._lib import check, lib, ptr, stream_handle

        T = partial.shape[0]
        if T > self.max_tokens:
            raise RuntimeError(f"{T} tokens exceed the exchange buffer's {self.max_tokens}")
        dev = partial.device
        f32 = dict(dtype=torch.float32, device=dev)
        logp, lse = torch.empty(T, **f32), torch.empty(T, **f32)
        ent = torch.empty(T if need_entropy else 0, **f32)
        rmax = torch.empty(T if need_minmax else 0, **f32)
        rmin = torch.empty(T if need_minmax else 0, **f32)
        self.epoch += 1
        st = stream_handle(dev)
        check(lib.b200rl_tp_push(ptr(partial), T, self.ptrs, self.tp, self.rank, self.max_tokens, self.epoch, st),
              "b200rl_tp_push")
        check(lib.b200rl_tp_combine(ptr(self.buf), self.tp, T, self.max_tokens, self.epoch, float(temperature),
                                    ptr(logp), ptr(ent) if need_entropy else None, ptr(lse),
                                    ptr(rmax) if need_minmax else None, ptr(rmin) if need_minmax else None,
                                    ptr(self.status), st), "b200rl_tp_combine")
        return logp, ent, lse, rmax, rmin

    def check(self) -> None:
        if int(self.status.item()) != 0:
            raise RuntimeError("a vocabulary-parallel combine timed out waiting for a peer's records")

    def close(self) -> None:
        for pb in self._peers:
            pb.close()
        self._peers = []


_EXCHANGES: Dict = {}


def enable_peer_exchange(group=None, max_tokens: int = 65536) -> PeerExchange:
    """Collective over ``group``: from now on ``vocab_parallel_*`` calls on this group exchange
    their records through peer memory (``PeerExchange``) instead of NCCL, for micro-batches of up
    to ``max_tokens`` tokens (larger ones fall back to the all-gather)."""
    group = dist.group.WORLD if group is None else group
    ex = PeerExchange(group, max_tokens)
    _EXCHANGES[group] = ex
    return ex


def disable_peer_exchange(group=None) -> None:
    group = dist.group.WORLD if group is None else group
    ex = _EXCHANGES.pop(group, None)
    if ex is not None:
        ex.close()


class _VocabParallelLogprobEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, shard, labels, temperature, group, vocab_start, need_entropy, need_minmax):
        local = (labels - vocab_start).contiguous()
        mm = bool(need_minmax)
        partial = ops._raw_logprob_shard_fwd(shard, local, temperature, need_entropy, mm)
        ex = _EXCHANGES.get(group) if group is not None else None
        if ex is not None and partial.shape[0] <= ex.max_tokens and partial.shape[0] <= 262144:
            logp, ent, lse, rmax, rmin = ex.exchange_and_combine(partial, temperature, need_entropy, mm)
        else:
            logp, ent, lse, rmax, rmin = ops._raw_logprob_shard_combine(
                gather_partials(partial, group), temperature, need_entropy, mm)
        ctx.save_for_backward(shard, local, lse, ent)
        ctx.temperature = temperature
        ctx.need_entropy = need_entropy
        ctx.set_materialize_grads(False)
        if need_minmax == "shard":  # this rank's slice only, as `logits.min(-1)` on the shard
            rmax, rmin = partial[:, 4].contiguous(), partial[:, 5].contiguous()
        ctx.mark_non_differentiable(rmax, rmin)
        return logp, ent, rmax, rmin

    @staticmethod
    def backward(ctx, g_logp, g_ent, *unused):
        shard, local, lse, ent = ctx.saved_tensors
        if g_logp is None and (g_ent is None or not ctx.need_entropy):
            return (None,) * 7
        if g_logp is None:
            g_logp = torch.zeros_like(lse)
        g_ent = g_ent if ctx.need_entropy else None
        d = ops._raw_logprob_entropy_bwd(shard, local, lse, ent, g_logp, g_ent, ctx.temperature)
        return (d,) + (None,) * 6


def _resolve(shard: torch.Tensor, group, vocab_start):
    if group is None:
        group = tp_group_from_megatron()
    if vocab_start is None:
        vocab_start = 0
        if group is not None:
            vocab_start, _ = vocab_range(shard.shape[-1], dist.get_rank(group), dist.get_world_size(group))
    return group, int(vocab_start)


def vocab_parallel_logprobs_entropy(
    vocab_parallel_logits: torch.Tensor,
    labels: torch.Tensor,
    temperature: float = 1.0,
    group=None,
    vocab_start: Optional[int] = None,
    need_minmax=False,
):
    """``(logprobs, entropy)`` of the full-vocabulary softmax from this rank's ``[..., V/tp]``
    shard; ``labels`` are global ids. ``need_minmax``: True / "global" also returns the row
    max / min over the whole vocabulary, "shard" over this rank's slice (what the reference's
    ``logits.min(-1)`` on the shard gives)."""
    group, vocab_start = _resolve(vocab_parallel_logits, group, vocab_start)
    lead = labels.shape
    l2 = vocab_parallel_logits.reshape(-1, vocab_parallel_logits.shape[-1])
    logp, ent, rmax, rmin = _VocabParallelLogprobEntropy.apply(
        l2, labels.reshape(-1), float(temperature), group, vocab_start, True,
        "shard" if need_minmax == "shard" else bool(need_minmax))
    if need_minmax:
        return logp.view(lead), ent.view(lead), rmax.view(lead), rmin.view(lead)
    return logp.view(lead), ent.view(lead)


def vocab_parallel_logprobs(
    vocab_parallel_logits: torch.Tensor,
    labels: torch.Tensor,
    temperature: float = 1.0,
    group=None,
    vocab_start: Optional[int] = None,
) -> torch.Tensor:
    """``-vocab_parallel_cross_entropy(logits.float() / temperature, labels)``
    (ref megatron_actor.py:60-66)."""
    group, vocab_start = _resolve(vocab_parallel_logits, group, vocab_start)
    lead = labels.shape
    l2 = vocab_parallel_logits.reshape(-1, vocab_parallel_logits.shape[-1])
    logp, _, _, _ = _VocabParallelLogprobEntropy.apply(
        l2, labels.reshape(-1), float(temperature), group, vocab_start, False, False)
    return logp.view(lead)


def vocab_parallel_entropy(vocab_parallel_logits: torch.Tensor, group=None) -> torch.Tensor:
    """``_VocabParallelEntropy.apply(logits)`` (ref mcore/functional.py:8-52)."""
    group, _ = _resolve(vocab_parallel_logits, group, 0)
    lead = vocab_parallel_logits.shape[:-1]
    l2 = vocab_parallel_logits.reshape(-1, vocab_parallel_logits.shape[-1])
    none = torch.full((l2.shape[0],), -1, dtype=torch.int64, device=l2.device)
    _, ent, _, _ = _VocabParallelLogprobEntropy.apply(l2, none, 1.0, group, 0, True, False)
    return ent.view(lead)


def grpo_loss_fn(
    logits: torch.Tensor,
    input_data: Dict,
    temperature: float,
    eps_clip: float,
    c_clip: float | None,
    behav_imp_weight_cap: float | None,
    tp_group=None,
    stats: Optional[Dict] = None,
):
    """The Megatron actor's loss (ref megatron_actor.py:309-392): vocabulary-parallel when a TP
    group exists, the whole-vocabulary kernel otherwise. ``stats`` (extension) receives the
    logged tensors; inside AReaL they go to ``stats_tracker`` under the reference's keys."""
    from .actor import _register_stats, _stats_tracker

    input_ids = input_data["input_ids"]
    old_logp = input_data["logprobs"]
    advantages = input_data["advantages"]
    loss_mask = input_data["loss_mask"].bool()
    prox_logp = input_data["prox_logp"]
    labels = torch.roll(input_ids, shifts=-1, dims=-1)
    logprobs, entropy, vmax, vmin = vocab_parallel_logprobs_entropy(
        logits, labels, temperature, group=tp_group, need_minmax="shard")  # ref :372-373
    entropy = entropy.detach()
    loss, stat = ppo_actor_loss_fn(
        logprobs=logprobs, old_logprobs=old_logp, advantages=advantages, eps_clip=eps_clip,
        loss_mask=loss_mask, c_clip=c_clip, proximal_logprobs=prox_logp,
        behav_imp_weight_cap=behav_imp_weight_cap)
    st = _stats_tracker()
    if stats is None and st is None:
        return loss
    with torch.no_grad():
        clip_mask = stat["clip_mask"]
        dual = stat["dual_clip_mask"]
        collected = dict(
            n_valid_tokens=loss_mask, clipped_tokens=clip_mask, dual_clipped_tokens=dual,
            importance_weight=stat["importance_weight"], approx_kl=stat["approx_kl"],
            new_logp=logprobs.detach(), old_logp=old_logp, entropy=entropy.float(),
            actor_loss=stat["loss"], clip_ratio=clip_mask.float(), dual_clip_ratio=dual.float(),
            unclipped_behave_tokens=stat["behave_mask"], behave_imp_weight=stat["behave_imp_weight"],
            behave_approx_kl=stat["behave_approx_kl"], vocab_min_logits=vmin.detach(),
            vocab_max_logits=vmax.detach(),
            clipped_new_logp=torch.where(clip_mask, logprobs.detach(), 0),
            clipped_old_logp=torch.where(clip_mask, old_logp, 0),
        )
        if stats is not None:
            stats.update(collected)
        _register_stats(st, logits, collected)
    return loss
