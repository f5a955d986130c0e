"""Data-parallel plumbing for the hot path: shard whole GRPO groups across ranks, and sum the few
global statistics with ONE small all-reduce.

The path shards by sequence (reference DP over sequences:
areal/engine/fsdp_engine.py:142-152). Group-level statistics never leave a rank because whole
groups are assigned to ranks. The only collectives are scalar sums — the three advantage moments
(ref areal/utils/functional.py:108-115 issues three all-reduces) and the loss normaliser
(ref areal/engine/fsdp_engine.py:491) — which are packed into a single fp64 vector here.
24-40 bytes over NVLink/NVSwitch is pure latency; NCCL through torch.distributed is the right
tool, a fused compute+collective kernel would buy nothing.
"""

from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

__all__ = ["shard_groups_by_tokens", "shard_sequences_by_tokens", "allreduce_stats",
           "global_advantage_stats", "sharded_group_advantage_norm"]


def shard_groups_by_tokens(
    seqlens: Sequence[int], group_size: int, world_size: int
) -> List[List[int]]:
    """Assign whole groups (``group_size`` adjacent sequences, as the reference assumes:
    areal/utils/functional.py:192) to ranks so that token counts are balanced.

    Longest-processing-time greedy: groups sorted by tokens descending, each placed on the
    currently lightest rank (reference analogue: allocate_balanced_mbs,
    areal/utils/data.py:219-243). Returns, per rank, the sorted list of SEQUENCE indices.
    Deterministic (ties broken by index), so every rank can compute it locally.
    """
    n = len(seqlens)
    if group_size <= 0 or n % group_size != 0:
        raise ValueError(f"{n} sequences are not a multiple of group_size={group_size}")
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    ngroups = n // group_size
    gtok = [sum(int(x) for x in seqlens[g * group_size:(g + 1) * group_size]) for g in range(ngroups)]
    order = sorted(range(ngroups), key=lambda g: (-gtok[g], g))
    load = [0] * world_size
    assign: List[List[int]] = [[] for _ in range(world_size)]
    for g in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        load[r] += gtok[g]
        assign[r].append(g)
    out = []
    for r in range(world_size):
        seqs: List[int] = []
        for g in sorted(assign[r]):
            seqs.extend(range(g * group_size, (g + 1) * group_size))
        out.append(seqs)
    return out


def shard_sequences_by_tokens(seqlens: Sequence[int], world_size: int) -> List[List[int]]:
    """Assign single sequences to ranks so that token counts are balanced (reference analogue:
    allocate_balanced_mbs, areal/utils/data.py:219-243, which balances by tokens at sequence
    granularity and knows nothing about GRPO groups).

    H1/H2 are per token and GAE is per sequence, so nothing on the path needs a GRPO group to sit
    on one rank; group-level advantage statistics are additive moments and are summed over the
    ranks by ``sharded_group_advantage_norm`` ([G, 3] fp64 in the step's one all-reduce). Whole
    groups of long sequences balance badly (15 groups on 8 ranks: max/mean 1.11); single sequences
    do not. Longest-processing-time greedy, then a pairwise-swap refinement; deterministic (ties
    by index), so every rank computes the same answer locally. Zero-length sequences are kept
    (assigned like any other) so the returned lists partition ``range(len(seqlens))``."""
    if world_size <= 0:
        raise ValueError("world_size must be positive")
    lens = [int(x) for x in seqlens]
    order = sorted(range(len(lens)), key=lambda i: (-lens[i], i))
    load = [0] * world_size
    assign: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        load[r] += lens[i]
        assign[r].append(i)
    # refinement: move or swap sequences between the heaviest and the lightest rank while that
    # lowers the maximum (LPT alone leaves ~ one shortest sequence of slack)
    for _ in range(4 * max(len(lens), 1)):
        hi = max(range(world_size), key=lambda k: (load[k], -k))
        lo = min(range(world_size), key=lambda k: (load[k], k))
        gap = load[hi] - load[lo]
        if gap <= 1:
            break
        best = None  # (new gap, i on hi, j on lo or None)
        for i in assign[hi]:
            if 0 < lens[i] < gap:
                cand = (abs(gap - 2 * lens[i]), i, None)
                best = cand if best is None or cand[:1] < best[:1] else best
            for j in assign[lo]:
                d = lens[i] - lens[j]
                if 0 < d < gap:
                    cand = (abs(gap - 2 * d), i, j)
                    best = cand if best is None or cand[:1] < best[:1] else best
        if best is None or best[0] >= gap:
            break
        _, i, j = best
        assign[hi].remove(i)
        assign[lo].append(i)
        load[hi] -= lens[i]
        load[lo] += lens[i]
        if j is not None:
            assign[lo].remove(j)
            assign[hi].append(j)
            load[lo] -= lens[j]
            load[hi] += lens[j]
    return [sorted(a) for a in assign]


def allreduce_stats(*tensors: torch.Tensor, group=None) -> Tuple[torch.Tensor, ...]:
    """Sum several small tensors over the data-parallel group with a single all-reduce.

    They are flattened into one fp64 vector (exact for counts, and what the reference uses for
    the advantage moments), reduced once, and split back. No-op without torch.distributed.
    """
    if not (dist.is_available() and dist.is_initialized()):
        return tuple(tensors)
    flat = torch.cat([t.reshape(-1).to(torch.float64) for t in tensors])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    out, off = [], 0
    for t in tensors:
        k = t.numel()
        out.append(flat[off:off + k].reshape(t.shape).to(t.dtype))
        off += k
    return tuple(out)


@torch.no_grad()
def global_advantage_stats(
    advantages: torch.Tensor,
    loss_mask: Optional[torch.Tensor],
    local_loss_weight: Optional[torch.Tensor] = None,
    eps: float = 1e-5,
    unbiased: bool = False,
    group=None,
):
    """Batch-level (mean, denom = std + eps) of the masked advantages over ALL ranks, and the
    global loss normaliser, with one all-reduce. Returns (mean[1], denom[1], total_loss_weight)."""
    from . import ops

    x = advantages.float().contiguous().view(-1)
    m = None if loss_mask is None else loss_mask.to(torch.float32).contiguous().view(-1)
    mom = ops.masked_moments(x, m, x.numel(), None, 1)
    if local_loss_weight is None:
        local_loss_weight = mom[0, 0:1].clone()
    mom, total_w = allreduce_stats(mom, local_loss_weight.reshape(-1).to(torch.float64), group=group)
    mean, denom = ops.moments_finalize(mom, False, bool(unbiased), float(eps))
    return mean, denom, total_w


@torch.no_grad()
def sharded_group_advantage_norm(
    advantages: torch.Tensor,
    loss_mask: Optional[torch.Tensor],
    row_group: torch.Tensor,
    n_groups: int,
    eps: float = 1e-5,
    unbiased: bool = False,
    group=None,
    extra: Sequence[torch.Tensor] = (),
    reduce=None,
):
    """AdvNorm(mean_level="group", std_level="group") when the rows of one GRPO group may live on
    DIFFERENT ranks (sequence-granular sharding, ``shard_sequences_by_tokens``).

    ``advantages`` / ``loss_mask``: this rank's padded ``[B_local, L]`` rows; ``row_group[b]`` is
    the GLOBAL group id of row b (int64, in ``[0, n_groups)``). The masked moments
    ``(sum m, sum x*m, sum (x*m)^2)`` of masked_normalization (ref areal/utils/functional.py:
    100-121) are additive, so each rank reduces its rows (one launch, fp64), adds them into a
    ``[n_groups, 3]`` table and the table is summed over the ranks in the SAME all-reduce as the
    ``extra`` tensors (loss normaliser etc.). Returns ``(normalised [B_local, L] fp32,
    reduced extras...)``; identical to ``AdvNorm`` on the gathered batch up to fp64 summation
    order. ``reduce`` (testing hook) replaces the all-reduce: ``reduce(table, *extra)`` ->
    tuple."""
    from . import ops

    x = advantages.float().contiguous()
    B, L = x.shape
    m = None if loss_mask is None else loss_mask.to(torch.float32).contiguous()
    table = torch.zeros(n_groups, 3, dtype=torch.float64, device=x.device)
    if x.numel():
        rows = ops.masked_moments(x.view(-1), None if m is None else m.view(-1), L, None, 1)
        table.index_add_(0, row_group.to(torch.int64), rows)
    red = (reduce or (lambda *ts: allreduce_stats(*ts, group=group)))(table, *extra)
    table = red[0]
    if x.numel() == 0:
        return (x,) + tuple(red[1:])
    mean, denom = ops.moments_finalize(table.contiguous(), False, bool(unbiased), float(eps))
    rg = row_group.to(torch.int64)
    out = ops.masked_normalize(x.view(-1), None if m is None else m.view(-1), mean[rg].contiguous(), L,
                               denom[rg].contiguous(), L, True).view(B, L)
    return (out,) + tuple(red[1:])


# ------------------------------------------------------------------------------------------------
# peer memory: let a kernel on this rank's GPU read another rank's flat parameter buffer
# ------------------------------------------------------------------------------------------------
class PeerBuffer:
    """A 1-D buffer that lives in ANOTHER process's GPU memory (same node), mapped into this
    process through CUDA IPC with this rank's GPU as the accessing device. Only what the
    interval kernels need: ``data_ptr() / numel() / dtype / element_size()``. Close it (or drop
    it) before the owner frees the memory."""

    def __init__(self, handle: bytes, offset_bytes: int, numel: int, dtype: torch.dtype,
                 device: torch.device):
        import ctypes

        from . import _lib

        self.dtype, self._numel, self.device = dtype, int(numel), torch.device(device)
        base = ctypes.c_void_p()
        buf = (ctypes.c_ubyte * 64).from_buffer_copy(handle)
        _lib.check(_lib.lib.b200rl_ipc_open(buf, self.device.index, ctypes.byref(base)), "b200rl_ipc_open")
        self._base = base.value
        self._ptr = self._base + int(offset_bytes)
        self.is_peer = True

    def data_ptr(self) -> int:
        return self._ptr

    def numel(self) -> int:
        return self._numel

    def element_size(self) -> int:
        return torch.empty((), dtype=self.dtype).element_size()

    def is_contiguous(self) -> bool:
        return True

    def close(self) -> None:
        if getattr(self, "_base", None):
            from . import _lib

            _lib.lib.b200rl_ipc_close(self._base, self.device.index)
            self._base = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def export_buffer(t: torch.Tensor) -> Tuple[bytes, int, int, torch.dtype]:
    """Picklable description of a contiguous CUDA tensor of THIS process for ``PeerBuffer``.
    The tensor must stay alive (and at the same address) while peers have it open. Works for
    ordinary caching-allocator memory (cudaMalloc blocks), not for expandable segments."""
    import ctypes

    from . import _lib

    if not (t.is_cuda and t.is_contiguous()):
        raise RuntimeError("only contiguous CUDA tensors can be exported")
    handle = (ctypes.c_ubyte * 64)()
    off = ctypes.c_int64()
    with torch.cuda.device(t.device):
        _lib.check(_lib.lib.b200rl_ipc_export(ctypes.c_void_p(t.data_ptr()), handle, ctypes.byref(off)),
                   "b200rl_ipc_export")
    return bytes(handle), int(off.value), t.numel(), t.dtype


def open_peer_buffer(t: Optional[torch.Tensor], src: int, group=None, device=None):
    """Collective over ``group``: rank ``src`` (a global rank) exports ``t``; every other rank
    returns a ``PeerBuffer`` aliasing it, readable by kernels on ``device`` (default: the current
    CUDA device). ``src`` itself gets ``t`` back."""
    me = dist.get_rank()
    box = [export_buffer(t) if me == src else None]
    dist.broadcast_object_list(box, src=src, group=group)
    if me == src:
        return t
    handle, off, numel, dtype = box[0]
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    return PeerBuffer(handle, off, numel, dtype, dev)
