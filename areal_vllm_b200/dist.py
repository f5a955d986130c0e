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

__all__ = ["shard_groups_by_tokens", "allreduce_stats", "global_advantage_stats"]


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
