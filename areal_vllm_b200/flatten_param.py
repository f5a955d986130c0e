"""Drop-in for the interval gather/scatter entry points of
``realhf/impl/model/nn/flatten_param.py:65-186`` and for the ``realhf._C.interval_op_cuda`` /
``realhf._C.interval_op`` bindings (csrc/interval_op/interval_op.cu:28-114, interval_op.cpp:4-25).

The reference falls back to python slicing below 1024 (slice) / 512 (set) intervals because its
kernel launches one thread per interval; the kernel here is balanced over bytes, so every size
goes through it. Results are bit-exact copies.
"""

from __future__ import annotations

import ctypes
from collections import OrderedDict
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import _lib, ops

__all__ = [
    "slice_intervals",
    "set_intervals",
    "merge_intervals",
    "slice_intervals_fp32",
    "slice_intervals_fp16",
    "slice_intervals_bf16",
    "set_intervals_fp32",
    "set_intervals_fp16",
    "set_intervals_bf16",
    "move_intervals",
    "clear_plan_cache",
]

_SUPPORTED = (torch.float32, torch.float16, torch.bfloat16)

# Plans (prefix sums of the interval lengths + tile->interval map, built on the device) are
# reused while the same intervals tensor is passed again, which is what the resharding loop of
# the reference does (intervals are uploaded once per plan: realhf/impl/model/comm/
# param_realloc.py:486-493). Key: storage pointer + version counter + shape + element size.
_PLAN_CACHE: "OrderedDict[tuple, Tuple[torch.Tensor, torch.Tensor]]" = OrderedDict()
_PLAN_CACHE_MAX = 32


def clear_plan_cache() -> None:
    _PLAN_CACHE.clear()


def _plan(intervals: torch.Tensor, packed_numel: int, elem_size: int, big_numel: int):
    """-> (workspace, reuse, key). A miss hands out a fresh workspace that is NOT yet in the cache:
    the caller registers it with ``_remember`` only after the op that builds the plan has returned
    without raising, so a failed call (strided tensor, misaligned pointer, launch error) can never
    leave an uninitialised plan behind for the next call to reuse."""
    key = (intervals.data_ptr(), intervals._version, intervals.shape[0], packed_numel, elem_size,
           big_numel, intervals.device.index)
    hit = _PLAN_CACHE.get(key)
    if hit is not None:
        _PLAN_CACHE.move_to_end(key)
        return hit[0], True, key
    nbytes = _lib.lib.b200rl_interval_workspace_bytes(intervals.shape[0], packed_numel, elem_size)
    return _lib.workspace(nbytes, intervals.device), False, key


def _remember(key, ws: torch.Tensor, keep_alive) -> None:
    """Cache a plan that a successful call has just built (``keep_alive``: the intervals tensors,
    so that their data pointers stay unique while the entry lives)."""
    _PLAN_CACHE[key] = (ws, keep_alive)
    while len(_PLAN_CACHE) > _PLAN_CACHE_MAX:
        _PLAN_CACHE.popitem(last=False)


def _to_cuda_intervals(intervals_cpu, device) -> torch.Tensor:
    arr = np.asarray(intervals_cpu, dtype=np.int64).reshape(-1, 2)
    return torch.from_numpy(arr).to(device, non_blocking=False)


def _slice_impl(src: torch.Tensor, intervals: torch.Tensor, total_size: int) -> torch.Tensor:
    ops._check_intervals(intervals)
    if intervals.shape[0] == 0 or total_size == 0:
        return torch.empty((total_size,), dtype=src.dtype, device=src.device)
    _lib.require_cuda(src=src)
    _lib.require_contiguous(src=src)  # all host-side validation before a plan is built or reused
    ws, reuse, key = _plan(intervals, total_size, src.element_size(), src.numel())
    out = ops.slice_intervals_op(src, intervals, int(total_size), ws, reuse)
    if not reuse:
        _remember(key, ws, intervals)
    return out


def _set_impl(src: torch.Tensor, dst: torch.Tensor, intervals: torch.Tensor) -> None:
    ops._check_intervals(intervals)
    if src.dtype != dst.dtype:
        raise RuntimeError("Source and destination tensors must have the same dtype")
    if intervals.shape[0] == 0 or src.numel() == 0:
        return None
    _lib.require_cuda(src=src, dst=dst)
    _lib.require_contiguous(src=src, dst=dst)
    ws, reuse, key = _plan(intervals, src.numel(), src.element_size(), dst.numel())
    ops.set_intervals_op(src, dst, intervals, ws, reuse)
    if not reuse:
        _remember(key, ws, intervals)
    return None


def slice_intervals(
    src: torch.Tensor,
    intervals_cpu: List[Tuple[int, int]] = None,
    intervals_cuda: torch.Tensor = None,
    output_size: int = None,
    max_interval_size: Optional[int] = None,
):
    """Concatenate ``src[s:e]`` for every interval, in the given order (ref flatten_param.py:97-143).

    ``max_interval_size`` only sized the reference's grid and is ignored. If only
    ``intervals_cpu`` is given it is uploaded; if ``output_size`` is missing it is computed from
    ``intervals_cpu`` (or, as a last resort, read back from the device)."""
    if not src.is_cuda:
        raise RuntimeError("src must be on CUDA (b200rl has no CPU path)")
    if src.dim() != 1:
        raise RuntimeError("src must be 1-D")
    if src.dtype not in _SUPPORTED:
        raise NotImplementedError(src.dtype)
    if intervals_cuda is None:
        if intervals_cpu is None:
            raise ValueError("either intervals_cpu or intervals_cuda must be given")
        intervals_cuda = _to_cuda_intervals(intervals_cpu, src.device)
    if output_size is None:
        if intervals_cpu is not None:
            arr = np.asarray(intervals_cpu, dtype=np.int64).reshape(-1, 2)
            output_size = int((arr[:, 1] - arr[:, 0]).sum())
        else:
            output_size = int((intervals_cuda[:, 1] - intervals_cuda[:, 0]).sum().item())
    return _slice_impl(src, intervals_cuda, int(output_size))


def set_intervals(
    src: torch.Tensor,
    dst: torch.Tensor,
    intervals_cpu: List[Tuple[int, int]] = None,
    intervals_cuda: torch.Tensor = None,
    max_interval_size: Optional[int] = None,
):
    """Scatter consecutive pieces of ``src`` into ``dst[s:e]``; in place, returns None
    (ref flatten_param.py:146-186)."""
    if not (src.is_cuda and dst.is_cuda):
        raise RuntimeError("src and dst must be on CUDA (b200rl has no CPU path)")
    if src.dim() != 1 or dst.dim() != 1:
        raise RuntimeError("src and dst must be 1-D")
    if src.dtype not in _SUPPORTED:
        raise NotImplementedError(src.dtype)
    if intervals_cuda is None:
        if intervals_cpu is None:
            raise ValueError("either intervals_cpu or intervals_cuda must be given")
        intervals_cuda = _to_cuda_intervals(intervals_cpu, src.device)
    return _set_impl(src, dst, intervals_cuda)


def merge_intervals(intervals: List[Tuple[int, int]]) -> List[Tuple[int, int]]:
    """Merge touching neighbours ``[a,b),[b,c) -> [a,c)`` in the order given, no sort
    (ref csrc/interval_op/interval_op.cpp:4-25). Host-side C++ in libb200rl.so."""
    n = len(intervals)
    if n == 0:
        return []
    arr = np.ascontiguousarray(np.asarray(intervals, dtype=np.uint64).reshape(n, 2))
    out = np.empty_like(arr)
    m = _lib.lib.b200rl_merge_intervals(
        arr.ctypes.data_as(ctypes.c_void_p), n, out.ctypes.data_as(ctypes.c_void_p)
    )
    return [(int(a), int(b)) for a, b in out[:m]]


# ---- the pybind names of realhf._C.interval_op_cuda (interval_op.cu:107-112) ----------------
def _typed_slice(dtype):
    def fn(src, intervals, total_size, max_interval_size):
        if src.dtype != dtype:
            raise RuntimeError(f"expected {dtype} tensor, got {src.dtype}")
        if not src.is_cuda:
            raise RuntimeError("src must be on CUDA")
        return _slice_impl(src, intervals, int(total_size))

    return fn


def _typed_set(dtype):
    def fn(src, dst, intervals, max_interval_size):
        if src.dtype != dtype:
            raise RuntimeError(f"expected {dtype} tensor, got {src.dtype}")
        return _set_impl(src, dst, intervals)

    return fn


slice_intervals_fp32 = _typed_slice(torch.float32)
slice_intervals_fp16 = _typed_slice(torch.float16)
slice_intervals_bf16 = _typed_slice(torch.bfloat16)
set_intervals_fp32 = _typed_set(torch.float32)
set_intervals_fp16 = _typed_set(torch.float16)
set_intervals_bf16 = _typed_set(torch.bfloat16)


def move_intervals(
    src: torch.Tensor,
    dst: torch.Tensor,
    src_intervals_cuda: torch.Tensor,
    dst_intervals_cuda: torch.Tensor,
    total_size: Optional[int] = None,
) -> None:
    """``set_intervals(slice_intervals(src, src_intervals), dst, dst_intervals)`` in one kernel
    and without the intermediate buffer (extension; the reference leaves it as a TODO at
    real_llm_api.py:719-720). Both interval lists must describe ``total_size`` elements in the
    same order; they may cut the stream differently. ``src`` may be a peer GPU's buffer opened
    with ``areal_vllm_b200.dist.open_peer_buffer`` - then this replaces the reference's
    slice -> chunked_bcast -> set sequence (real_llm_api.py:721-805) with one pass of NVLink
    loads; ordering against the owner of ``src`` stays with the caller. In place; returns None."""
    if src.dtype not in _SUPPORTED:
        raise NotImplementedError(src.dtype)
    ops._check_intervals(src_intervals_cuda)
    ops._check_intervals(dst_intervals_cuda)
    if total_size is None:
        total_size = int((src_intervals_cuda[:, 1] - src_intervals_cuda[:, 0]).sum().item())
    if total_size == 0:
        return None
    es = src.element_size()
    key = ("move", src_intervals_cuda.data_ptr(), src_intervals_cuda._version, src_intervals_cuda.shape[0],
           dst_intervals_cuda.data_ptr(), dst_intervals_cuda._version, dst_intervals_cuda.shape[0],
           total_size, es, src.numel(), dst.numel(), dst.device.index)
    hit = _PLAN_CACHE.get(key)
    if hit is not None:
        _PLAN_CACHE.move_to_end(key)
        ws, reuse = hit[0], True
    else:
        nbytes = _lib.lib.b200rl_move_workspace_bytes(src_intervals_cuda.shape[0], dst_intervals_cuda.shape[0],
                                                      total_size, es)
        ws, reuse = _lib.workspace(nbytes, dst.device), False
    ops.move_intervals_raw(src, dst, src_intervals_cuda, dst_intervals_cuda, total_size, ws, reuse)
    if not reuse:  # only a plan that a successful call has built may be reused
        _remember(key, ws, (src_intervals_cuda, dst_intervals_cuda))
    return None
