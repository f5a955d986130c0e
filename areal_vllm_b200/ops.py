"""torch.library custom ops over the C ABI (the "thin torch custom-op layer").

Each op validates, allocates outputs/workspace with torch (plumbing only) and launches exactly
the C-ABI entry point on the current CUDA stream. Fake (meta) implementations are registered so
the ops trace under torch.compile; none of them synchronises with the host.

Two ways in, one implementation: the ``b200rl::*`` custom ops (``*_op`` below; what a compiled
graph sees) and, for eager mode, plain functions of the same names without the ``_op`` suffix that
call the same launchers directly — through ``torch.autograd.Function`` where a gradient exists —
because the dispatcher round trip of a python custom op costs 50-150 us per call, which is more
than the [T]-sized kernels themselves take.
"""

from __future__ import annotations

from typing import List, Optional, Tuple

import ctypes
from ctypes import c_void_p

import torch

from . import _lib
from ._lib import check, lib, ptr, stream_handle

Tensor = torch.Tensor


class _NullCtx:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


_NULL = _NullCtx()


def _on(dev: torch.device):
    """Context that makes `dev` current for the launch (free when it already is)."""
    return _NULL if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


def _eager() -> bool:
    return not torch.compiler.is_compiling()


# --------------------------------------------------------------------------------------------
# H1 log-prob / entropy
# --------------------------------------------------------------------------------------------
def _check_logits(logits: Tensor, labels: Tensor) -> None:
    _lib.require_cuda(logits=logits, labels=labels)
    if logits.dim() != 2:
        raise RuntimeError(f"logits must be [T, V], got {tuple(logits.shape)}")
    if labels.dim() != 1 or labels.shape[0] != logits.shape[0]:
        raise RuntimeError(f"labels must be [T]={logits.shape[0]}, got {tuple(labels.shape)}")
    if labels.dtype != torch.int64:
        raise RuntimeError(f"labels must be int64, got {labels.dtype}")
    if logits.dtype not in _lib.DTYPE_CODE:
        raise NotImplementedError(logits.dtype)
    if logits.stride(1) != 1:
        raise RuntimeError("logits rows must be contiguous (stride(-1) == 1)")
    if not labels.is_contiguous():
        raise RuntimeError("labels must be contiguous")


def _raw_logprob_entropy_fwd(
    logits: Tensor, labels: Tensor, temperature: float, need_entropy: bool, need_minmax: bool
) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """-> (logp[T], entropy[T] or empty, lse[T], rowmax[T] or empty, rowmin[T] or empty), all fp32."""
    _check_logits(logits, labels)
    T, V = logits.shape
    dev = logits.device
    f32 = dict(dtype=torch.float32, device=dev)
    logp = torch.empty(T, **f32)
    lse = torch.empty(T, **f32)
    ent = torch.empty(T if need_entropy else 0, **f32)
    rmax = torch.empty(T if need_minmax else 0, **f32)
    rmin = torch.empty(T if need_minmax else 0, **f32)
    with _on(dev):
        rc = lib.b200rl_logprob_entropy_fwd(
            ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
            ptr(labels), T, V, float(temperature), ptr(logp),
            ptr(ent) if need_entropy else None, ptr(lse),
            ptr(rmax) if need_minmax else None, ptr(rmin) if need_minmax else None,
            stream_handle(dev),
        )
    check(rc, "b200rl_logprob_entropy_fwd")
    return logp, ent, lse, rmax, rmin


logprob_entropy_fwd_op = torch.library.custom_op("b200rl::logprob_entropy_fwd", _raw_logprob_entropy_fwd, mutates_args=())


@logprob_entropy_fwd_op.register_fake
def _(logits, labels, temperature, need_entropy, need_minmax):
    T = logits.shape[0]
    mk = lambda n: logits.new_empty((n,), dtype=torch.float32)
    return mk(T), mk(T if need_entropy else 0), mk(T), mk(T if need_minmax else 0), mk(T if need_minmax else 0)


def _raw_logprob_entropy_bwd(
    logits: Tensor, labels: Tensor, lse: Tensor, entropy: Tensor, g_logp: Tensor,
    g_ent: Optional[Tensor], temperature: float,
) -> Tensor:
    """dlogits[T, V] in the logits dtype (what autograd returns for `logits.float()`)."""
    _check_logits(logits, labels)
    T, V = logits.shape
    dev = logits.device
    g_logp = g_logp.to(torch.float32).contiguous()
    if g_ent is not None:
        g_ent = g_ent.to(torch.float32).contiguous()
        if entropy.numel() != T:
            raise RuntimeError("entropy gradient given but the forward did not compute the entropy")
    dlogits = torch.empty((T, V), dtype=logits.dtype, device=dev)
    with _on(dev):
        rc = lib.b200rl_logprob_entropy_bwd(
            ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
            ptr(labels), ptr(lse), ptr(entropy) if g_ent is not None else None, ptr(g_logp),
            ptr(g_ent), T, V, float(temperature), ptr(dlogits), V, stream_handle(dev),
        )
    check(rc, "b200rl_logprob_entropy_bwd")
    return dlogits


logprob_entropy_bwd_op = torch.library.custom_op("b200rl::logprob_entropy_bwd", _raw_logprob_entropy_bwd, mutates_args=())


@logprob_entropy_bwd_op.register_fake
def _(logits, labels, lse, entropy, g_logp, g_ent, temperature):
    return logits.new_empty(logits.shape)


def _lp_setup_context(ctx, inputs, output):
    logits, labels, temperature, need_entropy, _ = inputs
    _, ent, lse, _, _ = output
    ctx.save_for_backward(logits, labels, lse, ent)
    ctx.temperature = temperature
    ctx.need_entropy = need_entropy
    # an unused entropy must arrive as None, not as zeros: without an entropy gradient the backward
    # takes the cheaper kernel variant (same choice as the eager-mode node below)
    ctx.set_materialize_grads(False)


def _lp_backward(ctx, g_logp, g_ent, g_lse, g_max, g_min):
    logits, labels, lse, ent = ctx.saved_tensors
    if g_logp is None:
        g_logp = torch.zeros_like(lse)
    use_ent = ctx.need_entropy and g_ent is not None
    dlogits = logprob_entropy_bwd_op(
        logits, labels, lse, ent, g_logp, g_ent if use_ent else None, ctx.temperature
    )
    return dlogits, None, None, None, None


logprob_entropy_fwd_op.register_autograd(_lp_backward, setup_context=_lp_setup_context)


class _LogprobFn(torch.autograd.Function):
    """Eager-mode autograd node: same launchers as the custom op, no dispatcher round trip."""

    @staticmethod
    def forward(ctx, logits, labels, temperature, need_entropy, need_minmax):
        out = _raw_logprob_entropy_fwd(logits, labels, temperature, need_entropy, need_minmax)
        ctx.save_for_backward(logits, labels, out[2], out[1])
        ctx.temperature = temperature
        ctx.need_entropy = need_entropy
        ctx.set_materialize_grads(False)  # an unused entropy must arrive as None, not zeros
        ctx.mark_non_differentiable(out[2], out[3], out[4])
        return out

    @staticmethod
    def backward(ctx, g_logp, g_ent, g_lse, g_max, g_min):
        logits, labels, lse, ent = ctx.saved_tensors
        if g_logp is None:
            g_logp = torch.zeros_like(lse)
        use_ent = ctx.need_entropy and g_ent is not None
        dlogits = _raw_logprob_entropy_bwd(
            logits, labels, lse, ent, g_logp, g_ent if use_ent else None, ctx.temperature
        )
        return dlogits, None, None, None, None


def logprob_entropy_fwd(logits, labels, temperature, need_entropy, need_minmax):
    if not _eager():
        return logprob_entropy_fwd_op(logits, labels, temperature, need_entropy, need_minmax)
    if logits.requires_grad and torch.is_grad_enabled():
        return _LogprobFn.apply(logits, labels, temperature, need_entropy, need_minmax)
    return _raw_logprob_entropy_fwd(logits, labels, temperature, need_entropy, need_minmax)


def logprob_entropy_bwd(logits, labels, lse, entropy, g_logp, g_ent, temperature):
    if not _eager():
        return logprob_entropy_bwd_op(logits, labels, lse, entropy, g_logp, g_ent, temperature)
    return _raw_logprob_entropy_bwd(logits, labels, lse, entropy, g_logp, g_ent, temperature)


# --------------------------------------------------------------------------------------------
# H1 over a vocabulary-parallel shard: partial records + combine
# --------------------------------------------------------------------------------------------
PARTIAL_STRIDE = 8  # floats per row: (M, S, W, x_y, max, min, 0, 0)


def _raw_logprob_shard_fwd(logits: Tensor, local_labels: Tensor, temperature: float,
                           need_entropy: bool, need_minmax: bool) -> Tensor:
    """-> partial [T, 8] fp32 for this rank's [T, V_shard] slice (labels already shifted by the
    shard's first column; out-of-shard labels are fine)."""
    _check_logits(logits, local_labels)
    T, V = logits.shape
    dev = logits.device
    partial = torch.empty(T, PARTIAL_STRIDE, dtype=torch.float32, device=dev)
    with _on(dev):
        rc = lib.b200rl_logprob_shard_fwd(
            ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
            ptr(local_labels), T, V, float(temperature), int(need_entropy), int(need_minmax),
            ptr(partial), stream_handle(dev))
    check(rc, "b200rl_logprob_shard_fwd")
    return partial


def _raw_logprob_shard_combine(partials: Tensor, temperature: float, need_entropy: bool,
                               need_minmax: bool) -> Tuple[Tensor, Tensor, Tensor, Tensor, Tensor]:
    """partials [n_shards, T, 8] -> (logp, entropy or empty, lse, rowmax or empty, rowmin or empty)."""
    _lib.require_cuda(partials=partials)
    if partials.dim() != 3 or partials.shape[2] != PARTIAL_STRIDE or partials.dtype != torch.float32:
        raise RuntimeError(f"partials must be fp32 [n_shards, T, {PARTIAL_STRIDE}], got {tuple(partials.shape)}")
    partials = partials.contiguous()
    n, T, _ = partials.shape
    dev = partials.device
    f32 = dict(dtype=torch.float32, device=dev)
    logp, lse = torch.empty(T, **f32), torch.empty(T, **f32)
    ent = torch.empty(T if need_entropy else 0, **f32)
    rmax = torch.empty(T if need_minmax else 0, **f32)
    rmin = torch.empty(T if need_minmax else 0, **f32)
    with _on(dev):
        rc = lib.b200rl_logprob_shard_combine(
            ptr(partials), n, T, float(temperature), ptr(logp), ptr(ent) if need_entropy else None,
            ptr(lse), ptr(rmax) if need_minmax else None, ptr(rmin) if need_minmax else None,
            stream_handle(dev))
    check(rc, "b200rl_logprob_shard_combine")
    return logp, ent, lse, rmax, rmin


# --------------------------------------------------------------------------------------------
# H1+H2 fused GRPO step (eager only: it allocates and fills dlogits in the forward)
# --------------------------------------------------------------------------------------------
def _raw_grpo_fused_step(
    logits: Tensor, labels: Tensor, prox_logp: Tensor, old_logp: Tensor, adv: Tensor, mask: Tensor,
    temperature: float, clip_lo: float, clip_hi: float, c_clip: float, cap: float,
    upstream: Optional[Tensor], need_minmax: bool,
):
    """-> dict of outputs, or None when the kernel does not support the layout (unaligned rows)."""
    _check_logits(logits, labels)
    T, V = logits.shape
    dev = logits.device
    pv = 4 if logits.dtype == torch.float32 else 8
    if V % pv or logits.stride(0) % pv or logits.data_ptr() % 16:
        return None
    f32 = dict(dtype=torch.float32, device=dev)
    u8 = dict(dtype=torch.uint8, device=dev)

    def tokvec(t: Tensor, name: str) -> Tensor:
        if t.shape != (T,):
            raise RuntimeError(f"{name} must have shape ({T},), got {tuple(t.shape)}")
        return t.detach().to(torch.float32).contiguous()

    px, ol, ad = tokvec(prox_logp, "prox_logp"), tokvec(old_logp, "logprobs"), tokvec(adv, "advantages")
    if mask.shape != (T,):
        raise RuntimeError(f"loss_mask must have shape ({T},)")
    mk = (mask if mask.dtype == torch.bool else mask != 0).contiguous().view(torch.uint8)
    up = None if upstream is None else upstream.detach().to(torch.float32).reshape(1).contiguous()
    o = dict(
        logp=torch.empty(T, **f32), entropy=torch.empty(T, **f32), lse=torch.empty(T, **f32),
        rowmax=torch.empty(T if need_minmax else 0, **f32), rowmin=torch.empty(T if need_minmax else 0, **f32),
        s_loss=torch.empty(T, **f32), s_iw=torch.empty(T, **f32), s_kl=torch.empty(T, **f32),
        s_clip=torch.empty(T, **u8), s_dual=torch.empty(T, **u8), s_bw=torch.empty(T, **f32),
        s_bkl=torch.empty(T, **f32), s_bm=torch.empty(T, **u8),
        dlogits=torch.empty((T, V), dtype=logits.dtype, device=dev),
        loss=torch.empty((), **f32), denom=torch.empty(1, **f32),
    )
    ws = _lib.workspace(lib.b200rl_grpo_fused_workspace_bytes(T), dev)
    with _on(dev):
        rc = lib.b200rl_grpo_fused_step(
            ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
            ptr(labels), T, V, float(temperature), ptr(px), ptr(ol), ptr(ad), ptr(mk),
            float(clip_lo), float(clip_hi), float(c_clip), float(cap), ptr(up),
            ptr(o["logp"]), ptr(o["entropy"]), ptr(o["lse"]),
            ptr(o["rowmax"]) if need_minmax else None, ptr(o["rowmin"]) if need_minmax else None,
            ptr(o["s_loss"]), ptr(o["s_iw"]), ptr(o["s_kl"]), ptr(o["s_clip"]), ptr(o["s_dual"]),
            ptr(o["s_bw"]), ptr(o["s_bkl"]), ptr(o["s_bm"]), ptr(o["dlogits"]), V,
            ptr(o["loss"]), ptr(o["denom"]), ptr(ws), ws.numel(), stream_handle(dev),
        )
    if rc == -2:  # B200RL_EUNSUPPORTED
        return None
    check(rc, "b200rl_grpo_fused_step")
    return o


def _raw_grpo_fused_step_ids(
    logits: Tensor, input_ids: Tensor, roll: bool, prox_logp: Tensor, old_logp: Tensor, adv: Tensor,
    mask: Tensor, temperature: float, clip_lo: float, clip_hi: float, c_clip: float, cap: float,
    upstream: Optional[Tensor], want_stats: bool,
):
    """The fused step with the engine-facing conveniences (label roll inside the kernel, fp32 mask
    read as is, logged tensors written by the kernel) and slab outputs: one host call, two
    launches (mask count + the fused kernel), five allocations.
    -> (loss 0-d, dlogits, {row name: [T] view}) or None when the layout is not supported."""
    _check_logits(logits, input_ids)
    T, V = logits.shape
    dev = logits.device
    pv = 4 if logits.dtype == torch.float32 else 8
    if V % pv or logits.stride(0) % pv or logits.data_ptr() % 16:
        return None
    px, ol, ad = _tokvec(prox_logp.detach(), T, "prox_logp"), _tokvec(old_logp.detach(), T, "logprobs"), \
        _tokvec(adv.detach(), T, "advantages")
    if mask.shape != (T,) or not mask.is_cuda:
        raise RuntimeError(f"loss_mask must be a CUDA tensor of shape ({T},)")
    if mask.dtype == torch.float32:
        mk, mf = mask, 1
    else:
        mk, mf = (mask if mask.dtype == torch.bool else mask != 0), 0
    if not mk.is_contiguous():
        mk = mk.contiguous()
    up = None if upstream is None else upstream.detach().to(torch.float32).reshape(1)
    Tp = (T + 63) // 64 * 64
    names_f = [n for n in GL_F32 if n != "dlogp"]  # logp, lse, entropy, rowmax, rowmin, 5 stats, 4 derived
    fs = torch.empty((len(names_f), Tp), dtype=torch.float32, device=dev)
    bs = torch.empty((len(GL_U8), Tp), dtype=torch.bool, device=dev)
    sc = torch.empty(2, dtype=torch.float32, device=dev)
    dlogits = torch.empty((T, V), dtype=logits.dtype, device=dev)
    ws = torch.empty(max(lib.b200rl_grpo_fused_workspace_bytes(T), 256), dtype=torch.uint8, device=dev)
    fb, st4, bb, sp = fs.data_ptr(), Tp * 4, bs.data_ptr(), sc.data_ptr()
    idx = {n: i for i, n in enumerate(names_f)}
    f = lambda n: c_void_p(fb + idx[n] * st4) if (want_stats or n in ("logp", "lse", "entropy")) else None
    b = (lambda i: c_void_p(bb + i * Tp)) if want_stats else (lambda i: None)
    with _on(dev):
        rc = lib.b200rl_grpo_fused_step_ids(
            ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
            ptr(input_ids), int(roll), T, V, float(temperature), ptr(px), ptr(ol), ptr(ad), ptr(mk), mf,
            float(clip_lo), float(clip_hi), float(c_clip), float(cap), ptr(up),
            f("logp"), f("entropy"), f("lse"), f("rowmax"), f("rowmin"),
            f("s_loss"), f("s_iw"), f("s_kl"), b(0), b(1), f("s_bw"), f("s_bkl"), b(2),
            b(3), f("clip_ratio"), f("dual_clip_ratio"), f("clipped_new_logp"), f("clipped_old_logp"),
            ptr(dlogits), V, c_void_p(sp), c_void_p(sp + 4), ptr(ws), ws.numel(), stream_handle(dev))
    if rc == -2:  # B200RL_EUNSUPPORTED
        return None
    check(rc, "b200rl_grpo_fused_step_ids")
    rows = None
    if want_stats:
        fv, bv = (fs, bs) if Tp == T else (fs[:, :T], bs[:, :T])
        rows = dict(zip(names_f, fv.unbind(0)))
        rows.update(zip(GL_U8, bv.unbind(0)))
    return sc[0], dlogits, rows, up


def scale_inplace_(buf: Tensor, scale: Tensor) -> Tensor:
    """buf *= scale (device scalar); launches a kernel that exits at once when scale == 1."""
    _lib.require_cuda(buf=buf, scale=scale)
    _lib.require_contiguous(buf=buf)
    dev = buf.device
    sc = scale.detach().to(torch.float32).reshape(1).contiguous()
    with _on(dev):
        rc = lib.b200rl_scale_inplace(ptr(buf), _lib.DTYPE_CODE[buf.dtype], buf.numel(), ptr(sc), stream_handle(dev))
    check(rc, "b200rl_scale_inplace")
    return buf


class _GrpoFusedFn(torch.autograd.Function):
    """loss = upstream * PPO loss; the gradient w.r.t. logits was already written by the forward."""

    @staticmethod
    def forward(ctx, logits, loss_unscaled, dlogits, upstream):
        ctx.save_for_backward(dlogits)
        ctx.set_materialize_grads(False)
        ctx.consumed = False
        return loss_unscaled * upstream.reshape(()) if upstream is not None else loss_unscaled.clone()

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        if g is None:
            return None, None, None, None
        if ctx.consumed:
            # the stored gradient is handed out (and rescaled) IN PLACE; a second pass through this
            # node would scale it again and may alias what the first pass returned as logits.grad
            raise RuntimeError(
                "the fused GRPO step's gradient was already consumed by a previous backward(); "
                "retain_graph / a second backward through this node is not supported - use "
                "grpo_loss_fn (the two-kernel path) for that")
        ctx.consumed = True
        return scale_inplace_(dlogits, g), None, None, None


# --------------------------------------------------------------------------------------------
# H2-d composite: grpo_loss_fn as two launches forward, one launch backward (eager only)
# --------------------------------------------------------------------------------------------
# rows of the fp32 slab / of the bool slab that one call of b200rl_grpo_loss_fwd fills
GL_F32 = ("logp", "lse", "dlogp", "entropy", "rowmax", "rowmin", "s_loss", "s_iw", "s_kl", "s_bw", "s_bkl",
          "clip_ratio", "dual_clip_ratio", "clipped_new_logp", "clipped_old_logp")
GL_U8 = ("s_clip", "s_dual", "s_bm", "valid")
_GL_CORE = 3  # without stats only logp, lse, dlogp exist


def _tokvec(t: Tensor, T: int, name: str) -> Tensor:
    if t.shape != (T,):
        raise RuntimeError(f"{name} must have shape ({T},), got {tuple(t.shape)}")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be on CUDA (b200rl has no CPU path)")
    if t.dtype != torch.float32:
        t = t.to(torch.float32)
    return t if t.is_contiguous() else t.contiguous()


def _raw_grpo_loss_fwd(logits: Tensor, input_ids: Tensor, roll: bool, prox_logp: Tensor, old_logp: Tensor,
                       adv: Tensor, mask: Tensor, temperature: float, clip_lo: float, clip_hi: float,
                       c_clip: float, cap: float, want_stats: bool):
    """-> (scalars [2] fp32 = (loss, N), fp32 slab [15 or 3, Tp], bool slab [4, Tp] or None).
    One host call, two launches; every output of the step lives in the two slabs (rows GL_F32 /
    GL_U8, row stride Tp = T rounded up to 64) so the call allocates four tensors, not twenty."""
    _check_logits(logits, input_ids)
    T, V = logits.shape
    dev = logits.device
    px, ol, ad = _tokvec(prox_logp.detach(), T, "prox_logp"), _tokvec(old_logp.detach(), T, "logprobs"), \
        _tokvec(adv.detach(), T, "advantages")
    if mask.shape != (T,) or not mask.is_cuda:
        raise RuntimeError(f"loss_mask must be a CUDA tensor of shape ({T},)")
    if mask.dtype == torch.float32:
        mk, mf = mask, 1
    else:
        mk, mf = (mask if mask.dtype == torch.bool else mask != 0), 0
    if not mk.is_contiguous():
        mk = mk.contiguous()
    Tp = (T + 63) // 64 * 64
    nf = len(GL_F32) if want_stats else _GL_CORE
    fs = torch.empty((nf, Tp), dtype=torch.float32, device=dev)
    bs = torch.empty((len(GL_U8), Tp), dtype=torch.bool, device=dev) if want_stats else None
    sc = torch.empty(2, dtype=torch.float32, device=dev)
    ws = torch.empty(max(lib.b200rl_grpo_loss_workspace_bytes(T), 256), dtype=torch.uint8, device=dev)
    fb, st4 = fs.data_ptr(), Tp * 4
    f = lambda i: c_void_p(fb + i * st4) if i < nf else None
    bb = bs.data_ptr() if want_stats else 0
    b = (lambda i: c_void_p(bb + i * Tp)) if want_stats else (lambda i: None)
    sp = sc.data_ptr()
    with _on(dev):
        rc = lib.b200rl_grpo_loss_fwd(
            ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
            ptr(input_ids), int(roll), T, V, float(temperature), ptr(px), ptr(ol), ptr(ad), ptr(mk), mf,
            float(clip_lo), float(clip_hi), float(c_clip), float(cap),
            f(0), f(3), f(1), f(4), f(5),                      # logp, entropy, lse, rowmax, rowmin
            f(6), f(7), f(8), b(0), b(1), f(9), f(10), b(2),   # the reference's eight stats
            b(3), f(11), f(12), f(13), f(14),                  # valid, ratios, clipped logps
            f(2), c_void_p(sp), c_void_p(sp + 4), ptr(ws), ws.numel(), stream_handle(dev))
    check(rc, "b200rl_grpo_loss_fwd")
    return sc, fs, bs


class _GrpoLossFn(torch.autograd.Function):
    """loss = grpo_loss_fn(logits, ...); backward = one launch producing dlogits."""

    @staticmethod
    def forward(ctx, logits, input_ids, roll, prox_logp, old_logp, adv, mask, temperature, clip_lo, clip_hi,
                c_clip, cap, want_stats):
        sc, fs, bs = _raw_grpo_loss_fwd(logits, input_ids, roll, prox_logp, old_logp, adv, mask, temperature,
                                        clip_lo, clip_hi, c_clip, cap, want_stats)
        ctx.save_for_backward(logits, input_ids, fs, sc)
        ctx.roll, ctx.temperature = bool(roll), float(temperature)
        ctx.set_materialize_grads(False)
        loss = sc[0]
        if bs is None:
            ctx.mark_non_differentiable(fs)
            return loss, fs
        ctx.mark_non_differentiable(fs, bs)
        return loss, fs, bs

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 13
        logits, input_ids, fs, sc = ctx.saved_tensors
        T, V = logits.shape
        dev = logits.device
        gl = g_loss if g_loss.dtype == torch.float32 else g_loss.to(torch.float32)
        dlogits = torch.empty((T, V), dtype=logits.dtype, device=dev)
        fb, st4, sp = fs.data_ptr(), fs.stride(0) * 4, sc.data_ptr()
        with _on(dev):
            rc = lib.b200rl_grpo_loss_bwd(
                ptr(logits), _lib.DTYPE_CODE[logits.dtype], logits.stride(0) if T > 1 else max(V, logits.stride(0)),
                ptr(input_ids), int(ctx.roll), T, V, ctx.temperature, c_void_p(fb + st4), c_void_p(fb + 2 * st4),
                ptr(gl), c_void_p(sp + 4), ptr(dlogits), V, stream_handle(dev))
        check(rc, "b200rl_grpo_loss_bwd")
        return (dlogits,) + (None,) * 12


def grpo_loss(logits, input_ids, roll, prox_logp, old_logp, adv, mask, temperature, clip_lo, clip_hi, c_clip,
              cap, want_stats):
    """-> (loss, {name: [T] view}) ; eager mode only (the caller keeps the custom-op path for compiled
    graphs). The dict has the GL_F32 / GL_U8 rows (three of them without stats)."""
    if logits.requires_grad and torch.is_grad_enabled():
        out = _GrpoLossFn.apply(logits, input_ids, roll, prox_logp, old_logp, adv, mask, temperature, clip_lo,
                                clip_hi, c_clip, cap, want_stats)
        loss, fs, bs = out if want_stats else (out[0], out[1], None)
    else:
        sc, fs, bs = _raw_grpo_loss_fwd(logits, input_ids, roll, prox_logp, old_logp, adv, mask, temperature,
                                        clip_lo, clip_hi, c_clip, cap, want_stats)
        loss = sc[0]
    T = logits.shape[0]
    if fs.shape[1] != T:
        fs = fs[:, :T]
        bs = None if bs is None else bs[:, :T]
    rows = dict(zip(GL_F32, fs.unbind(0)))
    if bs is not None:
        rows.update(zip(GL_U8, bs.unbind(0)))
    return loss, rows


# --------------------------------------------------------------------------------------------
# H3 GAE
# --------------------------------------------------------------------------------------------
def _raw_gae_1d_nolp_misalign(
    rewards: Tensor, values: Tensor, cu_seqlens: Tensor, bootstrap: Tensor, gamma: float,
    lam: float, mode: int,
) -> Tuple[Tensor, Tensor]:
    return _gae_launch(rewards, values, cu_seqlens, bootstrap, gamma, lam, mode, False)


def _gae_launch(rewards: Tensor, values: Tensor, cu_seqlens: Tensor, bootstrap: Tensor, gamma: float,
                lam: float, mode: int, one_allocation: bool) -> Tuple[Tensor, Tensor]:
    # checks of csrc/cugae/gae.cu:36-50, minus the two .item() host syncs on cu_seqlens. Written out
    # inline and in this order because at BASELINE config 1 (256 sequences) the kernels take ~5 us
    # and everything the host does before the launch is on the measured path.
    if not (rewards.is_cuda and values.is_cuda and cu_seqlens.is_cuda and bootstrap.is_cuda):
        _lib.require_cuda(rewards=rewards, values=values, cu_seqlens=cu_seqlens, bootstrap=bootstrap)
    if not (rewards.is_contiguous() and values.is_contiguous() and cu_seqlens.is_contiguous()
            and bootstrap.is_contiguous()):
        _lib.require_contiguous(rewards=rewards, values=values, cu_seqlens=cu_seqlens, bootstrap=bootstrap)
    if not (rewards.dim() == values.dim() == cu_seqlens.dim() == 1):
        raise RuntimeError("rewards, values and cu_seqlens must be 1-D")
    batch = cu_seqlens.shape[0] - 1
    total = rewards.shape[0]
    if bootstrap.shape != (batch,):
        raise RuntimeError(f"bootstrap must have shape ({batch},)")
    if values.shape != (total + batch,):
        raise RuntimeError(f"values must have shape ({total + batch},)")
    if bootstrap.dtype != torch.bool:
        raise RuntimeError("bootstrap must be bool")
    if cu_seqlens.dtype != torch.int32:
        raise RuntimeError("cu_seqlens must be int32")
    if rewards.dtype != torch.float32 or values.dtype != torch.float32:
        raise RuntimeError("rewards and values must be float32")
    dev = rewards.device
    if batch <= 0 or total == 0:
        return torch.zeros_like(rewards), torch.zeros_like(rewards)
    # The reference zero-fills its outputs (gae.cu:53-54) and then asserts that cu_seqlens covers
    # [0, total) (gae.cu:48-49), so every element is overwritten; zero-filling 2 x 4 B/token first
    # would add 50% to the kernel's HBM traffic. Tokens outside [cu[0], cu[-1]) are undefined here.
    # Eager mode: adv and ret are the two rows of ONE allocation (row pitch a multiple of 16 bytes,
    # which the vectorised SCAN path wants); the custom op (compiled graphs) must not return
    # tensors that share storage and allocates them separately. Workspace size cached per shape.
    if one_allocation:
        Tp = (total + 3) & ~3
        out = torch.empty((2, Tp), dtype=torch.float32, device=dev)
        adv, ret = (out[0], out[1]) if Tp == total else (out[0, :total], out[1, :total])
        p_adv, p_ret = c_void_p(out.data_ptr()), c_void_p(out.data_ptr() + Tp * 4)
    else:
        adv, ret = torch.empty_like(rewards), torch.empty_like(rewards)
        p_adv, p_ret = ptr(adv), ptr(ret)
    key = (batch, total)
    ws_bytes = _GAE_WS_BYTES.get(key)
    if ws_bytes is None:
        if len(_GAE_WS_BYTES) > 4096:
            _GAE_WS_BYTES.clear()
        ws_bytes = _GAE_WS_BYTES[key] = max(int(lib.b200rl_gae_workspace_bytes(batch, total)), 256)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    with _on(dev):
        rc = lib.b200rl_gae_1d_nolp_misalign(
            ptr(rewards), ptr(values), ptr(cu_seqlens), ptr(bootstrap), p_adv, p_ret,
            batch, total, float(gamma), float(lam), int(mode), ptr(ws), ws_bytes,
            stream_handle(dev),
        )
    check(rc, "b200rl_gae_1d_nolp_misalign")
    return adv, ret


_GAE_WS_BYTES: dict = {}


gae_1d_nolp_misalign_op = torch.library.custom_op("b200rl::gae_1d_nolp_misalign", _raw_gae_1d_nolp_misalign, mutates_args=())


@gae_1d_nolp_misalign_op.register_fake
def _(rewards, values, cu_seqlens, bootstrap, gamma, lam, mode):
    return torch.empty_like(rewards), torch.empty_like(rewards)


def gae_1d_nolp_misalign(rewards, values, cu_seqlens, bootstrap, gamma, lam, mode):
    if _eager():
        return _gae_launch(rewards, values, cu_seqlens, bootstrap, gamma, lam, mode, True)
    return gae_1d_nolp_misalign_op(rewards, values, cu_seqlens, bootstrap, gamma, lam, mode)


# --------------------------------------------------------------------------------------------
# H2 PPO loss
# --------------------------------------------------------------------------------------------
def _raw_ppo_loss_fwd(
    logp: Tensor, prox_logp: Tensor, old_logp: Tensor, adv: Tensor, mask: Tensor,
    clip_lo: float, clip_hi: float, c_clip: float, cap: float,
    adv_mean: Optional[Tensor], adv_denom: Optional[Tensor], adv_group: Optional[Tensor],
) -> List[Tensor]:
    """-> [loss(0-d), denom(1), dlogp, adv_used, stat_loss, imp_w, approx_kl, clip_mask(u8),
    dual_clip_mask(u8), behave_imp_weight, behave_approx_kl, behave_mask(u8)]"""
    _lib.require_cuda(logp=logp, prox_logp=prox_logp, old_logp=old_logp, adv=adv, mask=mask)
    shape = logp.shape
    dev = logp.device

    def flat(t: Tensor, name: str) -> Tensor:
        if t.shape != shape:
            raise RuntimeError(f"{name} has shape {tuple(t.shape)}, expected {tuple(shape)}")
        if t.dtype != torch.float32:
            raise RuntimeError(f"{name} must be float32, got {t.dtype}")
        return t.detach().contiguous().view(-1)

    lp, px, ol, ad = flat(logp, "logprobs"), flat(prox_logp, "proximal_logprobs"), \
        flat(old_logp, "old_logprobs"), flat(adv, "advantages")
    if mask.shape != shape:
        raise RuntimeError(f"loss_mask has shape {tuple(mask.shape)}, expected {tuple(shape)}")
    mk = (mask if mask.dtype == torch.bool else mask != 0).contiguous().view(-1).view(torch.uint8)
    T = lp.numel()
    f32 = dict(dtype=torch.float32, device=dev)
    u8 = dict(dtype=torch.uint8, device=dev)
    loss = torch.empty((), **f32)
    denom = torch.empty(1, **f32)
    dlogp = torch.empty(T, **f32)
    fused = adv_mean is not None or adv_denom is not None
    adv_used = torch.empty(T if fused else 0, **f32)
    s_loss, s_iw, s_kl = torch.empty(T, **f32), torch.empty(T, **f32), torch.empty(T, **f32)
    s_clip, s_dual = torch.empty(T, **u8), torch.empty(T, **u8)
    s_bw, s_bkl, s_bm = torch.empty(T, **f32), torch.empty(T, **f32), torch.empty(T, **u8)
    if adv_mean is not None:
        adv_mean = adv_mean.to(torch.float64).contiguous()
    if adv_denom is not None:
        adv_denom = adv_denom.to(torch.float64).contiguous()
    if adv_group is not None:
        adv_group = adv_group.to(torch.int32).contiguous().view(-1)
        if adv_group.numel() != T:
            raise RuntimeError("adv_group must have one entry per token")
    ws = _lib.workspace(lib.b200rl_ppo_loss_workspace_bytes(T), dev)
    with _on(dev):
        rc = lib.b200rl_ppo_loss_fwd(
            ptr(lp), ptr(px), ptr(ol), ptr(ad), ptr(mk), T, float(clip_lo), float(clip_hi),
            float(c_clip), float(cap), ptr(adv_mean), ptr(adv_denom), ptr(adv_group),
            ptr(adv_used) if fused else None,
            ptr(s_loss), ptr(s_iw), ptr(s_kl), ptr(s_clip), ptr(s_dual), ptr(s_bw), ptr(s_bkl),
            ptr(s_bm), ptr(dlogp), ptr(loss), ptr(denom), ptr(ws), ws.numel(), stream_handle(dev),
        )
    check(rc, "b200rl_ppo_loss_fwd")
    rs = lambda t: t.view(shape)
    return [loss, denom, rs(dlogp), rs(adv_used) if fused else adv_used, rs(s_loss), rs(s_iw),
            rs(s_kl), rs(s_clip), rs(s_dual), rs(s_bw), rs(s_bkl), rs(s_bm)]


ppo_loss_fwd_op = torch.library.custom_op("b200rl::ppo_loss_fwd", _raw_ppo_loss_fwd, mutates_args=())


@ppo_loss_fwd_op.register_fake
def _(logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap, adv_mean, adv_denom, adv_group):
    f = lambda: torch.empty_like(logp, dtype=torch.float32)
    b = lambda: torch.empty_like(logp, dtype=torch.uint8)
    fused = adv_mean is not None or adv_denom is not None
    return [logp.new_empty((), dtype=torch.float32), logp.new_empty((1,), dtype=torch.float32), f(),
            f() if fused else logp.new_empty((0,), dtype=torch.float32), f(), f(), f(), b(), b(),
            f(), f(), b()]


def _raw_ppo_loss_bwd(dlogp: Tensor, grad_loss: Tensor, denom: Tensor) -> Tensor:
    _lib.require_cuda(dlogp=dlogp, grad_loss=grad_loss, denom=denom)
    dev = dlogp.device
    d = dlogp.contiguous()
    g = torch.empty_like(d)
    gl = grad_loss.to(torch.float32).reshape(1).contiguous()
    with _on(dev):
        rc = lib.b200rl_ppo_loss_bwd(ptr(d), ptr(gl), ptr(denom), ptr(g), d.numel(), stream_handle(dev))
    check(rc, "b200rl_ppo_loss_bwd")
    return g


ppo_loss_bwd_op = torch.library.custom_op("b200rl::ppo_loss_bwd", _raw_ppo_loss_bwd, mutates_args=())


@ppo_loss_bwd_op.register_fake
def _(dlogp, grad_loss, denom):
    return torch.empty_like(dlogp)


def _loss_setup_context(ctx, inputs, output):
    ctx.save_for_backward(output[2], output[1])


def _loss_backward(ctx, grads):
    dlogp, denom = ctx.saved_tensors
    g_loss = grads[0]
    g_lp = None
    if g_loss is not None:
        g_lp = ppo_loss_bwd_op(dlogp, g_loss, denom)
    # like the reference in practice: only `logprobs` carries a gradient
    return (g_lp,) + (None,) * 11


ppo_loss_fwd_op.register_autograd(_loss_backward, setup_context=_loss_setup_context)


def _ppo_loss_slabs(logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap):
    """Eager-mode launcher of b200rl_ppo_loss_fwd with every output in two slabs (one fp32
    [6, Tp], one byte [3, Tp]) and one [2] scalar pair: four allocations instead of fourteen, which
    is most of what a call costs on the host (the kernel itself runs 3-8 us).
    -> (loss, N, fs rows = dlogp, s_loss, s_iw, s_kl, s_bw, s_bkl; bs rows = clip, dual, bm)."""
    shape = logp.shape
    dev = logp.device
    T = logp.numel()

    def flat(t: Tensor, name: str) -> Tensor:
        if t.shape != shape:
            raise RuntimeError(f"{name} has shape {tuple(t.shape)}, expected {tuple(shape)}")
        if t.dtype != torch.float32:
            raise RuntimeError(f"{name} must be float32, got {t.dtype}")
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be on CUDA (b200rl has no CPU path)")
        return t if t.is_contiguous() else t.contiguous()

    lp, px, ol, ad = flat(logp.detach(), "logprobs"), flat(prox_logp, "proximal_logprobs"), \
        flat(old_logp, "old_logprobs"), flat(adv, "advantages")
    if mask.shape != shape:
        raise RuntimeError(f"loss_mask has shape {tuple(mask.shape)}, expected {tuple(shape)}")
    if not mask.is_cuda:
        raise RuntimeError("loss_mask must be on CUDA (b200rl has no CPU path)")
    mk = mask if mask.dtype == torch.bool else mask != 0
    if not mk.is_contiguous():
        mk = mk.contiguous()
    Tp = (T + 63) // 64 * 64
    fs = torch.empty((6, Tp), dtype=torch.float32, device=dev)
    bs = torch.empty((3, Tp), dtype=torch.bool, device=dev)  # the kernel writes 0/1 bytes
    loss = torch.empty((), dtype=torch.float32, device=dev)
    denom = torch.empty(1, dtype=torch.float32, device=dev)
    ws = torch.empty(max(lib.b200rl_ppo_loss_workspace_bytes(T), 256), dtype=torch.uint8, device=dev)
    fb, st4, bb = fs.data_ptr(), Tp * 4, bs.data_ptr()
    f = lambda i: c_void_p(fb + i * st4)
    b = lambda i: c_void_p(bb + i * Tp)
    with _on(dev):
        rc = lib.b200rl_ppo_loss_fwd(
            ptr(lp), ptr(px), ptr(ol), ptr(ad), ptr(mk), T, float(clip_lo), float(clip_hi), float(c_clip),
            float(cap), None, None, None, None, f(1), f(2), f(3), b(0), b(1), f(4), f(5), b(2), f(0),
            ptr(loss), ptr(denom), ptr(ws), ws.numel(), stream_handle(dev))
    check(rc, "b200rl_ppo_loss_fwd")
    return loss, denom, fs, bs


class _PpoLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap):
        loss, denom, fs, bs = _ppo_loss_slabs(logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap)
        ctx.save_for_backward(fs, denom)
        ctx.shape = logp.shape
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(fs, bs, denom)
        return loss, fs, bs, denom

    @staticmethod
    def backward(ctx, g_loss, *unused):
        if g_loss is None:
            return (None,) * 9
        fs, denom = ctx.saved_tensors
        T = 1
        for d in ctx.shape:
            T *= d
        dev = fs.device
        g = torch.empty(ctx.shape, dtype=torch.float32, device=dev)
        gl = g_loss if g_loss.dtype == torch.float32 else g_loss.to(torch.float32)
        with _on(dev):
            rc = lib.b200rl_ppo_loss_bwd(c_void_p(fs.data_ptr()), ptr(gl), ptr(denom), ptr(g), T,
                                         stream_handle(dev))
        check(rc, "b200rl_ppo_loss_bwd")
        return (g,) + (None,) * 8


def ppo_loss_fwd(logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap,
                 adv_mean, adv_denom, adv_group):
    args = (logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap, adv_mean, adv_denom,
            adv_group)
    if not _eager():
        return ppo_loss_fwd_op(*args)
    if adv_mean is not None or adv_denom is not None:  # fused advantage normalisation: the general launcher
        if logp.requires_grad and torch.is_grad_enabled():
            return list(_PpoLossFusedNormFn.apply(*args))
        return _raw_ppo_loss_fwd(*args)
    if logp.requires_grad and torch.is_grad_enabled():
        loss, fs, bs, denom = _PpoLossFn.apply(*args[:9])
    else:
        loss, denom, fs, bs = _ppo_loss_slabs(*args[:9])
    shape = logp.shape
    T = logp.numel()
    if fs.shape[1] != T:
        fs, bs = fs[:, :T], bs[:, :T]
    f = fs.unbind(0)
    b = bs.unbind(0)
    if len(shape) != 1:
        f = [t.reshape(shape) for t in f]
        b = [t.reshape(shape) for t in b]
    return [loss, denom, f[0], fs.new_empty(0), f[1], f[2], f[3], b[0], b[1], f[4], f[5], b[2]]


class _PpoLossFusedNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap,
                adv_mean, adv_denom, adv_group):
        out = _raw_ppo_loss_fwd(logp, prox_logp, old_logp, adv, mask, clip_lo, clip_hi, c_clip, cap,
                                adv_mean, adv_denom, adv_group)
        ctx.save_for_backward(out[2], out[1])
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(*out[1:])
        return tuple(out)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        dlogp, denom = ctx.saved_tensors
        g_lp = None if g_loss is None else _raw_ppo_loss_bwd(dlogp, g_loss, denom)
        return (g_lp,) + (None,) * 11


def ppo_loss_bwd(dlogp, grad_loss, denom):
    return _raw_ppo_loss_bwd(dlogp, grad_loss, denom) if _eager() else ppo_loss_bwd_op(dlogp, grad_loss, denom)


# --------------------------------------------------------------------------------------------
# critic (value) loss and packed rewards: companions of the GAE path (legacy PPO interface)
# --------------------------------------------------------------------------------------------
def _raw_critic_loss_fwd(value: Tensor, old_value: Tensor, target_value: Tensor,
                         mask: Optional[Tensor], eps: float, loss_type: int, delta: float) -> List[Tensor]:
    """-> [loss(0-d), denom(1), dvalue, stat_loss, clip_mask(u8)]"""
    _lib.require_cuda(value=value, old_value=old_value, target_value=target_value)
    shape, dev = value.shape, value.device

    def flat(t: Tensor, name: str) -> Tensor:
        if t.shape != shape:
            raise RuntimeError(f"{name} has shape {tuple(t.shape)}, expected {tuple(shape)}")
        if t.dtype != torch.float32:
            raise RuntimeError(f"{name} must be float32, got {t.dtype}")
        return t.detach().contiguous().view(-1)

    v, ov, tv = flat(value, "value"), flat(old_value, "old_value"), flat(target_value, "target_value")
    mk = None
    if mask is not None:
        if mask.shape != shape:
            raise RuntimeError(f"loss_mask has shape {tuple(mask.shape)}, expected {tuple(shape)}")
        _lib.require_cuda(loss_mask=mask)
        mk = (mask if mask.dtype == torch.bool else mask != 0).contiguous().view(-1).view(torch.uint8)
    T = v.numel()
    loss = torch.empty((), dtype=torch.float32, device=dev)
    denom = torch.empty(1, dtype=torch.float32, device=dev)
    dvalue, s_loss = torch.empty_like(v), torch.empty_like(v)
    s_clip = torch.empty(T, dtype=torch.uint8, device=dev)
    ws = _lib.workspace(lib.b200rl_critic_loss_workspace_bytes(T), dev)
    with _on(dev):
        rc = lib.b200rl_critic_loss_fwd(ptr(v), ptr(ov), ptr(tv), ptr(mk), T, float(eps), int(loss_type),
                                        float(delta), ptr(s_loss), ptr(s_clip), ptr(dvalue), ptr(loss),
                                        ptr(denom), ptr(ws), ws.numel(), stream_handle(dev))
    check(rc, "b200rl_critic_loss_fwd")
    return [loss, denom, dvalue.view(shape), s_loss.view(shape), s_clip.view(shape)]


critic_loss_fwd_op = torch.library.custom_op("b200rl::critic_loss_fwd", _raw_critic_loss_fwd, mutates_args=())


@critic_loss_fwd_op.register_fake
def _(value, old_value, target_value, mask, eps, loss_type, delta):
    return [value.new_empty((), dtype=torch.float32), value.new_empty((1,), dtype=torch.float32),
            torch.empty_like(value), torch.empty_like(value), torch.empty_like(value, dtype=torch.uint8)]


def _critic_backward(ctx, grads):
    dvalue, denom = ctx.saved_tensors
    g = None if grads[0] is None else ppo_loss_bwd_op(dvalue, grads[0], denom)
    return (g,) + (None,) * 6


critic_loss_fwd_op.register_autograd(
    _critic_backward, setup_context=lambda ctx, inputs, output: ctx.save_for_backward(output[2], output[1]))


class _CriticLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, value, old_value, target_value, mask, eps, loss_type, delta):
        out = _raw_critic_loss_fwd(value, old_value, target_value, mask, eps, loss_type, delta)
        ctx.save_for_backward(out[2], out[1])
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(*out[1:])
        return tuple(out)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        dvalue, denom = ctx.saved_tensors
        g = None if g_loss is None else _raw_ppo_loss_bwd(dvalue, g_loss, denom)
        return (g,) + (None,) * 6


def critic_loss_fwd(value, old_value, target_value, mask, eps, loss_type, delta):
    args = (value, old_value, target_value, mask, eps, loss_type, delta)
    if not _eager():
        return critic_loss_fwd_op(*args)
    if value.requires_grad and torch.is_grad_enabled():
        return list(_CriticLossFn.apply(*args))
    return _raw_critic_loss_fwd(*args)


def _raw_packed_rewards(log_probs: Tensor, ref_log_probs: Tensor, reward_score: Tensor,
                        short1cu_seqlens: Tensor, seq_no_eos_mask: Tensor, kl_ctl: float,
                        clip_reward_value: float, mask_no_eos_with_zero: bool) -> List[Tensor]:
    """-> [kl_rewards, tot_rewards]"""
    _lib.require_cuda(log_probs=log_probs, ref_log_probs=ref_log_probs, reward_score=reward_score,
                      short1cu_seqlens=short1cu_seqlens, seq_no_eos_mask=seq_no_eos_mask)
    dev = log_probs.device
    if log_probs.dim() != 1 or ref_log_probs.shape != log_probs.shape:
        raise RuntimeError("log_probs and ref_log_probs must be 1-D and of the same length")
    lp = log_probs.detach().to(torch.float32).contiguous()
    rf = ref_log_probs.detach().to(torch.float32).contiguous()
    sc = reward_score.detach().to(torch.float32).contiguous().view(-1)
    cu = short1cu_seqlens.to(torch.int32).contiguous().view(-1)
    ne = (seq_no_eos_mask if seq_no_eos_mask.dtype == torch.bool else seq_no_eos_mask != 0)
    ne = ne.contiguous().view(-1).view(torch.uint8)
    B = cu.numel() - 1
    if sc.numel() != B or ne.numel() != B:
        raise RuntimeError(f"reward_score/seq_no_eos_mask need {B} entries (one per sequence)")
    kl, tot = torch.empty_like(lp), torch.empty_like(lp)
    with _on(dev):
        rc = lib.b200rl_packed_rewards(ptr(lp), ptr(rf), ptr(sc), ptr(cu), ptr(ne), B, lp.numel(),
                                       float(kl_ctl), float(clip_reward_value),
                                       int(bool(mask_no_eos_with_zero)), ptr(kl), ptr(tot),
                                       stream_handle(dev))
    check(rc, "b200rl_packed_rewards")
    return [kl, tot]


packed_rewards_op = torch.library.custom_op("b200rl::packed_rewards", _raw_packed_rewards, mutates_args=())


@packed_rewards_op.register_fake
def _(log_probs, ref_log_probs, reward_score, short1cu_seqlens, seq_no_eos_mask, kl_ctl,
      clip_reward_value, mask_no_eos_with_zero):
    return [torch.empty_like(log_probs, dtype=torch.float32), torch.empty_like(log_probs, dtype=torch.float32)]


def packed_rewards(*args):
    return _raw_packed_rewards(*args) if _eager() else packed_rewards_op(*args)


# --------------------------------------------------------------------------------------------
# packed SFT loss tail
# --------------------------------------------------------------------------------------------
def _raw_sft_loss_fwd(logp: Tensor, mask: Tensor, cu_seqlens: Tensor) -> List[Tensor]:
    """-> [loss(0-d), denom(1), dlogp, masked_logp, seqlogp(f64 [B]), ppl(f32 [B])]"""
    _lib.require_cuda(logprobs=logp, loss_mask=mask, cu_seqlens=cu_seqlens)
    if logp.dim() != 1 or mask.shape != logp.shape:
        raise RuntimeError("logprobs and loss_mask must be 1-D and of the same length")
    if logp.dtype != torch.float32:
        raise RuntimeError(f"logprobs must be float32, got {logp.dtype}")
    if cu_seqlens.dtype not in (torch.int32, torch.int64) or cu_seqlens.dim() != 1:
        raise RuntimeError("cu_seqlens must be a 1-D int32 or int64 tensor")
    dev = logp.device
    lp = logp.detach().contiguous()
    mk = (mask if mask.dtype == torch.bool else mask != 0).contiguous().view(torch.uint8)
    cu = cu_seqlens.contiguous()
    T, B = lp.numel(), max(cu.numel() - 1, 0)
    loss = torch.empty((), dtype=torch.float32, device=dev)
    denom = torch.empty(1, dtype=torch.float32, device=dev)
    dlogp, masked = torch.empty_like(lp), torch.empty_like(lp)
    seqlogp = torch.empty(B, dtype=torch.float64, device=dev)
    ppl = torch.empty(B, dtype=torch.float32, device=dev)
    ws = _lib.workspace(lib.b200rl_sft_loss_workspace_bytes(T, B), dev)
    with _on(dev):
        rc = lib.b200rl_sft_loss_fwd(ptr(lp), ptr(mk), ptr(cu), int(cu.dtype == torch.int64), T, B,
                                     ptr(masked), ptr(dlogp), ptr(seqlogp), ptr(ppl), ptr(loss),
                                     ptr(denom), ptr(ws), ws.numel(), stream_handle(dev))
    check(rc, "b200rl_sft_loss_fwd")
    return [loss, denom, dlogp, masked, seqlogp, ppl]


sft_loss_fwd_op = torch.library.custom_op("b200rl::sft_loss_fwd", _raw_sft_loss_fwd, mutates_args=())


@sft_loss_fwd_op.register_fake
def _(logp, mask, cu_seqlens):
    B = max(cu_seqlens.shape[0] - 1, 0)
    return [logp.new_empty((), dtype=torch.float32), logp.new_empty((1,), dtype=torch.float32),
            torch.empty_like(logp), torch.empty_like(logp), logp.new_empty((B,), dtype=torch.float64),
            logp.new_empty((B,), dtype=torch.float32)]


def _sft_backward(ctx, grads):
    dlogp, denom = ctx.saved_tensors
    g = None if grads[0] is None else ppo_loss_bwd_op(dlogp, grads[0], denom)
    return g, None, None


sft_loss_fwd_op.register_autograd(
    _sft_backward, setup_context=lambda ctx, inputs, output: ctx.save_for_backward(output[2], output[1]))


class _SftLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logp, mask, cu_seqlens):
        out = _raw_sft_loss_fwd(logp, mask, cu_seqlens)
        ctx.save_for_backward(out[2], out[1])
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(*out[1:])
        return tuple(out)

    @staticmethod
    def backward(ctx, g_loss, *unused):
        dlogp, denom = ctx.saved_tensors
        g = None if g_loss is None else _raw_ppo_loss_bwd(dlogp, g_loss, denom)
        return g, None, None


def sft_loss_fwd(logp, mask, cu_seqlens):
    if not _eager():
        return sft_loss_fwd_op(logp, mask, cu_seqlens)
    if logp.requires_grad and torch.is_grad_enabled():
        return list(_SftLossFn.apply(logp, mask, cu_seqlens))
    return _raw_sft_loss_fwd(logp, mask, cu_seqlens)


# --------------------------------------------------------------------------------------------
# H2 masked moments / normalisation
# --------------------------------------------------------------------------------------------
def _raw_masked_moments(
    x: Tensor, mask: Optional[Tensor], group_elems: int, center: Optional[Tensor],
    center_group_elems: int,
) -> Tensor:
    """x, mask: flat fp32 -> moments [G, 3] fp64 = (sum mask, sum x*mask, sum ((x-center)*mask)^2)."""
    _lib.require_cuda(x=x, mask=mask, center=center)
    _lib.require_contiguous(x=x, mask=mask, center=center)
    n = x.numel()
    dev = x.device
    G = n // group_elems if group_elems > 0 else 0
    mom = torch.empty((G, 3), dtype=torch.float64, device=dev)
    if n == 0:
        return mom
    ws = _lib.workspace(lib.b200rl_moments_workspace_bytes(n, group_elems), dev)
    with _on(dev):
        rc = lib.b200rl_masked_moments(
            ptr(x), ptr(mask), n, group_elems, ptr(center), center_group_elems, ptr(mom),
            ptr(ws), ws.numel(), stream_handle(dev),
        )
    check(rc, "b200rl_masked_moments")
    return mom


masked_moments_op = torch.library.custom_op("b200rl::masked_moments", _raw_masked_moments, mutates_args=())


@masked_moments_op.register_fake
def _(x, mask, group_elems, center, center_group_elems):
    return x.new_empty((x.numel() // group_elems, 3), dtype=torch.float64)


def masked_moments(x, mask, group_elems, center, center_group_elems):
    if _eager():
        return _raw_masked_moments(x, mask, group_elems, center, center_group_elems)
    return masked_moments_op(x, mask, group_elems, center, center_group_elems)


def _raw_moments_finalize(moments: Tensor, centered: bool, unbiased: bool, eps: float) -> Tuple[Tensor, Tensor]:
    """-> (mean[G], denom[G] = sqrt(var) + eps), fp64."""
    _lib.require_cuda(moments=moments)
    _lib.require_contiguous(moments=moments)
    G = moments.shape[0]
    dev = moments.device
    mean = torch.empty(G, dtype=torch.float64, device=dev)
    denom = torch.empty(G, dtype=torch.float64, device=dev)
    with _on(dev):
        rc = lib.b200rl_moments_finalize(ptr(moments), G, int(centered), int(unbiased), float(eps),
                                         ptr(mean), ptr(denom), stream_handle(dev))
    check(rc, "b200rl_moments_finalize")
    return mean, denom


moments_finalize_op = torch.library.custom_op("b200rl::moments_finalize", _raw_moments_finalize, mutates_args=())


@moments_finalize_op.register_fake
def _(moments, centered, unbiased, eps):
    G = moments.shape[0]
    return moments.new_empty((G,)), moments.new_empty((G,))


def moments_finalize(moments, centered, unbiased, eps):
    if _eager():
        return _raw_moments_finalize(moments, centered, unbiased, eps)
    return moments_finalize_op(moments, centered, unbiased, eps)


def _raw_masked_normalize(
    x: Tensor, mask: Optional[Tensor], mean: Optional[Tensor], mean_group_elems: int,
    denom: Optional[Tensor], denom_group_elems: int, mask_x: bool,
) -> Tensor:
    _lib.require_cuda(x=x, mask=mask, mean=mean, denom=denom)
    _lib.require_contiguous(x=x, mask=mask, mean=mean, denom=denom)
    out = torch.empty_like(x, dtype=torch.float32)
    dev = x.device
    with _on(dev):
        rc = lib.b200rl_masked_normalize(
            ptr(x), ptr(mask), x.numel(), ptr(mean), mean_group_elems, ptr(denom),
            denom_group_elems, int(mask_x), ptr(out), stream_handle(dev),
        )
    check(rc, "b200rl_masked_normalize")
    return out


masked_normalize_op = torch.library.custom_op("b200rl::masked_normalize", _raw_masked_normalize, mutates_args=())


@masked_normalize_op.register_fake
def _(x, mask, mean, mean_group_elems, denom, denom_group_elems, mask_x):
    return torch.empty_like(x, dtype=torch.float32)


def masked_normalize(x, mask, mean, mean_group_elems, denom, denom_group_elems, mask_x):
    if _eager():
        return _raw_masked_normalize(x, mask, mean, mean_group_elems, denom, denom_group_elems, mask_x)
    return masked_normalize_op(x, mask, mean, mean_group_elems, denom, denom_group_elems, mask_x)


def _raw_masked_normalize_groups(x: Tensor, mask: Optional[Tensor], group_elems: int, unbiased: bool,
                                 eps: float) -> Tensor:
    """(x*mask - mean[g]) / (std[g] + eps) per group of `group_elems` consecutive elements, same
    level for mean and std, no collective: one host call (three launches). x, mask flat fp32."""
    _lib.require_cuda(x=x, mask=mask)
    _lib.require_contiguous(x=x, mask=mask)
    n, dev = x.numel(), x.device
    out = torch.empty_like(x, dtype=torch.float32)
    if n == 0:
        return out
    G = n // group_elems
    st = torch.empty((G, 5), dtype=torch.float64, device=dev)  # moments [G,3] | mean [G] | denom [G]
    ws = torch.empty(max(lib.b200rl_moments_workspace_bytes(n, group_elems), 256), dtype=torch.uint8, device=dev)
    sb = st.data_ptr()
    with _on(dev):
        rc = lib.b200rl_masked_normalize_groups(
            ptr(x), ptr(mask), n, group_elems, int(unbiased), float(eps), ptr(out), c_void_p(sb),
            c_void_p(sb + G * 24), c_void_p(sb + G * 32), ptr(ws), ws.numel(), stream_handle(dev))
    check(rc, "b200rl_masked_normalize_groups")
    return out


# --------------------------------------------------------------------------------------------
# H4 interval gather / scatter
# --------------------------------------------------------------------------------------------
def _check_intervals(intervals: Tensor) -> int:
    if not intervals.is_cuda:
        raise RuntimeError("intervals must be on CUDA")
    if intervals.dtype != torch.int64:
        raise RuntimeError("intervals must be of type long")
    if intervals.dim() != 2 or intervals.shape[1] != 2:
        raise RuntimeError("intervals must have shape (N, 2)")
    if not intervals.is_contiguous():
        raise RuntimeError("intervals must be contiguous")
    return intervals.shape[0]


@torch.library.custom_op("b200rl::slice_intervals", mutates_args=("workspace",))
def slice_intervals_op(
    src: Tensor, intervals: Tensor, total_size: int, workspace: Tensor, reuse_plan: bool
) -> Tensor:
    _lib.require_cuda(src=src, workspace=workspace)
    _lib.require_contiguous(src=src)
    n = _check_intervals(intervals)
    dev = src.device
    dst = torch.empty((total_size,), dtype=src.dtype, device=dev)  # interval_op.cu:82
    with _on(dev):
        rc = lib.b200rl_slice_intervals(
            ptr(src), src.numel(), ptr(dst), total_size, ptr(intervals), n, src.element_size(),
            ptr(workspace), workspace.numel(), int(reuse_plan), stream_handle(dev),
        )
    check(rc, "b200rl_slice_intervals")
    return dst


@slice_intervals_op.register_fake
def _(src, intervals, total_size, workspace, reuse_plan):
    return src.new_empty((total_size,))


@torch.library.custom_op("b200rl::set_intervals", mutates_args=("dst", "workspace"))
def set_intervals_op(
    src: Tensor, dst: Tensor, intervals: Tensor, workspace: Tensor, reuse_plan: bool
) -> None:
    _lib.require_cuda(src=src, dst=dst, workspace=workspace)
    _lib.require_contiguous(src=src, dst=dst)
    if src.dtype != dst.dtype:
        raise RuntimeError("Source and destination tensors must have the same dtype")
    n = _check_intervals(intervals)
    dev = src.device
    with _on(dev):
        rc = lib.b200rl_set_intervals(
            ptr(src), src.numel(), ptr(dst), dst.numel(), ptr(intervals), n, src.element_size(),
            ptr(workspace), workspace.numel(), int(reuse_plan), stream_handle(dev),
        )
    check(rc, "b200rl_set_intervals")


def move_intervals_raw(src, dst: Tensor, src_intervals: Tensor, dst_intervals: Tensor,
                       total_size: int, workspace: Tensor, reuse_plan: bool) -> None:
    """dst[dst_intervals] <- src[src_intervals] with no packed staging buffer. ``src`` is a tensor
    on ``dst``'s device or a ``dist.PeerBuffer`` (another rank's memory, read over NVLink);
    everything else is on ``dst``'s device, where the kernel runs. Eager only (in-place on dst)."""
    peer = bool(getattr(src, "is_peer", False))
    _lib.require_cuda(dst=dst, workspace=workspace)
    if not peer:
        _lib.require_cuda(src=src)
    _lib.require_contiguous(src=src, dst=dst)
    if src.dtype != dst.dtype:
        raise RuntimeError("Source and destination tensors must have the same dtype")
    ns, nd = _check_intervals(src_intervals), _check_intervals(dst_intervals)
    dev = dst.device
    if src_intervals.device != dev or dst_intervals.device != dev or workspace.device != dev:
        raise RuntimeError("intervals and workspace must be on the destination's device")
    if src.device != dev:
        if peer:
            raise RuntimeError(f"the peer buffer was opened for {src.device}, the destination is on {dev}")
        # a tensor of this process on another of its GPUs
        check(lib.b200rl_enable_peer_access(dev.index, src.device.index), "b200rl_enable_peer_access")
    with _on(dev):
        rc = lib.b200rl_move_intervals(
            ctypes.c_void_p(src.data_ptr()), src.numel(), ptr(dst), dst.numel(), ptr(src_intervals), ns,
            ptr(dst_intervals), nd, int(total_size), src.element_size(), ptr(workspace),
            workspace.numel(), int(reuse_plan), stream_handle(dev))
    check(rc, "b200rl_move_intervals")
