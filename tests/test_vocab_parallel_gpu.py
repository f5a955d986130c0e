"""Vocabulary-parallel log-prob / entropy (ref megatron_actor.py:60-66,323-331,
mcore/functional.py:8-52). megatron.core is not under the reference tree; what it computes is
log_softmax(cat(shards)/T) at the label and its entropy, so the checks are against the
whole-vocabulary oracle (pinned by the h1_* fixtures) and against the whole-vocabulary kernel.

One GPU plays every tensor-parallel rank in turn (the all-gather becomes a torch.stack); the NCCL
path itself is exercised by tests/test_vocab_parallel_dist.py and scripts/tp_check.py.
Tolerances: logp / entropy 1e-3 relative (north_star; observed ~1e-6), gradients one ulp of the
logits dtype, row min/max exact."""

import numpy as np
import pytest
import torch

from conftest import golden_logits, load_golden
from oracle import torch_ref

pytestmark = pytest.mark.gpu


def _shards(logits, tp):
    Vs = logits.shape[1] // tp
    return [logits[:, k * Vs:(k + 1) * Vs].contiguous() for k in range(tp)], Vs


def _run_sharded(logits, labels, temp, tp, need_ent=True, need_mm=True):
    from areal_vllm_b200 import ops

    shards, Vs = _shards(logits, tp)
    parts = [ops._raw_logprob_shard_fwd(sh, (labels - k * Vs).contiguous(), temp, need_ent, need_mm)
             for k, sh in enumerate(shards)]
    out = ops._raw_logprob_shard_combine(torch.stack(parts), temp, need_ent, need_mm)
    return shards, Vs, parts, out


@pytest.mark.parametrize("name,tp", [("h1_bf16_v1024_t1", 2), ("h1_bf16_v1024_t07", 4),
                                     ("h1_fp32_v777", 7), ("h1_fp16_v2048", 8),
                                     ("h1_bf16_v151936", 2), ("h1_bf16_v151936", 8)])
def test_golden_sharded(name, tp, cuda):
    from areal_vllm_b200 import ops

    g = load_golden(name)
    logits, dref = golden_logits(g)
    logits = logits.to(cuda)
    labels = torch.from_numpy(g["labels"]).to(cuda)
    temp = float(g["temperature"])
    shards, Vs, _, (logp, ent, lse, rmax, rmin) = _run_sharded(logits, labels, temp, tp)
    np.testing.assert_allclose(logp.cpu().numpy(), g["logp"], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(ent.cpu().numpy(), g["entropy"], rtol=1e-4, atol=2e-5)
    assert np.array_equal(rmax.cpu().numpy(), g["vmax"])
    assert np.array_equal(rmin.cpu().numpy(), g["vmin"])
    g_lp = torch.from_numpy(g["g_logp"]).to(cuda)
    g_ent = torch.from_numpy(g["g_ent"]).to(cuda) if bool(g["has_ent_grad"]) else None
    d = torch.cat([ops._raw_logprob_entropy_bwd(sh, (labels - k * Vs).contiguous(), lse, ent, g_lp, g_ent, temp)
                   for k, sh in enumerate(shards)], dim=1)
    tol = 2e-5 if logits.dtype == torch.float32 else 8e-3
    np.testing.assert_allclose(d.float().cpu().numpy(), dref.float().numpy(), rtol=tol, atol=2e-7)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("tp,V", [(1, 4096), (2, 151936), (4, 151936), (8, 151936), (3, 3000), (2, 1002)])
def test_sharded_matches_whole(dtype, tp, V, cuda):
    """shards of widths that hit the TMA kernel (V/tp a multiple of the 16-byte vector) and the
    plain kernel (3000/3 = 1000 bf16 is, 1002/2 = 501 is not)."""
    from areal_vllm_b200 import ops

    T = 67
    gen = torch.Generator(device="cpu").manual_seed(V + tp)
    logits = (torch.randn(T, V, generator=gen) * 3).to(dtype)
    labels = torch.randint(0, V, (T,), generator=gen)
    labels[0], labels[1] = 0, V - 1
    logits[2, labels[2]] = 40.0   # label is the (large) row maximum
    logits[3] -= 300.0            # far-negative row
    temp = 0.8
    x = logits.to(cuda)
    y = labels.to(cuda)
    whole = ops._raw_logprob_entropy_fwd(x, y, temp, True, True)
    shards, Vs, parts, out = _run_sharded(x, y, temp, tp)
    for a, b in zip(out, whole):
        np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=2e-6, atol=2e-6)
    assert torch.equal(out[3], whole[3]) and torch.equal(out[4], whole[4])
    lp64, ent64 = torch_ref.logprobs_entropy_f64(logits, labels, temp)
    np.testing.assert_allclose(out[0].cpu().numpy(), lp64.numpy(), rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(out[1].cpu().numpy(), ent64.numpy(), rtol=1e-4, atol=2e-5)
    # exactly one shard owns each label
    owner = torch.stack([~torch.isnan(p[:, 3]) for p in parts]).sum(0)
    assert torch.all(owner == 1)
    # without entropy / minmax
    lp_only = _run_sharded(x, y, temp, tp, need_ent=False, need_mm=False)[3]
    assert lp_only[1].numel() == 0 and lp_only[3].numel() == 0
    np.testing.assert_allclose(lp_only[0].cpu().numpy(), out[0].cpu().numpy(), rtol=1e-6, atol=1e-6)


def test_autograd_single_rank_and_explicit_start(cuda):
    """group=None: one shard, behaves like gather_logprobs_entropy; explicit vocab_start lets one
    process stand in for a rank (used by engines that pass their own offset)."""
    from areal_vllm_b200.functional import gather_logprobs_entropy
    from areal_vllm_b200.mcore_functional import (vocab_parallel_entropy, vocab_parallel_logprobs,
                                                  vocab_parallel_logprobs_entropy)

    gen = torch.Generator().manual_seed(3)
    T, V = 50, 2048
    logits = (torch.randn(T, V, generator=gen) * 2).bfloat16().to(cuda)
    labels = torch.randint(0, V, (T,), generator=gen).to(cuda)
    g1 = torch.randn(T, generator=gen).to(cuda)
    g2 = torch.randn(T, generator=gen).to(cuda) * 0.1
    a = logits.clone().requires_grad_(True)
    lp_a, ent_a = gather_logprobs_entropy(a, labels, 0.9)
    ((lp_a * g1).sum() + (ent_a * g2).sum()).backward()
    b = logits.clone().requires_grad_(True)
    lp_b, ent_b = vocab_parallel_logprobs_entropy(b, labels, 0.9)
    ((lp_b * g1).sum() + (ent_b * g2).sum()).backward()
    np.testing.assert_allclose(lp_b.detach().cpu().numpy(), lp_a.detach().cpu().numpy(), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(ent_b.detach().cpu().numpy(), ent_a.detach().cpu().numpy(), rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(b.grad.float().cpu().numpy(), a.grad.float().cpu().numpy(), rtol=8e-3, atol=1e-7)
    c = logits.clone().requires_grad_(True)
    lp_c = vocab_parallel_logprobs(c, labels, 0.9)
    (lp_c * g1).sum().backward()
    np.testing.assert_allclose(lp_c.detach().cpu().numpy(), lp_a.detach().cpu().numpy(), rtol=1e-6, atol=1e-6)
    # entropy-only twin of _VocabParallelEntropy.apply (no temperature argument)
    e = vocab_parallel_entropy(logits)
    ref = torch_ref.logprobs_entropy_f64(logits.cpu(), labels.cpu(), 1.0)[1]
    np.testing.assert_allclose(e.cpu().numpy(), ref.numpy(), rtol=1e-4, atol=2e-5)
    # 2-D leading shape
    lp2, ent2 = vocab_parallel_logprobs_entropy(logits.view(5, 10, V), labels.view(5, 10), 0.9)
    assert lp2.shape == (5, 10) and torch.equal(lp2.reshape(-1), lp_b.detach())


def test_megatron_grpo_loss_single_rank(cuda):
    """the Megatron actor's loss fn (no eps_clip_higher) equals the FSDP actor's on one rank."""
    from areal_vllm_b200 import actor, mcore_functional

    gen = torch.Generator().manual_seed(11)
    T, V = 300, 4096
    logits = (torch.randn(T, V, generator=gen) * 2).bfloat16().to(cuda)
    data = dict(input_ids=torch.randint(0, V, (T,), generator=gen).to(cuda),
                logprobs=(-torch.rand(T, generator=gen) * 8).to(cuda),
                advantages=torch.randn(T, generator=gen).to(cuda),
                loss_mask=(torch.rand(T, generator=gen) < 0.7).to(cuda),
                prox_logp=(-torch.rand(T, generator=gen) * 8).to(cuda))
    a = logits.clone().requires_grad_(True)
    sa = {}
    la = actor.grpo_loss_fn(a, data, 1.0, 0.2, None, 3.0, 5.0, stats=sa)
    la.backward()
    b = logits.clone().requires_grad_(True)
    sb = {}
    lb = mcore_functional.grpo_loss_fn(b, data, 1.0, 0.2, 3.0, 5.0, stats=sb)
    lb.backward()
    np.testing.assert_allclose(lb.item(), la.item(), rtol=1e-6)
    np.testing.assert_allclose(b.grad.float().cpu().numpy(), a.grad.float().cpu().numpy(), rtol=8e-3, atol=1e-9)
    assert set(sb) == set(sa)
    for k in sa:
        va, vb = sa[k], sb[k]
        if va.dtype == torch.bool:
            assert torch.equal(va, vb), k
        else:
            np.testing.assert_allclose(vb.float().cpu().numpy(), va.float().cpu().numpy(), rtol=1e-5, atol=1e-6, err_msg=k)


def test_shard_errors(cuda):
    from areal_vllm_b200 import ops

    x = torch.zeros(4, 64, device=cuda, dtype=torch.bfloat16)
    y = torch.zeros(4, dtype=torch.int64, device=cuda)
    with pytest.raises(RuntimeError):
        ops._raw_logprob_shard_fwd(x, y.int(), 1.0, True, True)
    with pytest.raises(RuntimeError):
        ops._raw_logprob_shard_fwd(x, y, 0.0, True, True)
    with pytest.raises(RuntimeError):
        ops._raw_logprob_shard_combine(torch.zeros(2, 4, 7, device=cuda), 1.0, True, True)
    # a label owned by no shard gives nan, like an out-of-range label in the whole-vocabulary op
    p = ops._raw_logprob_shard_fwd(x, y - 5, 1.0, True, True)
    out = ops._raw_logprob_shard_combine(p.unsqueeze(0), 1.0, True, True)
    assert torch.isnan(out[0]).all() and torch.isfinite(out[1]).all()


@pytest.mark.parametrize("tp,T", [(2, 300), (8, 1000), (3, 1)])
def test_peer_store_exchange_single_process(tp, T, cuda):
    """b200rl_tp_push / b200rl_tp_combine with one GPU playing every rank: each "rank" owns an
    exchange buffer, pushes its records into all of them, and every buffer's combine must give the
    same bits as b200rl_logprob_shard_combine on the stacked records - on two calls in a row (the two
    slots alternate), with the second call shorter than the first (stale rows stay behind)."""
    import ctypes

    from areal_vllm_b200 import _lib, ops

    lib = _lib.lib
    V, max_tokens = 4096, 1024
    Vs = (V // tp) // 8 * 8
    nbytes = lib.b200rl_tp_exchange_bytes(tp, max_tokens)
    assert nbytes > 0 and lib.b200rl_tp_exchange_bytes(17, 10) == 0
    bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=cuda) for _ in range(tp)]
    ptrs = (ctypes.c_void_p * tp)(*[b.data_ptr() for b in bufs])
    status = torch.zeros(1, dtype=torch.int32, device=cuda)
    st = _lib.stream_handle(cuda)
    gen = torch.Generator().manual_seed(tp * 1000 + T)
    for epoch, Tc in ((1, T), (2, max(T // 2, 1)), (3, T)):
        x = (torch.randn(Tc, Vs * tp, generator=gen) * 2).bfloat16().to(cuda)
        y = torch.randint(0, Vs * tp, (Tc,), generator=gen).to(cuda)
        parts = [ops._raw_logprob_shard_fwd(x[:, r * Vs:(r + 1) * Vs].contiguous(), (y - r * Vs).contiguous(), 0.8, True, True)
                 for r in range(tp)]
        want = ops._raw_logprob_shard_combine(torch.stack(parts), 0.8, True, True)
        for r in range(tp):
            rc = lib.b200rl_tp_push(_lib.ptr(parts[r]), Tc, ptrs, tp, r, max_tokens, epoch, st)
            assert rc == 0, _lib.last_error()
        for r in range(tp):
            outs = [torch.empty(Tc, device=cuda) for _ in range(5)]
            rc = lib.b200rl_tp_combine(_lib.ptr(bufs[r]), tp, Tc, max_tokens, epoch, 0.8, *[_lib.ptr(o) for o in outs],
                                       _lib.ptr(status), st)
            assert rc == 0, _lib.last_error()
            for got, ref in zip(outs, want):
                assert torch.equal(got, ref)
    assert int(status.item()) == 0
    # argument errors: too many rows for the buffer, epoch 0
    assert lib.b200rl_tp_push(_lib.ptr(parts[0]), max_tokens + 1, ptrs, tp, 0, max_tokens, 4, st) == -1
    assert lib.b200rl_tp_push(_lib.ptr(parts[0]), T, ptrs, tp, 0, max_tokens, 0, st) == -1
