"""The torch.library custom ops (what a compiled graph sees) against the eager fast path, and the
whole drop-in path under torch.compile(backend="eager") (exercises the fake impls + autograd
registration without Inductor code generation)."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(cuda, T=96, V=2048):
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(T, V, generator=g) * 2).bfloat16().to(cuda)
    y = torch.randint(0, V, (T,), generator=g).to(cuda)
    prox = -torch.rand(T, generator=g).to(cuda) * 3
    old = prox + 0.1 * torch.randn(T, generator=g).to(cuda)
    adv = torch.randn(T, generator=g).to(cuda)
    mask = (torch.rand(T, generator=g) < 0.7).to(cuda)
    return x, y, prox, old, adv, mask


def test_custom_ops_match_eager_path(cuda):
    from areal_vllm_b200 import ops

    x, y, prox, old, adv, mask = _inputs(cuda)
    a = ops.logprob_entropy_fwd_op(x, y, 0.9, True, True)
    b = ops.logprob_entropy_fwd(x, y, 0.9, True, True)
    for u, v in zip(a, b):
        assert torch.equal(u, v)
    g = torch.randn_like(a[0])
    d1 = ops.logprob_entropy_bwd_op(x, y, a[2], a[1], g, 0.1 * g, 0.9)
    d2 = ops.logprob_entropy_bwd(x, y, a[2], a[1], g, 0.1 * g, 0.9)
    assert torch.equal(d1, d2)
    l1 = ops.ppo_loss_fwd_op(a[0], prox, old, adv, mask, 0.8, 1.2, 3.0, 5.0, None, None, None)
    l2 = ops.ppo_loss_fwd(a[0], prox, old, adv, mask, 0.8, 1.2, 3.0, 5.0, None, None, None)
    for u, v in zip(l1, l2):
        assert torch.equal(u, v)
    # autograd through the custom ops == autograd through the eager nodes
    for fwd, loss in ((ops.logprob_entropy_fwd_op, ops.ppo_loss_fwd_op), (ops.logprob_entropy_fwd, ops.ppo_loss_fwd)):
        xg = x.clone().requires_grad_(True)
        lp, ent, *_ = fwd(xg, y, 0.9, True, False)
        out = loss(lp, prox, old, adv, mask, 0.8, 1.2, -1.0, -1.0, None, None, None)
        (out[0] * 2 + 0.01 * ent.sum()).backward()
        if fwd is ops.logprob_entropy_fwd_op:
            ref = xg.grad.clone()
        else:
            assert torch.equal(xg.grad, ref)


def test_opcheck(cuda):
    from areal_vllm_b200 import ops

    x, y, prox, old, adv, mask = _inputs(cuda, T=8, V=256)
    torch.library.opcheck(ops.logprob_entropy_fwd_op, (x, y, 1.0, True, False),
                          test_utils=("test_schema", "test_faketensor"))
    lp = ops.logprob_entropy_fwd(x, y, 1.0, False, False)[0]
    torch.library.opcheck(ops.ppo_loss_fwd_op, (lp, prox, old, adv, mask, 0.8, 1.2, -1.0, 5.0, None, None, None),
                          test_utils=("test_schema", "test_faketensor"))
    r = torch.randn(50, device=cuda); v = torch.randn(53, device=cuda)
    cu = torch.tensor([0, 10, 30, 50], dtype=torch.int32, device=cuda)
    b = torch.ones(3, dtype=torch.bool, device=cuda)
    torch.library.opcheck(ops.gae_1d_nolp_misalign_op, (r, v, cu, b, 0.99, 0.95, 0),
                          test_utils=("test_schema", "test_faketensor"))


def test_dropin_under_torch_compile_eager_backend(cuda):
    from areal_vllm_b200 import functional as BF

    x, y, prox, old, adv, mask = _inputs(cuda)

    def step(logits):
        lp, ent = BF.gather_logprobs_entropy(logits, y, 1.0)
        loss, stat = BF.ppo_actor_loss_fn(lp, prox, old, adv, 0.2, mask, c_clip=3.0, behav_imp_weight_cap=5.0)
        return loss, ent, stat["clip_mask"]

    xe = x.clone().requires_grad_(True)
    le, ee, ce = step(xe)
    le.backward()
    xc = x.clone().requires_grad_(True)
    lc, ec, cc = torch.compile(step, backend="eager", fullgraph=True)(xc)
    lc.backward()
    assert torch.equal(le, lc) and torch.equal(ee, ec) and torch.equal(ce, cc)
    assert torch.equal(xe.grad, xc.grad)


def test_opcheck_new_ops(cuda):
    from areal_vllm_b200 import ops

    g = torch.Generator().manual_seed(2)
    v = torch.randn(64, generator=g).to(cuda)
    m = (torch.rand(64, generator=g) < 0.7).to(cuda)
    torch.library.opcheck(ops.critic_loss_fwd_op, (v, v + 0.1, v - 1.0, m, 0.2, 0, 10.0),
                          test_utils=("test_schema", "test_faketensor"))
    torch.library.opcheck(ops.critic_loss_fwd_op, (v, v + 0.1, v - 1.0, None, 0.2, 1, 10.0),
                          test_utils=("test_schema", "test_faketensor"))
    cu = torch.tensor([0, 10, 30, 64], dtype=torch.int32, device=cuda)
    torch.library.opcheck(ops.sft_loss_fwd_op, (v, m, cu), test_utils=("test_schema", "test_faketensor"))
    sc = torch.randn(3, device=cuda)
    ne = torch.tensor([True, False, True], device=cuda)
    torch.library.opcheck(ops.packed_rewards_op, (v, v + 0.1, sc, cu, ne, 0.1, 5.0, False),
                          test_utils=("test_schema", "test_faketensor"))


def test_new_dropins_under_torch_compile_eager_backend(cuda):
    from areal_vllm_b200 import ppo_functional as PF
    from areal_vllm_b200.lm_engine import compute_packed_sft_loss

    x, y, prox, old, adv, mask = _inputs(cuda)
    cu = torch.tensor([0, 40, 96], dtype=torch.int32, device=cuda)

    def step(logits, value):
        l1 = compute_packed_sft_loss(logits, dict(input_ids=y, cu_seqlens=cu, loss_mask=mask))
        l2, st = PF.critic_loss_fn(value, old, adv, 0.2, mask, "huber")
        return l1 + l2, st["clip_mask"]

    xe, ve = x.clone().requires_grad_(True), prox.clone().requires_grad_(True)
    le, ce = step(xe, ve)
    le.backward()
    xc, vc = x.clone().requires_grad_(True), prox.clone().requires_grad_(True)
    lc, cc = torch.compile(step, backend="eager", fullgraph=True)(xc, vc)
    lc.backward()
    assert torch.equal(le, lc) and torch.equal(ce, cc)
    assert torch.equal(xe.grad, xc.grad) and torch.equal(ve.grad, vc.grad)


def test_cuda_graph_capture_and_replay(cuda):
    """No host synchronisation and no allocation outside torch's allocator on the loss path: a
    forward+backward captured in a CUDA graph replays on new data with the eager result."""
    from areal_vllm_b200 import actor

    x, y, prox, old, adv, mask = _inputs(cuda, T=200, V=4096)
    data = dict(input_ids=y.clone(), logprobs=old.clone(), advantages=adv.clone(),
                loss_mask=mask.clone(), prox_logp=prox.clone())
    static_x = x.clone().requires_grad_(True)

    def step():
        static_x.grad = None
        loss = actor.grpo_loss_fn(static_x, data, 1.0, 0.2, None, 3.0, 5.0)
        loss.backward()
        return loss

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):  # warm-up: kernel attributes, allocator
            step()
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        static_loss = step()
        static_grad = static_x.grad
    # new inputs, same shapes
    g = torch.Generator().manual_seed(99)
    new_x = (torch.randn(200, 4096, generator=g) * 2).bfloat16().to(cuda)
    new_adv = torch.randn(200, generator=g).to(cuda)
    with torch.no_grad():
        static_x.copy_(new_x)
        data["advantages"].copy_(new_adv)
    graph.replay()
    torch.cuda.synchronize()
    xe = new_x.clone().requires_grad_(True)
    le = actor.grpo_loss_fn(xe, data, 1.0, 0.2, None, 3.0, 5.0)
    le.backward()
    assert torch.equal(static_loss, le.detach())
    assert torch.equal(static_grad, xe.grad)
