"""The fused GRPO step (one pass over HBM: forward statistics, loss, and dlogits out of L2) against
the separate forward / loss / backward path of the same library and against the CPU oracle."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(T, V, seed, dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(T, V, generator=g) * 2).to(dtype)
    ids = torch.randint(0, V, (T,), generator=g)
    data = dict(input_ids=ids, logprobs=-torch.rand(T, generator=g) * 6, advantages=torch.randn(T, generator=g),
                loss_mask=(torch.rand(T, generator=g) < 0.7).float(), prox_logp=-torch.rand(T, generator=g) * 6)
    return logits, data


@pytest.mark.parametrize("T,V", [(1, 64), (37, 4104), (300, 32768), (333, 151936)])
@pytest.mark.parametrize("kw", [dict(eps_clip=0.4, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=5.0),
                                dict(eps_clip=0.2, eps_clip_higher=0.28, c_clip=3.0, behav_imp_weight_cap=None)])
def test_fused_equals_separate_path(T, V, kw, cuda):
    from areal_vllm_b200.actor import grpo_loss_fn, grpo_loss_fn_fused

    logits, data = _batch(T, V, 17 + T)
    dd = {k: v.to(cuda) for k, v in data.items()}
    scale = torch.tensor(0.37, device=cuda)
    a = logits.to(cuda).requires_grad_(True)
    sa = {}
    la = grpo_loss_fn(a, dd, 0.9, kw["eps_clip"], kw["eps_clip_higher"], kw["c_clip"], kw["behav_imp_weight_cap"], stats=sa)
    (la * scale).backward()
    b = logits.to(cuda).requires_grad_(True)
    sb = {}
    lb = grpo_loss_fn_fused(b, dd, 0.9, kw["eps_clip"], kw["eps_clip_higher"], kw["c_clip"],
                            kw["behav_imp_weight_cap"], loss_scale=scale, stats=sb)
    lb.backward()
    # same arithmetic, but the row statistics are reduced over a different warp partition (16 vs
    # 24 warps): log-probs agree to an ulp or two, gradients to one ulp of bf16
    np.testing.assert_allclose(lb.item(), (la * scale).item(), rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(b.grad.float().cpu().numpy(), a.grad.float().cpu().numpy(), rtol=2.0 ** -7, atol=1e-7)
    assert set(sa) == set(sb)
    for k in sa:
        assert sa[k].dtype == sb[k].dtype and sa[k].shape == sb[k].shape, k
        if sa[k].dtype == torch.bool:
            assert torch.equal(sa[k], sb[k]), k
        else:
            np.testing.assert_allclose(sb[k].cpu().numpy(), sa[k].cpu().numpy(), rtol=1e-5, atol=2e-6, err_msg=k)


def test_fused_vs_oracle_and_other_upstream_gradients(cuda):
    from areal_vllm_b200.actor import grpo_loss_fn_fused
    from oracle import torch_ref

    T, V = 150, 8192
    logits, data = _batch(T, V, 5)
    dd = {k: v.to(cuda) for k, v in data.items()}
    xc = logits.clone().requires_grad_(True)
    lp, ent = torch_ref.logprobs_entropy(xc, torch.roll(data["input_ids"], -1), 1.0)
    loss_r, _ = torch_ref.ppo_actor_loss(lp, data["prox_logp"], data["logprobs"], data["advantages"], 0.2,
                                         data["loss_mask"].bool(), 0.28, 3.0, 5.0)
    (loss_r * 2.5).backward()
    # loss_scale given up front (the fast case), plus a non-unit gradient arriving at backward()
    xg = logits.to(cuda).requires_grad_(True)
    loss = grpo_loss_fn_fused(xg, dd, 1.0, 0.2, 0.28, 3.0, 5.0, loss_scale=torch.tensor(0.5, device=cuda))
    (loss * 5.0).backward()  # total factor 2.5: the stored gradient is rescaled in place by 5
    np.testing.assert_allclose(loss.item(), 0.5 * loss_r.item(), rtol=1e-3, atol=1e-7)
    np.testing.assert_allclose(xg.grad.float().cpu().numpy(), xc.grad.float().numpy(), rtol=2.0 ** -6, atol=1e-6)
    # no loss_scale at all
    xg2 = logits.to(cuda).requires_grad_(True)
    loss2 = grpo_loss_fn_fused(xg2, dd, 1.0, 0.2, 0.28, 3.0, 5.0)
    loss2.backward()
    np.testing.assert_allclose(loss2.item(), loss_r.item(), rtol=1e-3, atol=1e-7)
    np.testing.assert_allclose(xg2.grad.float().cpu().numpy() * 2.5, xc.grad.float().numpy(), rtol=2.0 ** -6, atol=1e-6)


def test_fused_fallbacks_and_dtypes(cuda):
    from areal_vllm_b200.actor import grpo_loss_fn, grpo_loss_fn_fused

    # unaligned vocabulary -> falls back to the separate path, same numbers
    logits, data = _batch(20, 1001, 3)
    dd = {k: v.to(cuda) for k, v in data.items()}
    a = logits.to(cuda).requires_grad_(True)
    b = logits.to(cuda).requires_grad_(True)
    la = grpo_loss_fn(a, dd, 1.0, 0.2, None, None, None)
    lb = grpo_loss_fn_fused(b, dd, 1.0, 0.2, None, None, None)
    la.backward(); lb.backward()
    assert torch.equal(la, lb) and torch.equal(a.grad, b.grad)
    # fp32 and fp16 logits through the fused kernel
    for dtype in (torch.float32, torch.float16):
        logits, data = _batch(33, 2048, 4, dtype)
        dd = {k: v.to(cuda) for k, v in data.items()}
        a = logits.to(cuda).requires_grad_(True)
        b = logits.to(cuda).requires_grad_(True)
        la = grpo_loss_fn(a, dd, 1.0, 0.2, None, 2.0, 3.0)
        lb = grpo_loss_fn_fused(b, dd, 1.0, 0.2, None, 2.0, 3.0)
        la.backward(); lb.backward()
        np.testing.assert_allclose(lb.item(), la.item(), rtol=1e-5)
        assert b.grad.dtype == dtype
        np.testing.assert_allclose(b.grad.float().cpu().numpy(), a.grad.float().cpu().numpy(),
                                   rtol=1e-4 if dtype == torch.float32 else 2.0 ** -10, atol=1e-7)
    # no grad requested -> plain forward
    logits, data = _batch(9, 512, 6)
    with torch.no_grad():
        l = grpo_loss_fn_fused(logits.to(cuda), {k: v.to(cuda) for k, v in data.items()}, 1.0, 0.2, None, None, None)
    assert l.dim() == 0
