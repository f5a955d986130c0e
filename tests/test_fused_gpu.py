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


def test_fused_full_size_cfg2_microbatch(cuda):
    """One full GRPO micro-batch of BASELINE config 2: 10,240 tokens x V=151,936 bf16 (3.1 GB).
    The one-pass kernel against the separate drop-in path on the whole tensor (loss, every stat,
    dlogits within one ulp of bf16), and a row subset of dlogits against the oracle's autograd."""
    from areal_vllm_b200 import actor
    from oracle import torch_ref

    T, V = 10240, 151936
    free, _ = torch.cuda.mem_get_info()
    if free < 16 * 2**30:
        pytest.skip("needs ~13 GB of free HBM")
    gen = torch.Generator(device=cuda).manual_seed(2)
    x = torch.empty(T, V, dtype=torch.bfloat16, device=cuda)
    for lo in range(0, T, 2048):
        x[lo:lo + 2048] = (torch.randn(2048, V, generator=gen, device=cuda) * 2).bfloat16()
    ids = torch.randint(0, V, (T,), generator=gen, device=cuda)
    labels = torch.roll(ids, -1)
    rows = torch.arange(0, T, 2, device=cuda)
    x[rows, labels[rows]] += 6.0
    lp0 = torch_ref.logprobs_entropy(x[:64], labels[:64], 1.0)[0].mean().item()
    data = dict(input_ids=ids, logprobs=torch.full((T,), lp0, device=cuda) + 0.05 * torch.randn(T, generator=gen, device=cuda),
                advantages=torch.randn(T, generator=gen, device=cuda),
                loss_mask=torch.rand(T, generator=gen, device=cuda) < 0.8,
                prox_logp=torch.full((T,), lp0, device=cuda) + 0.02 * torch.randn(T, generator=gen, device=cuda))
    a = x.detach().requires_grad_(True)
    sa = {}
    la = actor.grpo_loss_fn(a, data, 1.0, 0.4, None, None, 5.0, stats=sa)
    (ga,) = torch.autograd.grad(la, a)
    b = x.detach().requires_grad_(True)
    sb = {}
    lb = actor.grpo_loss_fn_fused(b, data, 1.0, 0.4, None, None, 5.0, stats=sb)
    (gb,) = torch.autograd.grad(lb, b)
    np.testing.assert_allclose(lb.item(), la.item(), rtol=1e-5)
    for k in sa:
        if sa[k].dtype == torch.bool:
            assert torch.equal(sa[k], sb[k]), k
        else:
            assert torch.allclose(sa[k].float(), sb[k].float(), rtol=1e-5, atol=1e-6), k
    diff = (ga.float() - gb.float()).abs()
    tol = 8e-3 * ga.float().abs() + 1e-9
    assert bool((diff <= tol).all())
    del diff, tol
    # rows of the gradient sum to ~0 (softmax minus one-hot) and a subset matches the oracle
    pick = torch.randperm(T, generator=gen, device=cuda)[:64].sort().values
    xs = x.detach()[pick].clone().requires_grad_(True)
    lp_r, _ = torch_ref.logprobs_entropy(xs, labels[pick], 1.0)
    # d loss / d logp for the picked rows from the separate path's autograd
    full_lp = sa["new_logp"].clone()
    from areal_vllm_b200 import functional as BF
    lp_all = full_lp.clone().requires_grad_(True)
    loss_all, _ = BF.ppo_actor_loss_fn(lp_all, data["prox_logp"], data["logprobs"], data["advantages"], 0.4,
                                       data["loss_mask"], behav_imp_weight_cap=5.0)
    (g_lp,) = torch.autograd.grad(loss_all, lp_all)
    assert float(gb[pick].float().sum(-1).abs().max()) < 2e-2 * float(g_lp.abs().max())  # bf16 rounding only
    (lp_r * g_lp[pick]).sum().backward()
    np.testing.assert_allclose(gb[pick].float().cpu().numpy(), xs.grad.float().cpu().numpy(), rtol=8e-3, atol=1e-9)


def test_fused_second_backward_is_refused(cuda):
    """The fused node hands out (and rescales) its stored gradient in place, so a second pass
    through it would scale twice: it must raise instead of returning silently wrong gradients."""
    from areal_vllm_b200.actor import grpo_loss_fn_fused

    logits, data = _batch(64, 4096, 9)
    dd = {k: v.to(cuda) for k, v in data.items()}
    xg = logits.to(cuda).requires_grad_(True)
    loss = grpo_loss_fn_fused(xg, dd, 1.0, 0.2, None, None, 5.0)
    (loss * 2.0).backward(retain_graph=True)
    first = xg.grad.clone()
    with pytest.raises(RuntimeError, match="already consumed"):
        (loss * 2.0).backward()
    assert torch.equal(xg.grad, first)


def test_fused_pair_mode_bit_identical_to_one_row_per_cta(cuda):
    """The default launch runs clusters of two CTAs sharing two row streams (csrc/grpo_fused.cu,
    pair mode); B200RL_FUSED_PAIR=0 selects the one-row-per-CTA form. Same arithmetic in the same
    order: every output of every shape must agree bit for bit (the mode is read once per process,
    hence the two subprocesses)."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for mode in ("0", "1"):
        env = dict(os.environ, B200RL_FUSED_PAIR=mode)
        r = subprocess.run([sys.executable, os.path.join(root, "scripts", "fused_pair_check.py")], env=env,
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout)
    assert outs[0].count("\n") >= 7 and outs[0] == outs[1]
