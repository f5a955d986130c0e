"""compute_packed_sft_loss on the GPU (ref areal/engine/sft/lm_engine.py:45-83) against fixtures
made by the real reference (incl. everything it registers with stats_tracker) and against the
oracle on seeded inputs. Tolerances: loss, per-token loss, perplexity within 1e-3 relative
(north_star; observed ~1e-6), gradient within one ulp of the logits dtype, row min/max exact."""

import numpy as np
import pytest
import torch

from conftest import golden_logits, load_golden
from oracle import torch_ref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", range(2))
def test_sft_loss_golden(k, cuda):
    from areal_vllm_b200.lm_engine import compute_packed_sft_loss

    g = load_golden(f"sft_{k}")
    logits, dref = golden_logits(g)
    x = logits.to(cuda).requires_grad_(True)
    stats = {}
    loss = compute_packed_sft_loss(
        x, dict(input_ids=torch.from_numpy(g["input_ids"]).to(cuda), cu_seqlens=torch.from_numpy(g["cu"]).to(cuda),
                loss_mask=torch.from_numpy(g["loss_mask"]).to(cuda)), stats=stats)
    (loss * 2.5).backward()
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=1e-5)
    tol = 2e-5 if logits.dtype == torch.float32 else 8e-3
    np.testing.assert_allclose(x.grad.float().cpu().numpy(), dref.float().numpy(), rtol=tol, atol=1e-7)
    np.testing.assert_allclose(stats["ppl"].cpu().numpy(), g["stat_ppl"], rtol=1e-4)
    assert stats["ppl"].dtype == torch.float32
    np.testing.assert_allclose(stats["loss"].cpu().numpy(), g["stat_loss"], rtol=1e-5, atol=1e-6)
    assert np.array_equal(stats["vocab_min_logits"].cpu().numpy(), g["stat_vocab_min_logits"])
    assert np.array_equal(stats["vocab_max_logits"].cpu().numpy(), g["stat_vocab_max_logits"])
    assert np.array_equal(stats["n_valid_tokens"].cpu().numpy(), g["denom_n_valid_tokens"])


@pytest.mark.parametrize("B,maxlen,V", [(1, 9, 512), (37, 300, 4096), (600, 40, 1024)])
def test_sft_loss_seeded(B, maxlen, V, cuda):
    from areal_vllm_b200.lm_engine import compute_packed_sft_loss

    gen = torch.Generator().manual_seed(B)
    lens = torch.randint(3, maxlen + 1, (B,), generator=gen)
    cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
    T = int(cu[-1])
    logits = (torch.randn(T, V, generator=gen) * 2).bfloat16()
    ids = torch.randint(0, V, (T,), generator=gen)
    mask = (torch.rand(T, generator=gen) < 0.6).int()
    mask[1] = 1
    if B > 2:  # a sequence with no valid token after the roll: 0/0 = nan in both
        mask[int(cu[1]) + 1: int(cu[2]) + 1] = 0
    x_ref = logits.clone().requires_grad_(True)
    l_ref, s_ref = torch_ref.packed_sft_loss(x_ref, ids, cu, mask)
    l_ref.backward()
    x = logits.to(cuda).requires_grad_(True)
    stats = {}
    loss = compute_packed_sft_loss(x, dict(input_ids=ids.to(cuda), cu_seqlens=cu.to(cuda),
                                           loss_mask=mask.to(cuda)), stats=stats)
    loss.backward()
    np.testing.assert_allclose(loss.item(), l_ref.item(), rtol=1e-5)
    np.testing.assert_allclose(stats["seqlogp"].cpu().numpy(), s_ref["seqlogp"].numpy(), rtol=1e-5,
                               equal_nan=True)
    np.testing.assert_allclose(stats["ppl"].cpu().numpy(), s_ref["ppl"].numpy(), rtol=1e-4, equal_nan=True)
    if B > 2:
        assert torch.isnan(stats["seqlogp"][1])
    np.testing.assert_allclose(stats["loss"].cpu().numpy(), s_ref["loss"].numpy(), rtol=1e-5, atol=1e-6)
    assert torch.equal(stats["vocab_max_logits"].cpu(), s_ref["vocab_max_logits"])
    assert torch.equal(stats["vocab_min_logits"].cpu(), s_ref["vocab_min_logits"])
    np.testing.assert_allclose(x.grad.float().cpu().numpy(), x_ref.grad.float().numpy(), rtol=8e-3, atol=1e-7)


def test_sft_loss_all_masked_and_errors(cuda):
    from areal_vllm_b200 import ops
    from areal_vllm_b200.lm_engine import compute_packed_sft_loss

    logits = torch.randn(10, 256, device=cuda).bfloat16()
    inp = dict(input_ids=torch.zeros(10, dtype=torch.long, device=cuda),
               cu_seqlens=torch.tensor([0, 4, 10], dtype=torch.int32, device=cuda),
               loss_mask=torch.zeros(10, dtype=torch.int32, device=cuda))
    assert torch.isnan(compute_packed_sft_loss(logits, inp))  # reference: 0 / 0
    lp = torch.zeros(10, device=cuda)
    with pytest.raises(RuntimeError):
        ops.sft_loss_fwd(lp.double(), inp["loss_mask"], inp["cu_seqlens"])
    with pytest.raises(RuntimeError):
        ops.sft_loss_fwd(lp, inp["loss_mask"][:5], inp["cu_seqlens"])
    with pytest.raises(RuntimeError):
        ops.sft_loss_fwd(lp, inp["loss_mask"], inp["cu_seqlens"].float())
    # dispatcher path
    lp2 = torch.randn(10, device=cuda, requires_grad=True)
    m = torch.ones(10, dtype=torch.bool, device=cuda)
    out = ops.sft_loss_fwd_op(lp2, m, inp["cu_seqlens"])
    out[0].backward()
    np.testing.assert_allclose(lp2.grad.cpu().numpy(), np.full(10, -0.1, np.float32), rtol=1e-6)


@pytest.mark.parametrize("k", range(2))
def test_sft_fused_golden(k, cuda):
    """the one-pass SFT step against the fixture of the real compute_packed_sft_loss + backward"""
    from areal_vllm_b200.lm_engine import compute_packed_sft_loss_fused

    g = load_golden(f"sft_{k}")
    logits, dref = golden_logits(g)
    x = logits.to(cuda).requires_grad_(True)
    stats = {}
    inp = dict(input_ids=torch.from_numpy(g["input_ids"]).to(cuda), cu_seqlens=torch.from_numpy(g["cu"]).to(cuda),
               loss_mask=torch.from_numpy(g["loss_mask"]).to(cuda))
    loss = compute_packed_sft_loss_fused(x, inp, loss_scale=2.5, stats=stats)  # fixture: (loss * 2.5).backward()
    loss.backward()
    np.testing.assert_allclose(loss.item() / 2.5, g["loss"], rtol=1e-5)
    tol = 2e-5 if logits.dtype == torch.float32 else 8e-3
    np.testing.assert_allclose(x.grad.float().cpu().numpy(), dref.float().numpy(), rtol=tol, atol=1e-7)
    np.testing.assert_allclose(stats["ppl"].cpu().numpy(), g["stat_ppl"], rtol=1e-4)
    np.testing.assert_allclose(stats["loss"].cpu().numpy(), g["stat_loss"], rtol=1e-5, atol=1e-6)
    assert np.array_equal(stats["vocab_min_logits"].cpu().numpy(), g["stat_vocab_min_logits"])
    assert np.array_equal(stats["vocab_max_logits"].cpu().numpy(), g["stat_vocab_max_logits"])


def test_sft_fused_equals_separate(cuda):
    from areal_vllm_b200.lm_engine import compute_packed_sft_loss, compute_packed_sft_loss_fused

    gen = torch.Generator().manual_seed(8)
    T, V = 700, 8192
    logits = (torch.randn(T, V, generator=gen) * 2).bfloat16().to(cuda)
    inp = dict(input_ids=torch.randint(0, V, (T,), generator=gen).to(cuda),
               cu_seqlens=torch.tensor([0, 100, 350, 700], dtype=torch.int32, device=cuda),
               loss_mask=(torch.rand(T, generator=gen) < 0.6).int().to(cuda))
    a = logits.clone().requires_grad_(True)
    sa = {}
    la = compute_packed_sft_loss(a, inp, stats=sa)
    (la * 0.37).backward()
    b = logits.clone().requires_grad_(True)
    sb = {}
    lb = compute_packed_sft_loss_fused(b, inp, loss_scale=torch.tensor(0.37, device=cuda), stats=sb)
    lb.backward()
    np.testing.assert_allclose(lb.item(), la.item() * 0.37, rtol=1e-5)
    np.testing.assert_allclose(b.grad.float().cpu().numpy(), a.grad.float().cpu().numpy(), rtol=8e-3, atol=1e-9)
    for key in ("ppl", "loss", "seqlogp", "vocab_min_logits", "vocab_max_logits"):
        np.testing.assert_allclose(sb[key].float().cpu().numpy(), sa[key].float().cpu().numpy(), rtol=1e-5, atol=1e-6, err_msg=key)
    # no scale given and a different upstream gradient: still right (rescaled in place)
    c = logits.clone().requires_grad_(True)
    lc = compute_packed_sft_loss_fused(c, inp)
    (lc * 0.37).backward()
    np.testing.assert_allclose(c.grad.float().cpu().numpy(), a.grad.float().cpu().numpy(), rtol=8e-3, atol=1e-9)
    # all-false mask: nan like the reference (no `or 1`)
    inp0 = dict(inp, loss_mask=torch.zeros(T, dtype=torch.int32, device=cuda))
    d = logits.clone().requires_grad_(True)
    assert torch.isnan(compute_packed_sft_loss_fused(d, inp0))
