"""Companions of the GAE path in the legacy PPO interface, on the GPU: clipped value loss
(ref realhf/impl/model/utils/ppo_functional.py:171-225) and KL-shaped packed rewards (:228-253),
against fixtures made by the real reference and against the oracle on seeded inputs.

Tolerances: loss / per-token loss / gradient within 1e-5 relative (north_star asks 1e-3); the
packed rewards are plain fp32 multiply/add in the reference's order and must match bit for bit;
clip_mask must match exactly."""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import torch_ref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", range(4))
def test_critic_loss_golden(k, cuda):
    from areal_vllm_b200.ppo_functional import critic_loss_fn

    g = load_golden(f"critic_{k}")
    value = torch.from_numpy(g["value"]).to(cuda).requires_grad_(True)
    mask = torch.from_numpy(g["mask"]).to(cuda) if bool(g["use_mask"]) else None
    loss, stat = critic_loss_fn(value, torch.from_numpy(g["old"]).to(cuda),
                                torch.from_numpy(g["target"]).to(cuda), float(g["eps"]),
                                loss_mask=mask, loss_fn_type="huber" if bool(g["huber"]) else "mse")
    (loss * 1.7).backward()
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=1e-5)
    np.testing.assert_allclose(value.grad.cpu().numpy(), g["g_value"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(stat["loss"].cpu().numpy(), g["stat_loss"], rtol=1e-6)
    assert stat["clip_mask"].dtype == torch.bool
    assert np.array_equal(stat["clip_mask"].cpu().numpy(), g["stat_clip_mask"])


@pytest.mark.parametrize("kind", ["mse", "huber"])
@pytest.mark.parametrize("T", [1, 1000, 1 << 20])
@pytest.mark.parametrize("masked", [True, False])
def test_critic_loss_seeded(kind, T, masked, cuda):
    from areal_vllm_b200.ppo_functional import critic_loss_fn

    gen = torch.Generator().manual_seed(T + masked)
    spread = 12.0 if kind == "huber" else 1.0
    old = torch.randn(T, generator=gen) * spread
    value = old + 0.4 * torch.randn(T, generator=gen)
    value[::7] = old[::7] + 0.2  # exactly on the clip edge (bit-identical in fp32 or not: either side)
    value[::11] = old[::11]      # zero step: both branches tie -> same gradient either way
    target = old + torch.randn(T, generator=gen) * spread
    mask = (torch.rand(T, generator=gen) < 0.7) if masked else None
    if masked:
        mask[0] = True
    v_ref = value.clone().requires_grad_(True)
    l_ref, s_ref = torch_ref.critic_loss(v_ref, old, target, 0.25, mask, kind)
    l_ref.backward()
    v = value.to(cuda).requires_grad_(True)
    loss, stat = critic_loss_fn(v, old.to(cuda), target.to(cuda), 0.25,
                                None if mask is None else mask.to(cuda), kind)
    loss.backward()
    np.testing.assert_allclose(loss.item(), l_ref.item(), rtol=1e-5)
    np.testing.assert_allclose(v.grad.cpu().numpy(), v_ref.grad.numpy(), rtol=1e-5, atol=1e-12)
    np.testing.assert_allclose(stat["loss"].cpu().numpy(), s_ref["loss"].numpy(), rtol=1e-6)
    assert torch.equal(stat["clip_mask"].cpu(), s_ref["clip_mask"])


def test_critic_loss_edge_cases(cuda):
    from areal_vllm_b200.ppo_functional import critic_loss_fn

    z = torch.zeros(8, device=cuda)
    # all-false mask: the reference divides by count_nonzero == 0 -> nan
    loss, _ = critic_loss_fn(z, z, z + 1, 0.2, torch.zeros(8, dtype=torch.bool, device=cuda))
    assert torch.isnan(loss)
    # 2-D inputs keep their shape in the stats
    v = torch.randn(4, 6, device=cuda, requires_grad=True)
    loss, stat = critic_loss_fn(v, v.detach() + 0.1, v.detach() - 1, 0.2)
    assert stat["loss"].shape == (4, 6) and stat["clip_mask"].shape == (4, 6)
    loss.backward()
    assert v.grad.shape == (4, 6)
    with pytest.raises(NotImplementedError):
        critic_loss_fn(z, z, z, 0.2, loss_fn_type="l1")
    with pytest.raises(AssertionError):
        critic_loss_fn(z.double(), z, z, 0.2)
    with pytest.raises(RuntimeError):
        critic_loss_fn(z, z[:4], z, 0.2)
    # through the dispatcher (custom op) as well
    from areal_vllm_b200 import ops
    v2 = v.detach().clone().requires_grad_(True)
    out = ops.critic_loss_fwd_op(v2, v.detach() + 0.1, v.detach() - 1, None, 0.2, 0, 10.0)
    out[0].backward()
    assert torch.equal(v2.grad, v.grad)


@pytest.mark.parametrize("k", range(2))
def test_packed_rewards_golden(k, cuda):
    from areal_vllm_b200.ppo_functional import get_packed_rewards

    g = load_golden(f"rewards_{k}")
    kl, tot = get_packed_rewards(float(g["kl_ctl"]), float(g["clip"]),
                                 torch.from_numpy(g["lp"]).to(cuda), torch.from_numpy(g["ref"]).to(cuda),
                                 torch.from_numpy(g["score"]).to(cuda), torch.from_numpy(g["cu"]).to(cuda),
                                 torch.from_numpy(g["no_eos"]).to(cuda),
                                 mask_no_eos_with_zero=bool(g["mask_zero"]))
    assert np.array_equal(kl.cpu().numpy(), g["kl"])
    assert np.array_equal(tot.cpu().numpy(), g["tot"])


@pytest.mark.parametrize("B,maxlen", [(1, 5), (512, 2048), (4096, 3)])
@pytest.mark.parametrize("mz", [False, True])
def test_packed_rewards_seeded(B, maxlen, mz, cuda):
    from areal_vllm_b200.ppo_functional import get_packed_rewards

    gen = torch.Generator().manual_seed(B)
    lens = torch.randint(1, maxlen + 1, (B,), generator=gen)
    cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
    T = int(cu[-1])
    lp = -torch.rand(T, generator=gen) * 3
    ref = lp + 0.3 * torch.randn(T, generator=gen)
    score = torch.randn(B, generator=gen) * 4
    no_eos = torch.rand(B, generator=gen) < 0.3
    kl_r, tot_r = torch_ref.packed_rewards(0.07, 2.5, lp, ref, score, cu, no_eos, mz)
    kl, tot = get_packed_rewards(0.07, 2.5, lp.to(cuda), ref.to(cuda), score.to(cuda), cu.to(cuda),
                                 no_eos.to(cuda), mask_no_eos_with_zero=mz)
    assert torch.equal(kl.cpu(), kl_r)
    assert torch.equal(tot.cpu(), tot_r)


def test_rewards_into_gae_pipeline(cuda):
    """The legacy interface's order: packed rewards -> GAE -> critic loss on the returns."""
    from areal_vllm_b200.ppo_functional import (critic_loss_fn, get_packed_advantages_and_returns,
                                                get_packed_rewards)
    from oracle import cpu_ref

    gen = torch.Generator().manual_seed(5)
    B = 64
    lens = torch.randint(2, 300, (B,), generator=gen)
    cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
    T = int(cu[-1])
    lp = -torch.rand(T, generator=gen)
    ref = lp + 0.1 * torch.randn(T, generator=gen)
    score = torch.randn(B, generator=gen)
    no_eos = torch.rand(B, generator=gen) < 0.3
    values = torch.randn(T + B, generator=gen)
    _, tot = get_packed_rewards(0.1, 5.0, lp.to(cuda), ref.to(cuda), score.to(cuda), cu.to(cuda),
                                no_eos.to(cuda))
    adv, ret = get_packed_advantages_and_returns(0.99, 0.95, values.to(cuda), tot, cu.to(cuda),
                                                 no_eos.to(cuda))
    _, tot_r = torch_ref.packed_rewards(0.1, 5.0, lp, ref, score, cu, no_eos)
    adv_r, ret_r = cpu_ref.gae_py(tot_r.numpy(), values.numpy(), cu.numpy(), no_eos.numpy(), 0.99, 0.95)
    np.testing.assert_allclose(adv.cpu().numpy(), adv_r, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ret.cpu().numpy(), ret_r, rtol=1e-5, atol=1e-5)
    # values without the bootstrap slot of each sequence
    keep = torch.ones(T + B, dtype=torch.bool)
    keep[(cu[1:].long() + torch.arange(B))] = False
    v = values[keep].to(cuda).requires_grad_(True)
    loss, _ = critic_loss_fn(v, v.detach() + 0.05, ret, 0.2)
    loss.backward()
    assert torch.isfinite(loss) and torch.isfinite(v.grad).all()
