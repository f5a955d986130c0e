"""Oracle vs the LIVE reference (imported from /root/reference) on fresh seeds. Runs only in the
build container; on the GPU box the tree does not exist and these tests skip. Complements the
committed fixtures (tests/golden) with inputs the fixtures did not see."""

import numpy as np
import pytest
import torch

from oracle import cpu_ref, ref_loader as R, torch_ref

pytestmark = pytest.mark.skipif(not R.available(), reason="reference tree not present")
needs_realhf = pytest.mark.skipif(not R.realhf_available(), reason="realhf tree not present (pip-installed areal only)")


@pytest.mark.parametrize("seed", range(4))
def test_h1_h2_fuzz(seed):
    F = R.areal_functional()
    gen = torch.Generator().manual_seed(500 + seed)
    T, V = 40 + seed, 3000 + 17 * seed
    temp = [1.0, 0.7, 1.5, 0.3][seed]
    x = (torch.randn(T, V, generator=gen) * 3).bfloat16()
    y = torch.randint(0, V, (T,), generator=gen)
    a = x.clone().requires_grad_(True)
    b = x.clone().requires_grad_(True)
    lp_r, ent_r = F.gather_logprobs_entropy(a, y, temp)
    lp_o, ent_o = torch_ref.logprobs_entropy(b, y, temp)
    assert torch.equal(lp_r, lp_o) and torch.equal(ent_r, ent_o)
    prox = lp_r.detach() + 0.2 * torch.randn(T, generator=gen)
    old = prox + 0.2 * torch.randn(T, generator=gen)
    adv = torch.randn(T, generator=gen)
    mask = torch.rand(T, generator=gen) < 0.6
    kw = [dict(), dict(c_clip=2.0), dict(eps_clip_higher=0.3, behav_imp_weight_cap=1.2),
          dict(eps_clip_higher=0.25, c_clip=3.0, behav_imp_weight_cap=5.0)][seed]
    l_r, s_r = F.ppo_actor_loss_fn(lp_r, prox, old, adv, 0.2, mask, **kw)
    l_o, s_o = torch_ref.ppo_actor_loss(lp_o, prox, old, adv, 0.2, mask, **kw)
    (l_r + 0.01 * ent_r.mean()).backward()
    (l_o + 0.01 * ent_o.mean()).backward()
    assert torch.equal(l_r, l_o) and torch.equal(a.grad, b.grad)
    for k in s_r:
        assert torch.equal(s_r[k], s_o[k]), k


@needs_realhf
@pytest.mark.parametrize("seed", range(3))
def test_gae_fuzz(seed):
    P = R.realhf_ppo_functional()
    rng = np.random.default_rng(700 + seed)
    B = 5 + seed
    lens = rng.integers(1, 200, B)
    total = int(lens.sum())
    r = rng.standard_normal(total).astype(np.float32)
    v = rng.standard_normal(total + B).astype(np.float32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    b = (rng.random(B) < 0.5).astype(np.float32)
    gamma, lam = [(1.0, 1.0), (0.99, 0.95), (0.9, 1.0)][seed]
    adv, ret = P.pygae1d_nolp_misalign(torch.from_numpy(r), torch.from_numpy(v.copy()), torch.from_numpy(cu),
                                       torch.from_numpy(b), gamma, lam)
    a_o, r_o = cpu_ref.gae_py(r, v, cu, b, gamma, lam)
    assert np.array_equal(adv.numpy(), a_o) and np.array_equal(ret.numpy(), r_o)


def test_advnorm_fuzz():
    AdvNorm = R.adv_norm_class()

    class Cfg:
        def __init__(self, m, s, g):
            self.mean_level, self.std_level, self.group_size = m, s, g

    gen = torch.Generator().manual_seed(9)
    x = torch.randn(24, 7, generator=gen) * 2 - 1
    mask = (torch.rand(24, 7, generator=gen) < 0.7).float()
    for m in ("batch", "group", "none"):
        for s in ("batch", "group", "none"):
            ref = AdvNorm(Cfg(m, s, 8))(x.clone(), mask.clone())
            out = torch_ref.adv_norm(x, mask, m, s, 8)
            assert torch.equal(ref, out), (m, s)


@needs_realhf
def test_intervals_and_merge_fuzz():
    fp = R.flatten_param_py()
    merge_ref = R.merge_intervals_ext()
    rng = np.random.default_rng(3)
    size = 5000
    starts = np.sort(rng.choice(size - 50, 60, replace=False))
    iv = [(int(s), int(s) + int(rng.integers(0, min(40, (starts[i + 1] if i + 1 < len(starts) else size) - s + 1))))
          for i, s in enumerate(starts)]
    rng.shuffle(iv)
    src = torch.from_numpy(rng.integers(-32768, 32767, size, dtype=np.int16)).view(torch.bfloat16)
    got = fp["_slice_intervals_py"](src, iv)
    mine = cpu_ref.slice_intervals(src.view(torch.int16).numpy().view(np.uint16), np.array(iv))
    assert np.array_equal(got.contiguous().view(torch.int16).numpy().view(np.uint16), mine)
    dst = torch.zeros(size, dtype=torch.bfloat16)
    fp["_set_intervals_py"](got, dst, iv)
    d2 = np.zeros(size, np.uint16)
    cpu_ref.set_intervals(mine, d2, np.array(iv))
    assert np.array_equal(dst.view(torch.int16).numpy().view(np.uint16), d2)
    if merge_ref is not None:
        pairs = [(int(min(a, b)), int(max(a, b))) for a, b in rng.integers(0, 30, size=(200, 2))]
        assert cpu_ref.merge_intervals(pairs) == merge_ref(pairs)


@needs_realhf
@pytest.mark.parametrize("seed", range(4))
def test_critic_rewards_fuzz(seed):
    P = R.realhf_ppo_functional()
    gen = torch.Generator().manual_seed(700 + seed)
    T = 200 + 13 * seed
    kind = ["mse", "huber", "huber", "mse"][seed]
    spread = 12.0 if kind == "huber" else 1.0
    old = torch.randn(T, generator=gen) * spread
    val = old + 0.5 * torch.randn(T, generator=gen)
    tgt = old + torch.randn(T, generator=gen) * spread
    mask = (torch.rand(T, generator=gen) < 0.5) if seed % 2 == 0 else None
    a, b = val.clone().requires_grad_(True), val.clone().requires_grad_(True)
    l_r, s_r = P.critic_loss_fn(a, old, tgt, 0.3, loss_mask=mask, loss_fn_type=kind)
    l_o, s_o = torch_ref.critic_loss(b, old, tgt, 0.3, mask, kind)
    l_r.backward()
    l_o.backward()
    assert torch.equal(l_r, l_o) and torch.equal(a.grad, b.grad)
    assert torch.equal(s_r["loss"], s_o["loss"]) and torch.equal(s_r["clip_mask"], s_o["clip_mask"])
    B = 9 + seed
    lens = torch.randint(1, 30, (B,), generator=gen)
    cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
    n = int(cu[-1])
    lp = -torch.rand(n, generator=gen)
    rf = lp + 0.1 * torch.randn(n, generator=gen)
    sc = torch.randn(B, generator=gen) * 4
    ne = torch.rand(B, generator=gen) < 0.4
    for mz in (False, True):
        k_r, t_r = P.get_packed_rewards(0.1, 2.0, lp, rf, sc, cu, ne, mask_no_eos_with_zero=mz)
        k_o, t_o = torch_ref.packed_rewards(0.1, 2.0, lp, rf, sc, cu, ne, mz)
        assert torch.equal(k_r, k_o) and torch.equal(t_r, t_o)


@pytest.mark.parametrize("seed", range(2))
def test_sft_loss_fuzz(seed):
    fn, rec = R.sft_loss_fn()
    gen = torch.Generator().manual_seed(800 + seed)
    B, V = 4 + seed, 500 + 11 * seed
    lens = torch.randint(3, 25, (B,), generator=gen)
    cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
    T = int(cu[-1])
    x = (torch.randn(T, V, generator=gen) * 2).bfloat16()
    ids = torch.randint(0, V, (T,), generator=gen)
    mask = (torch.rand(T, generator=gen) < 0.7).int()
    mask[1] = 1
    a, b = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    l_r = fn(a, dict(input_ids=ids, cu_seqlens=cu, loss_mask=mask))
    l_o, s_o = torch_ref.packed_sft_loss(b, ids, cu, mask)
    l_r.backward()
    l_o.backward()
    assert torch.equal(l_r, l_o) and torch.equal(a.grad, b.grad)
    np.testing.assert_array_equal(rec.stats["ppl"][0].numpy(), s_o["ppl"].numpy())
    assert torch.equal(rec.stats["loss"][0], s_o["loss"])
    assert torch.equal(rec.stats["vocab_min_logits"][0], s_o["vocab_min_logits"])
