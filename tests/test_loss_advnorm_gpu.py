"""H2 parity on the GPU: fused PPO loss (+ gradient), masked_normalization, AdvNorm (all nine
level combinations), group reward norm, and the fused normalise-inside-the-loss entry point.

Tolerances: loss and per-token stats within 1e-3 relative of the reference (north_star) — they
agree far better in practice, so the tests ask for rtol=1e-5/atol=1e-6; advantages within 1e-5.
Boolean stats must match exactly."""

import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _kw(g, key):
    v = float(g["kw_" + key])
    return None if np.isnan(v) else v


def _cmp_stat(stat, g, cuda):
    for key in ("loss", "importance_weight", "approx_kl", "clip_mask", "dual_clip_mask",
                "behave_imp_weight", "behave_approx_kl", "behave_mask"):
        ref = g.get("stat_" + key)
        if ref is None:
            assert key not in stat
            continue
        v = stat[key]
        assert v.shape == ref.shape, key
        if ref.dtype == np.bool_:
            assert v.dtype == torch.bool, key  # stats_tracker.denominator needs bool (stats_tracker.py:85)
            assert np.array_equal(v.cpu().numpy(), ref), key
        else:
            assert v.dtype == torch.float32, key  # stats_tracker.stat needs fp32 (:111)
            np.testing.assert_allclose(v.cpu().numpy(), ref, rtol=1e-5, atol=1e-6, err_msg=key)


@pytest.mark.parametrize("k", range(5))
def test_loss_golden(k, cuda):
    from areal_vllm_b200 import functional as BF

    g = load_golden(f"h2_loss_{k}")
    lp = torch.from_numpy(g["logp"]).to(cuda).requires_grad_(True)
    loss, stat = BF.ppo_actor_loss_fn(
        lp, torch.from_numpy(g["prox"]).to(cuda), torch.from_numpy(g["old"]).to(cuda),
        torch.from_numpy(g["adv"]).to(cuda), _kw(g, "eps_clip"), torch.from_numpy(g["mask"]).to(cuda),
        eps_clip_higher=_kw(g, "eps_clip_higher"), c_clip=_kw(g, "c_clip"),
        behav_imp_weight_cap=_kw(g, "behav_imp_weight_cap"))
    assert loss.dim() == 0 and loss.dtype == torch.float32
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=1e-5, atol=1e-7)
    _cmp_stat(stat, g, cuda)
    (loss * 3.0).backward()  # engine scales the loss before backward (fsdp_engine.py:543-550)
    np.testing.assert_allclose(lp.grad.cpu().numpy(), 3.0 * g["g_logp"], rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("k", range(3))
def test_legacy_actor_loss_golden(k, cuda):
    from areal_vllm_b200 import ppo_functional as PF

    g = load_golden(f"h2_legacy_{k}")
    t = lambda a: torch.from_numpy(a).to(cuda)
    lp = t(g["logp"]).requires_grad_(True)
    cap, c_clip = float(g["cap"]), float(g["c_clip"])
    loss, stat = PF.actor_loss_fn(
        lp, t(g["old"]), t(g["adv"]), 0.2, loss_mask=t(g["mask"]) if bool(g["use_mask"]) else None,
        c_clip=None if np.isnan(c_clip) else c_clip,
        proximal_logprobs=t(g["prox"]) if bool(g["use_prox"]) else None,
        behav_imp_weight_cap=None if np.isnan(cap) else cap)
    np.testing.assert_allclose(loss.item(), g["loss"], rtol=1e-5, atol=1e-7)
    _cmp_stat(stat, g, cuda)
    loss.backward()
    np.testing.assert_allclose(lp.grad.cpu().numpy(), g["g_logp"], rtol=1e-5, atol=1e-9)


def test_loss_vs_oracle_ties_and_sizes(cuda):
    """Ratios exactly on / inside / outside the clip bounds, zero advantages, large T (multi-CTA
    reduction), dual clip ties; value and gradient vs the torch oracle."""
    from areal_vllm_b200 import functional as BF
    from oracle import torch_ref

    for T in (1, 31, 4096, 300001):
        gen = torch.Generator().manual_seed(T)
        prox = -torch.rand(T, generator=gen) * 2
        lp = prox.clone()
        lp[1::4] += 0.5 * torch.randn(T, generator=gen)[1::4]
        lp[2::4] += np.log(1.2) + 0.01   # just outside the upper bound
        lp[3::4] += np.log(0.8) - 0.01   # just outside the lower bound
        old = prox + 0.1 * torch.randn(T, generator=gen)
        adv = torch.randn(T, generator=gen)
        adv[::7] = 0.0
        mask = torch.rand(T, generator=gen) < 0.8
        for kw in (dict(), dict(c_clip=3.0, behav_imp_weight_cap=1.1), dict(eps_clip_higher=0.28)):
            a = lp.clone().requires_grad_(True)
            l_r, s_r = torch_ref.ppo_actor_loss(a, prox, old, adv, 0.2, mask, **kw)
            l_r.backward()
            b = lp.clone().to(cuda).requires_grad_(True)
            l_g, s_g = BF.ppo_actor_loss_fn(b, prox.to(cuda), old.to(cuda), adv.to(cuda), 0.2,
                                            mask.to(cuda), **kw)
            l_g.backward()
            np.testing.assert_allclose(l_g.item(), l_r.item(), rtol=2e-5, atol=1e-7)
            np.testing.assert_allclose(b.grad.cpu().numpy(), a.grad.numpy(), rtol=1e-5, atol=1e-10)
            for key in s_r:
                if s_r[key].dtype == torch.bool:
                    assert torch.equal(s_g[key].cpu(), s_r[key]), key
                else:
                    np.testing.assert_allclose(s_g[key].cpu().numpy(), s_r[key].numpy(), rtol=1e-5, atol=1e-6)
    # exact ties: lp == prox gives ratio == 1.0 on both sides; with eps_clip = 0 the ratio sits ON both
    # clamp bounds (inclusive -> gradient passes) and pg_loss1 == pg_loss2 (max splits 1/2 + 1/2)
    T = 64
    gen = torch.Generator().manual_seed(1)
    prox = -torch.rand(T, generator=gen)
    adv = torch.randn(T, generator=gen)
    mask = torch.ones(T, dtype=torch.bool)
    a = prox.clone().requires_grad_(True)
    l_r, _ = torch_ref.ppo_actor_loss(a, prox, prox, adv, 0.0, mask, c_clip=1.5)
    l_r.backward()
    b = prox.clone().to(cuda).requires_grad_(True)
    l_g, _ = BF.ppo_actor_loss_fn(b, prox.to(cuda), prox.to(cuda), adv.to(cuda), 0.0, mask.to(cuda), c_clip=1.5)
    l_g.backward()
    np.testing.assert_allclose(b.grad.cpu().numpy(), a.grad.numpy(), rtol=1e-6, atol=0)
    assert float(a.grad.abs().sum()) > 0
    # 2-D inputs keep their shape
    x = torch.randn(4, 9, device=cuda)
    l, s = BF.ppo_actor_loss_fn(x, x, x, x, 0.2, torch.ones(4, 9, dtype=torch.bool, device=cuda))
    assert s["loss"].shape == (4, 9) and s["clip_mask"].shape == (4, 9)


def test_advnorm_golden_all_levels(cuda):
    from areal_vllm_b200 import functional as BF
    from areal_vllm_b200.actor import AdvNorm

    class Cfg:
        def __init__(self, m, s, g):
            self.mean_level, self.std_level, self.group_size = m, s, g

    g = load_golden("h2_advnorm")
    for tag in ("a", "b"):
        x, mask = torch.from_numpy(g[f"x_{tag}"]).to(cuda), torch.from_numpy(g[f"mask_{tag}"]).to(cuda)
        for m in ("batch", "group", "none"):
            for s in ("batch", "group", "none"):
                for use_mask in (1, 0):
                    for unbiased in (0, 1):
                        ref = g[f"out_{tag}_{m}_{s}_{use_mask}_{unbiased}"]
                        out = AdvNorm(Cfg(m, s, 4))(x, mask if use_mask else None, unbiased=bool(unbiased))
                        assert out.dtype == torch.float32 and tuple(out.shape) == ref.shape, (tag, m, s)
                        # trailing partial group + group-level std divides by eps=1e-5: values ~1e5
                        np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=1e-5, atol=1e-5,
                                                   err_msg=f"{tag} {m} {s} {use_mask} {unbiased}")
    x, mask = torch.from_numpy(g["mn_x"]).to(cuda), torch.from_numpy(g["mn_mask"]).to(cuda)
    cmp = lambda out, key: np.testing.assert_allclose(out.cpu().numpy(), g[key], rtol=1e-5, atol=1e-5, err_msg=key)
    cmp(BF.masked_normalization(x, mask), "mn_all")
    cmp(BF.masked_normalization(x), "mn_nomask")
    cmp(BF.masked_normalization(x, mask, dim=(1,)), "mn_dim1")
    cmp(BF.masked_normalization(x, mask, dim=(0,), unbiased=True), "mn_dim0_unbiased")
    cmp(BF.masked_normalization(x, mask, high_precision=False), "mn_lowprec")
    z = BF.masked_normalization(torch.zeros(4, 5, device=cuda), torch.ones(4, 5, device=cuda))
    assert torch.equal(z.cpu(), torch.zeros(4, 5))  # areal/tests/test_adv_norm_config.py:243-254


def test_advnorm_cfg2_shape_and_fused_loss(cuda):
    """[512, 1024] advantages in groups of 8 (BASELINE config 2), group/group AdvNorm vs the oracle,
    then the fused kernel (normalise inside the loss) vs AdvNorm followed by the plain loss."""
    from areal_vllm_b200 import functional as BF
    from areal_vllm_b200.actor import (AdvNorm, group_advantage_stats, group_reward_norm,
                                       ppo_actor_loss_fn_fused_advnorm)
    from oracle import torch_ref

    class Cfg:
        mean_level, std_level, group_size = "group", "group", 8

    B, L, G = 512, 1024, 8
    gen = torch.Generator().manual_seed(2)
    lens = torch.randint(16, L, (B,), generator=gen)
    mask = (torch.arange(L)[None, :] < lens[:, None]).float()
    adv = (torch.randn(B, 1, generator=gen) * 5).expand(B, L).contiguous() * mask + 0.01 * torch.randn(B, L, generator=gen)
    ref = torch_ref.adv_norm(adv, mask, "group", "group", G)
    out = AdvNorm(Cfg)(adv.to(cuda), mask.to(cuda))
    np.testing.assert_allclose(out.cpu().numpy(), ref.numpy(), rtol=1e-5, atol=1e-5)

    # group reward normalisation (actor.py:94-98)
    rew = torch.randint(0, 2, (B,), generator=gen).float() * 10 - 5
    np.testing.assert_allclose(group_reward_norm(rew.to(cuda), G).cpu().numpy(),
                               torch_ref.group_reward_norm(rew, G).numpy(), rtol=1e-5, atol=1e-5)

    # fused: raw advantages + per-group (mean, denom) + token->group map, packed layout
    mean, denom = group_advantage_stats(adv.to(cuda), mask.to(cuda), G)
    keep = mask.bool()
    tok_group = (torch.arange(B)[:, None] // G).expand(B, L)[keep].int()
    raw = adv[keep]
    T = raw.numel()
    lp = -torch.rand(T, generator=gen)
    prox = lp + 0.05 * torch.randn(T, generator=gen)
    old = prox + 0.05 * torch.randn(T, generator=gen)
    lmask = torch.rand(T, generator=gen) < 0.9
    a = lp.clone().to(cuda).requires_grad_(True)
    loss_f, stat_f, adv_used = ppo_actor_loss_fn_fused_advnorm(
        a, prox.to(cuda), old.to(cuda), raw.to(cuda), 0.4, lmask.to(cuda), mean, denom,
        tok_group.to(cuda), behav_imp_weight_cap=5.0)
    loss_f.backward()
    # the fused kernel masks the advantage with the LOSS mask of the packed call
    norm_packed = ((raw.double() * lmask.double() - mean.cpu()[tok_group.long()]) / denom.cpu()[tok_group.long()]).float()
    np.testing.assert_allclose(adv_used.cpu().numpy(), norm_packed.numpy(), rtol=1e-6, atol=1e-6)
    b = lp.clone().requires_grad_(True)
    loss_r, stat_r = torch_ref.ppo_actor_loss(b, prox, old, norm_packed, 0.4, lmask, behav_imp_weight_cap=5.0)
    loss_r.backward()
    np.testing.assert_allclose(loss_f.item(), loss_r.item(), rtol=2e-5, atol=1e-7)
    np.testing.assert_allclose(a.grad.cpu().numpy(), b.grad.numpy(), rtol=1e-5, atol=1e-10)
    assert torch.equal(stat_f["clip_mask"].cpu(), stat_r["clip_mask"])


def test_grpo_loss_fn_end_to_end(cuda):
    """The composite the engine calls (actor.py:292-367): logits -> loss -> backward, vs the oracle."""
    from areal_vllm_b200.actor import grpo_loss_fn
    from oracle import torch_ref

    T, V = 300, 4096
    gen = torch.Generator().manual_seed(4)
    logits = (torch.randn(T, V, generator=gen) * 2).bfloat16()
    ids = torch.randint(0, V, (T,), generator=gen)
    data = dict(input_ids=ids, logprobs=-torch.rand(T, generator=gen) * 8, advantages=torch.randn(T, generator=gen),
                loss_mask=(torch.rand(T, generator=gen) < 0.7).float(), prox_logp=-torch.rand(T, generator=gen) * 8)
    xg = logits.to(cuda).requires_grad_(True)
    stats = {}
    loss = grpo_loss_fn(xg, {k: v.to(cuda) for k, v in data.items()}, 1.0, 0.2, 0.28, 3.0, 5.0, stats=stats)
    loss.backward()
    xc = logits.clone().requires_grad_(True)
    lp, ent = torch_ref.logprobs_entropy(xc, torch.roll(ids, -1), 1.0)
    loss_r, _ = torch_ref.ppo_actor_loss(lp, data["prox_logp"], data["logprobs"], data["advantages"], 0.2,
                                         data["loss_mask"].bool(), 0.28, 3.0, 5.0)
    loss_r.backward()
    np.testing.assert_allclose(loss.item(), loss_r.item(), rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(xg.grad.float().cpu().numpy(), xc.grad.float().numpy(), rtol=2.0**-7, atol=1e-6)
    assert np.array_equal(stats["vocab_max_logits"].cpu().numpy(), logits.float().max(-1).values.numpy())
    assert np.array_equal(stats["vocab_min_logits"].cpu().numpy(), logits.float().min(-1).values.numpy())
    np.testing.assert_allclose(stats["entropy"].cpu().numpy(), ent.detach().numpy(), rtol=1e-3, atol=2e-5)


def test_padded_rows_get_exact_zero_gradient(cuda):
    """INTEGRATION.md 'Avoiding the padded-gradient copy': running the loss on the padded logits
    with loss_mask = 0 on the pad rows gives the same loss, the same gradient on the real rows and
    exact zeros on the pad rows."""
    from areal_vllm_b200 import actor

    gen = torch.Generator().manual_seed(21)
    T, Tp, V = 300, 512, 4096
    logits = (torch.randn(Tp, V, generator=gen) * 2).bfloat16().to(cuda)
    mask = (torch.rand(T, generator=gen) < 0.7)
    mask[-1] = False  # its label comes from the roll over the stream end (garbage by design)
    data = dict(input_ids=torch.randint(0, V, (T,), generator=gen).to(cuda),
                logprobs=(-torch.rand(T, generator=gen) * 8).to(cuda),
                advantages=torch.randn(T, generator=gen).to(cuda), loss_mask=mask.to(cuda),
                prox_logp=(-torch.rand(T, generator=gen) * 8).to(cuda))
    a = logits[:T].clone().requires_grad_(True)
    la = actor.grpo_loss_fn(a, data, 1.0, 0.2, None, None, 5.0)
    la.backward()
    pad = lambda t, fill=0: torch.cat([t, torch.full((Tp - T,), fill, dtype=t.dtype, device=t.device)])
    datap = {k: pad(v) for k, v in data.items()}
    b = logits.clone().requires_grad_(True)
    lb = actor.grpo_loss_fn(b, datap, 1.0, 0.2, None, None, 5.0)
    lb.backward()
    assert torch.equal(la, lb)
    assert torch.equal(b.grad[:T], a.grad)
    assert int((b.grad[T:].view(torch.int16) & 0x7fff).max()) == 0  # +-0 only


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_group_advantage_norm_matches_advnorm(world, cuda):
    """Rows of one GRPO group on different ranks (sequence-granular sharding): the [G, 3] moment
    tables of the simulated ranks are summed (what the all-reduce does) and every rank's rows must
    equal AdvNorm(group, group) on the whole batch (ref actor.py:412-470)."""
    from areal_vllm_b200 import dist as bdist
    from areal_vllm_b200.actor import AdvNorm

    gen = torch.Generator().manual_seed(5)
    B, L, gs = 64, 96, 8
    lens = torch.randint(4, L + 1, (B,), generator=gen)
    mask = (torch.arange(L)[None, :] < lens[:, None]).float()
    adv = torch.randn(B, L, generator=gen) * 3 + 1
    adv, mask = adv.to(cuda), mask.to(cuda)

    class Cfg:
        mean_level, std_level, group_size = "group", "group", gs

    for unbiased in (False, True):
        full = AdvNorm(Cfg)(adv, mask, unbiased=unbiased)
        parts = bdist.shard_sequences_by_tokens(lens.tolist(), world)
        idx = [torch.tensor(p, device=cuda, dtype=torch.int64) for p in parts]
        tables = []
        for r in range(world):  # pass 1: every rank's local table
            bdist.sharded_group_advantage_norm(adv[idx[r]], mask[idx[r]], idx[r] // gs, B // gs, unbiased=unbiased,
                                               reduce=lambda t, *e: (tables.append(t.clone()) or (t,) + e))
        total = torch.stack(tables).sum(0)
        for r in range(world):  # pass 2: with the summed table, as after the all-reduce
            out, extra = bdist.sharded_group_advantage_norm(
                adv[idx[r]], mask[idx[r]], idx[r] // gs, B // gs, unbiased=unbiased,
                extra=(torch.ones(1, device=cuda),), reduce=lambda t, *e: (total,) + e)
            assert extra.item() == 1.0
            if len(parts[r]):
                torch.testing.assert_close(out, full[idx[r]], rtol=1e-5, atol=1e-5)
            else:
                assert out.shape == (0, L)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("T,V,mask_kind,want", [(300, 4096, "float", True), (64, 4100, "bool", True),
                                                (1, 512, "float", True), (257, 151936, "bool", False),
                                                (130, 1003, "int", True)])
def test_grpo_loss_fast_path_equals_op_path(T, V, mask_kind, want, dtype, cuda):
    """The two-launch composite (label roll as an index, fp32/bool mask read as is, logging tensors
    written by the loss kernel, loss backward folded into the log-prob backward) against the
    custom-op path it replaces in eager mode: same loss bits, same dlogits bits, same stats."""
    from areal_vllm_b200 import actor

    gen = torch.Generator().manual_seed(T + V)
    logits = (torch.randn(T, V, generator=gen) * 2).to(dtype).to(cuda)
    m = torch.rand(T, generator=gen) < 0.7
    mask = {"float": m.float(), "bool": m, "int": m.long() * 3}[mask_kind]
    data = {k: v.to(cuda) for k, v in dict(
        input_ids=torch.randint(0, V, (T,), generator=gen), logprobs=-torch.rand(T, generator=gen) * 4,
        advantages=torch.randn(T, generator=gen), loss_mask=mask, prox_logp=-torch.rand(T, generator=gen) * 4).items()}
    scale = torch.tensor(0.37, device=cuda)
    res = {}
    for fast in (True, False):
        actor.FAST_PATH = fast
        try:
            x = logits.clone().requires_grad_(True)
            st = {} if want else None
            loss = actor.grpo_loss_fn(x, data, 0.8, 0.2, 0.28, 3.0, 5.0, stats=st)
            (loss * scale).backward()
            with torch.no_grad():
                loss_ng = actor.grpo_loss_fn(x.detach(), data, 0.8, 0.2, 0.28, 3.0, 5.0)
            res[fast] = (loss.detach().clone(), x.grad.clone(), st, loss_ng.clone())
        finally:
            actor.FAST_PATH = True
    (la, ga, sa, na), (lb, gb, sb, nb) = res[True], res[False]
    assert torch.equal(la, lb) and torch.equal(na, nb) and torch.equal(la, na)
    assert torch.equal(ga.view(torch.int16 if dtype != torch.float32 else torch.int32),
                       gb.view(torch.int16 if dtype != torch.float32 else torch.int32))
    if want:
        assert set(sa) == set(sb)
        for k in sa:
            assert sa[k].dtype == sb[k].dtype and sa[k].shape == sb[k].shape, k
            assert torch.equal(sa[k], sb[k]), k


def test_grpo_loss_fast_path_empty_and_errors(cuda):
    """T = 0 (an empty micro-batch after dynamic filtering) and the argument errors of the composite."""
    from areal_vllm_b200 import actor

    V = 1024
    empty = dict(input_ids=torch.zeros(0, dtype=torch.int64, device=cuda), logprobs=torch.zeros(0, device=cuda),
                 advantages=torch.zeros(0, device=cuda), loss_mask=torch.zeros(0, device=cuda),
                 prox_logp=torch.zeros(0, device=cuda))
    x = torch.zeros(0, V, dtype=torch.bfloat16, device=cuda, requires_grad=True)
    st = {}
    loss = actor.grpo_loss_fn(x, empty, 1.0, 0.2, None, None, 5.0, stats=st)
    assert loss.item() == 0.0 and st["new_logp"].shape == (0,) and st["n_valid_tokens"].dtype == torch.bool
    loss.backward()
    assert x.grad.shape == (0, V)
    # wrong lengths / devices raise RuntimeError like the reference's TORCH_CHECKs
    T = 16
    good = dict(input_ids=torch.zeros(T, dtype=torch.int64, device=cuda), logprobs=torch.zeros(T, device=cuda),
                advantages=torch.zeros(T, device=cuda), loss_mask=torch.ones(T, device=cuda),
                prox_logp=torch.zeros(T, device=cuda))
    xl = torch.randn(T, V, device=cuda).bfloat16()
    for key, bad in (("advantages", torch.zeros(T + 1, device=cuda)), ("loss_mask", torch.ones(T - 1, device=cuda)),
                     ("prox_logp", torch.zeros(T))):
        with pytest.raises(RuntimeError):
            actor.grpo_loss_fn(xl, {**good, key: bad}, 1.0, 0.2, None, None, 5.0)
    with pytest.raises(AssertionError):
        actor.grpo_loss_fn(xl, good, 1.0, 0.2, None, 0.5, 5.0)  # c_clip must be > 1 (functional.py:157)
