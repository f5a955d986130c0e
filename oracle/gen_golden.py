"""Generates tests/golden/*.npz by running the REAL reference from /root/reference on seeded
inputs (build container only; the fixtures are committed so the GPU box never needs the tree).

    python -m oracle.gen_golden            # regenerate everything

TEST INFRASTRUCTURE. Every fixture stores inputs AND reference outputs, so both the CPU oracle
(tests/test_oracle_golden.py) and the CUDA kernels (tests/test_*_gpu.py) are checked against the
reference itself, not against each other. bf16/fp16 arrays are stored as their uint16 bit
patterns (suffix `_bits`).
"""

from __future__ import annotations

import ast
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_loader as R  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def make_logits(gen, T, V, dtype, peaked=True):
    x = torch.randn(T, V, generator=gen) * 2.0
    labels = torch.randint(0, V, (T,), generator=gen)
    if peaked:
        rows = torch.arange(0, T, 2)
        x[rows, labels[rows]] += 6.0
    return x.to(dtype), labels


def gen_h1(F):
    cases = [
        # name, T, V, dtype, temperature, with entropy grad
        ("h1_bf16_v1024_t1", 24, 1024, torch.bfloat16, 1.0, True),
        ("h1_bf16_v1024_t07", 24, 1024, torch.bfloat16, 0.7, False),
        ("h1_bf16_v4104_ragged", 5, 4104, torch.bfloat16, 1.3, True),   # V % 8 == 0, V/8 % 16 != 0
        ("h1_bf16_v1000_unaligned", 7, 1000, torch.bfloat16, 1.0, True),  # V % 8 != 0 -> plain kernel
        ("h1_fp32_v777", 6, 777, torch.float32, 0.9, True),
        ("h1_fp16_v2048", 6, 2048, torch.float16, 1.0, False),
        ("h1_bf16_v151936", 2, 151936, torch.bfloat16, 1.0, False),  # Qwen2 vocabulary, 2 rows
    ]
    for k, (name, T, V, dtype, temp, ent_grad) in enumerate(cases):
        gen = torch.Generator().manual_seed(100 + k)
        logits, labels = make_logits(gen, T, V, dtype)
        logits.requires_grad_(True)
        lp, ent = F.gather_logprobs_entropy(logits, labels, temp)
        g_lp = torch.randn(T, generator=gen)
        g_ent = torch.randn(T, generator=gen) * (0.1 if ent_grad else 0.0)
        (lp * g_lp).sum().add((ent * g_ent).sum()).backward()
        lp_only = F.gather_logprobs(logits.detach(), labels, temp)
        assert torch.equal(lp_only, lp.detach())
        d = dict(
            labels=labels.numpy(), temperature=np.float64(temp), logp=lp.detach().numpy(),
            entropy=ent.detach().numpy(), g_logp=g_lp.numpy(), g_ent=g_ent.numpy(),
            has_ent_grad=np.bool_(ent_grad),
            vmax=logits.detach().float().max(-1).values.numpy(),
            vmin=logits.detach().float().min(-1).values.numpy(),
        )
        if dtype == torch.float32:
            d["logits_f32"] = logits.detach().numpy()
            d["dlogits_f32"] = logits.grad.numpy()
        else:
            d["logits_bits"] = bits(logits.detach())
            d["dlogits_bits"] = bits(logits.grad)
            d["dtype"] = np.str_("bf16" if dtype == torch.bfloat16 else "fp16")
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print("wrote", name)


def gen_h2_loss(F):
    gen = torch.Generator().manual_seed(7)
    T = 517
    combos = [
        dict(eps_clip=0.2, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=None),
        dict(eps_clip=0.4, eps_clip_higher=None, c_clip=None, behav_imp_weight_cap=5.0),   # gsm8k_grpo.yaml
        dict(eps_clip=0.2, eps_clip_higher=0.28, c_clip=3.0, behav_imp_weight_cap=5.0),    # cfg3
        dict(eps_clip=0.1, eps_clip_higher=0.3, c_clip=1.5, behav_imp_weight_cap=1.05),
        dict(eps_clip=0.2, eps_clip_higher=None, c_clip=2.0, behav_imp_weight_cap=None),
    ]
    for k, kw in enumerate(combos):
        lp = (-torch.rand(T, generator=gen) * 3).requires_grad_(True)
        prox = lp.detach() + 0.3 * torch.randn(T, generator=gen)
        old = prox + 0.2 * torch.randn(T, generator=gen)
        adv = torch.randn(T, generator=gen)
        adv[::17] = 0.0
        mask = torch.rand(T, generator=gen) < 0.7
        if k == 4:
            mask[:] = False  # empty mask: denominator falls back to 1
        # make some ratios sit exactly on the clip bounds / inside / outside
        loss, stat = F.ppo_actor_loss_fn(lp, prox, old, adv, loss_mask=mask, **kw)
        loss.backward()
        d = dict(logp=lp.detach().numpy(), prox=prox.numpy(), old=old.numpy(), adv=adv.numpy(),
                 mask=mask.numpy(), loss=loss.detach().numpy(), g_logp=lp.grad.numpy())
        for key, v in kw.items():
            d["kw_" + key] = np.float64(np.nan if v is None else v)
        for key, v in stat.items():
            d["stat_" + key] = v.detach().numpy()
        np.savez_compressed(os.path.join(OUT, f"h2_loss_{k}.npz"), **d)
        print("wrote h2_loss", k)


def gen_h2_legacy(P):
    """realhf actor_loss_fn (ppo_functional.py:51-149), with and without proximal_logprobs."""
    gen = torch.Generator().manual_seed(8)
    T = 300
    for k, (use_prox, cap, c_clip, use_mask) in enumerate(
        [(False, None, None, True), (True, 2.0, 3.0, True), (False, None, 1.5, False)]
    ):
        lp = (-torch.rand(T, generator=gen) * 3).requires_grad_(True)
        old = lp.detach() + 0.3 * torch.randn(T, generator=gen)
        prox = old + 0.2 * torch.randn(T, generator=gen) if use_prox else None
        adv = torch.randn(T, generator=gen)
        mask = (torch.rand(T, generator=gen) < 0.6) if use_mask else None
        loss, stat = P.actor_loss_fn(lp, old, adv, 0.2, loss_mask=mask, c_clip=c_clip,
                                     proximal_logprobs=prox, behav_imp_weight_cap=cap)
        loss.backward()
        d = dict(logp=lp.detach().numpy(), old=old.numpy(), adv=adv.numpy(),
                 loss=loss.detach().numpy(), g_logp=lp.grad.numpy(),
                 use_prox=np.bool_(use_prox), use_mask=np.bool_(use_mask),
                 cap=np.float64(np.nan if cap is None else cap),
                 c_clip=np.float64(np.nan if c_clip is None else c_clip))
        if use_prox:
            d["prox"] = prox.numpy()
        if use_mask:
            d["mask"] = mask.numpy()
        for key, v in stat.items():
            d["stat_" + key] = v.detach().numpy()
        np.savez_compressed(os.path.join(OUT, f"h2_legacy_{k}.npz"), **d)
        print("wrote h2_legacy", k)


def gen_critic_rewards(P):
    """realhf critic_loss_fn (ppo_functional.py:171-225) and get_packed_rewards (:228-253)."""
    gen = torch.Generator().manual_seed(21)
    T = 400
    for k, (kind, use_mask, eps, spread) in enumerate(
        [("mse", True, 0.2, 1.0), ("huber", True, 0.5, 12.0), ("mse", False, 0.1, 1.0),
         ("huber", False, 0.2, 12.0)]
    ):
        old = torch.randn(T, generator=gen) * spread
        value = (old + 0.4 * torch.randn(T, generator=gen)).requires_grad_(True)
        target = old + torch.randn(T, generator=gen) * spread
        mask = (torch.rand(T, generator=gen) < 0.7) if use_mask else None
        loss, stat = P.critic_loss_fn(value, old, target, eps, loss_mask=mask, loss_fn_type=kind)
        (loss * 1.7).backward()
        d = dict(value=value.detach().numpy(), old=old.numpy(), target=target.numpy(),
                 eps=np.float64(eps), huber=np.bool_(kind == "huber"), use_mask=np.bool_(use_mask),
                 loss=loss.detach().numpy(), g_value=value.grad.numpy(),
                 stat_loss=stat["loss"].numpy(), stat_clip_mask=stat["clip_mask"].numpy())
        if use_mask:
            d["mask"] = mask.numpy()
        np.savez_compressed(os.path.join(OUT, f"critic_{k}.npz"), **d)
        print("wrote critic", k)
    for k, (mz, kl_ctl, clipv) in enumerate([(False, 0.1, 5.0), (True, 0.05, 1.0)]):
        lens = torch.randint(1, 40, (24,), generator=gen)
        cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
        Tn = int(cu[-1])
        lp = -torch.rand(Tn, generator=gen) * 3
        ref = lp + 0.3 * torch.randn(Tn, generator=gen)
        score = torch.randn(24, generator=gen) * 3
        no_eos = torch.rand(24, generator=gen) < 0.3
        kl, tot = P.get_packed_rewards(kl_ctl, clipv, lp, ref, score, cu, no_eos, mask_no_eos_with_zero=mz)
        np.savez_compressed(os.path.join(OUT, f"rewards_{k}.npz"), lp=lp.numpy(), ref=ref.numpy(),
                            score=score.numpy(), cu=cu.numpy(), no_eos=no_eos.numpy(),
                            kl_ctl=np.float64(kl_ctl), clip=np.float64(clipv), mask_zero=np.bool_(mz),
                            kl=kl.numpy(), tot=tot.numpy())
        print("wrote rewards", k)


def gen_sft():
    """compute_packed_sft_loss (lm_engine.py:45-83) incl. what it registers with stats_tracker."""
    fn, rec = R.sft_loss_fn()
    gen = torch.Generator().manual_seed(31)
    for k, (B, V, dtype, cu_dtype) in enumerate(
        [(6, 1024, torch.bfloat16, torch.int32), (5, 777, torch.float32, torch.int64)]
    ):
        lens = torch.randint(4, 60, (B,), generator=gen)
        cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).to(cu_dtype)
        T = int(cu[-1])
        logits, nxt = make_logits(gen, T, V, dtype)
        logits.requires_grad_(True)
        ids = torch.roll(nxt, 1)  # so that roll(ids, -1) hits the peaked columns
        mask = torch.zeros(T, dtype=torch.int32)
        for i in range(B):  # prompt tokens masked out, like an SFT batch
            p = int(torch.randint(1, int(lens[i]) - 1, (1,), generator=gen))
            mask[int(cu[i]) + p: int(cu[i + 1])] = 1
        loss = fn(logits, dict(input_ids=ids, cu_seqlens=cu, loss_mask=mask))
        (loss * 2.5).backward()
        d = dict(input_ids=ids.numpy(), cu=cu.numpy(), loss_mask=mask.numpy(),
                 loss=loss.detach().numpy())
        if dtype == torch.float32:
            d["logits_f32"] = logits.detach().numpy()
            d["dlogits_f32"] = logits.grad.numpy()
        else:
            d["logits_bits"] = bits(logits.detach())
            d["dlogits_bits"] = bits(logits.grad)
            d["dtype"] = np.str_("bf16")
        for key, (v, den) in rec.stats.items():
            d["stat_" + key] = v.detach().numpy()
            d["den_" + key] = np.array(den)
        for key, v in rec.denominators.items():
            d["denom_" + key] = v.detach().numpy()
        np.savez_compressed(os.path.join(OUT, f"sft_{k}.npz"), **d)
        print("wrote sft", k, float(loss.detach()))


def gen_advnorm(F, AdvNorm):
    class Cfg:
        def __init__(self, m, s, g):
            self.mean_level, self.std_level, self.group_size = m, s, g

    gen = torch.Generator().manual_seed(9)
    d = {}
    shapes = {"a": (16, 12), "b": (18, 12)}  # b: batch not a multiple of the group size
    for tag, (B, L) in shapes.items():
        x = torch.randn(B, L, generator=gen) * 3 + 0.5
        mask = (torch.rand(B, L, generator=gen) < 0.6).float()
        d[f"x_{tag}"] = x.numpy()
        d[f"mask_{tag}"] = mask.numpy()
        for m in ("batch", "group", "none"):
            for s in ("batch", "group", "none"):
                for use_mask in (True, False):
                    for unbiased in (False, True):
                        out = AdvNorm(Cfg(m, s, 4))(x.clone(), mask.clone() if use_mask else None,
                                                   unbiased=unbiased)
                        d[f"out_{tag}_{m}_{s}_{int(use_mask)}_{int(unbiased)}"] = out.numpy()
    # masked_normalization on its own, incl. a `dim` argument and fp32 statistics
    x = torch.randn(6, 10, generator=gen)
    mask = (torch.rand(6, 10, generator=gen) < 0.5).float()
    d["mn_x"], d["mn_mask"] = x.numpy(), mask.numpy()
    d["mn_all"] = F.masked_normalization(x, mask).numpy()
    d["mn_nomask"] = F.masked_normalization(x).numpy()
    d["mn_dim1"] = F.masked_normalization(x, mask, dim=(1,)).numpy()
    d["mn_dim0_unbiased"] = F.masked_normalization(x, mask, dim=(0,), unbiased=True).numpy()
    d["mn_lowprec"] = F.masked_normalization(x, mask, high_precision=False).numpy()
    d["mn_zero"] = F.masked_normalization(torch.zeros(4, 5), torch.ones(4, 5)).numpy()
    np.savez_compressed(os.path.join(OUT, "h2_advnorm.npz"), **d)
    print("wrote h2_advnorm")


def gen_gae(P):
    cases = [
        # name, batch, len range, (gamma, lam), bootstrap kind
        ("gae_small_09405", 6, (1, 40), (0.99, 0.95), "ones"),
        ("gae_small_11", 6, (1, 40), (1.0, 1.0), "mixed"),
        ("gae_small_0905", 5, (1, 130), (0.9, 0.5), "zeros"),
        ("gae_mid_11", 12, (64, 2048), (1.0, 1.0), "ones"),       # cfg1 length law, fewer sequences
        ("gae_mid_09405", 12, (64, 2048), (0.99, 0.95), "mixed"),
        ("gae_long_11", 2, (3000, 5000), (1.0, 1.0), "mixed"),
    ]
    for k, (name, B, (lo, hi), (gamma, lam), bk) in enumerate(cases):
        gen = torch.Generator().manual_seed(200 + k)
        lens = torch.randint(lo, hi + 1, (B,), generator=gen)
        total = int(lens.sum())
        rewards = torch.randn(total, generator=gen)
        values = torch.randn(total + B, generator=gen)
        cu = torch.nn.functional.pad(lens.cumsum(0), (1, 0)).int()
        if bk == "ones":
            boot = torch.ones(B)
        elif bk == "zeros":
            boot = torch.zeros(B)
        else:
            boot = (torch.rand(B, generator=gen) < 0.5).float()
        adv, ret = P.pygae1d_nolp_misalign(rewards, values.clone(), cu, boot, gamma, lam)
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"), rewards=rewards.numpy(), values=values.numpy(),
            cu_seqlens=cu.numpy(), bootstrap=boot.numpy(), gamma=np.float64(gamma),
            lam=np.float64(lam), adv=adv.numpy(), ret=ret.numpy(),
        )
        print("wrote", name, total, "tokens")


def gen_padded_gae():
    """The padded GAE loop of PPOActor.compute_advantages (areal/engine/ppo/actor.py:131-148),
    executed from the reference source lines themselves."""
    import textwrap

    lines = open(os.path.join(R.REF_ROOT, "areal/engine/ppo/actor.py")).read().splitlines()[130:148]
    code = textwrap.dedent("\n".join(lines))
    d = {}
    for k, (B, L, gamma, lam) in enumerate([(6, 33, 1.0, 1.0), (9, 130, 0.99, 0.95)]):
        gen = torch.Generator().manual_seed(300 + k)
        lens = torch.randint(3, L + 1, (B,), generator=gen)
        lens[0] = L
        att = torch.arange(L)[None, :] < lens[:, None]
        rewards = torch.randn(B, L, generator=gen) * att
        values = torch.randn(B, L, generator=gen) * att
        no_eos = lens == L

        class Self:
            discount, gae_lambda = gamma, lam

        ns = dict(torch=torch, rewards=rewards, bs=B, max_seqlen=L, seq_no_eos_mask=no_eos,
                  self=Self, data={"values": values.clone()})
        exec(code, ns)
        d[f"rewards_{k}"], d[f"values_{k}"], d[f"no_eos_{k}"] = rewards.numpy(), values.numpy(), no_eos.numpy()
        d[f"gamma_{k}"], d[f"lam_{k}"] = np.float64(gamma), np.float64(lam)
        d[f"adv_{k}"] = ns["advantages"].numpy()
    np.savez_compressed(os.path.join(OUT, "gae_padded.npz"), **d)
    print("wrote gae_padded")


def _ref_make_intervals():
    path = os.path.join(R.REF_ROOT, "realhf/tests/cpp_extensions/test_interval_ops.py")
    tree = ast.parse(open(path).read())
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "make_intervals"]
    g = dict(np=np)
    exec(compile(ast.Module(body=fn, type_ignores=[]), path, "exec"), g)
    return g["make_intervals"]


def gen_intervals():
    fp = R.flatten_param_py()
    make_intervals = _ref_make_intervals()  # realhf/tests/cpp_extensions/test_interval_ops.py:20-31
    merge_ref = R.merge_intervals_ext()
    d = {}
    for tag, (size, n, dtype) in {
        "bf16_a": (40000, 1000, torch.bfloat16),
        "fp32_b": (9000, 100, torch.float32),
        "fp16_c": (5000, 1, torch.float16),
    }.items():
        np.random.seed(len(tag) + size)
        iv, out_size, _ = make_intervals(size, n)
        gen = torch.Generator().manual_seed(size)
        src_bits = torch.randint(-32768, 32767, (size * (2 if dtype == torch.float32 else 1),),
                                 generator=gen, dtype=torch.int16)
        src = src_bits.view(dtype)
        got = fp["_slice_intervals_py"](src, [tuple(x) for x in iv.tolist()])
        dst = torch.zeros(size, dtype=dtype)
        fp["_set_intervals_py"](got, dst, [tuple(x) for x in iv.tolist()])
        d[f"{tag}_intervals"] = iv
        d[f"{tag}_src_bits"] = src_bits.numpy().view(np.uint16)
        d[f"{tag}_slice_bits"] = got.contiguous().view(torch.int16).numpy().view(np.uint16)
        d[f"{tag}_set_bits"] = dst.view(torch.int16).numpy().view(np.uint16)
    # adversarial: odd offsets, tiny and empty intervals, one long interval, arbitrary order
    rng = np.random.default_rng(5)
    size = 70001
    starts = np.sort(rng.choice(size - 9000, 300, replace=False))
    iv = []
    for i, s in enumerate(starts):
        lim = (starts[i + 1] if i + 1 < len(starts) else size) - s
        ln = int(min(lim, rng.integers(0, 8)))
        iv.append((int(s), int(s) + ln))
    iv.append((size - 8191, size))  # long one at an odd offset
    rng.shuffle(iv)
    iv = np.array(iv, dtype=np.int64)
    src_bits = torch.from_numpy(rng.integers(-32768, 32767, size, dtype=np.int16))
    src = src_bits.view(torch.bfloat16)
    got = fp["_slice_intervals_py"](src, [tuple(x) for x in iv.tolist()])
    dst = torch.zeros(size, dtype=torch.bfloat16)
    fp["_set_intervals_py"](got, dst, [tuple(x) for x in iv.tolist()])
    d["adv_intervals"] = iv
    d["adv_src_bits"] = src_bits.numpy().view(np.uint16)
    d["adv_slice_bits"] = got.contiguous().view(torch.int16).numpy().view(np.uint16)
    d["adv_set_bits"] = dst.view(torch.int16).numpy().view(np.uint16)
    # merge_intervals through the reference's own compiled C++
    m_in = [(0, 4), (4, 9), (9, 10), (12, 15), (15, 15), (15, 20), (3, 5), (5, 6), (100, 101)]
    d["merge_in"] = np.array(m_in, dtype=np.int64)
    d["merge_out"] = np.array(merge_ref(m_in), dtype=np.int64)
    rngi = rng.integers(0, 50, size=(400, 2))
    m_in2 = [(int(min(a, b)), int(max(a, b))) for a, b in rngi]
    d["merge_in2"] = np.array(m_in2, dtype=np.int64)
    d["merge_out2"] = np.array(merge_ref(m_in2), dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "h4_intervals.npz"), **d)
    print("wrote h4_intervals")


def main():
    assert R.available(), "the reference tree is required to (re)generate goldens"
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    F = R.areal_functional()
    P = R.realhf_ppo_functional()
    gen_h1(F)
    gen_h2_loss(F)
    gen_h2_legacy(P)
    gen_critic_rewards(P)
    gen_sft()
    gen_advnorm(F, R.adv_norm_class())
    gen_gae(P)
    gen_padded_gae()
    gen_intervals()


if __name__ == "__main__":
    main()
