"""The CPU oracle against the fixtures that oracle/gen_golden.py produced from the REAL reference
(areal/utils/functional.py, areal/engine/ppo/actor.py, realhf/.../ppo_functional.py,
realhf/.../flatten_param.py, csrc/interval_op/interval_op.cpp). This pins the oracle."""

import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, from_bits, golden_logits, load_golden
from oracle import cpu_ref, torch_ref

H1 = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "h1_*.npz")))
GAE = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "gae_*.npz"))
             if "padded" not in p)


@pytest.mark.parametrize("name", H1)
def test_h1_forward_and_grad(name):
    g = load_golden(name)
    logits, dref = golden_logits(g)
    labels = torch.from_numpy(g["labels"])
    temp = float(g["temperature"])
    x = logits.clone().requires_grad_(True)
    lp, ent = torch_ref.logprobs_entropy(x, labels, temp)
    # same torch ops as the reference: identical up to libm/threading differences between hosts
    np.testing.assert_allclose(lp.detach().numpy(), g["logp"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(ent.detach().numpy(), g["entropy"], rtol=1e-6, atol=1e-6)
    (lp * torch.from_numpy(g["g_logp"])).sum().add((ent * torch.from_numpy(g["g_ent"])).sum()).backward()
    tol = 1e-6 if logits.dtype == torch.float32 else 8e-3  # one ulp of the 16-bit grad dtype
    np.testing.assert_allclose(x.grad.float().numpy(), dref.float().numpy(), rtol=tol, atol=1e-7)
    # fp64 ground truth agrees with the fp32 reference at the parity tolerance
    lp64, ent64 = torch_ref.logprobs_entropy_f64(logits, labels, temp)
    np.testing.assert_allclose(lp64.numpy(), g["logp"], rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(ent64.numpy(), g["entropy"], rtol=1e-3, atol=1e-5)


def _kw(g, key):
    v = float(g["kw_" + key])
    return None if np.isnan(v) else v


@pytest.mark.parametrize("k", range(5))
def test_h2_loss(k):
    g = load_golden(f"h2_loss_{k}")
    lp = torch.from_numpy(g["logp"]).requires_grad_(True)
    loss, stat = torch_ref.ppo_actor_loss(
        lp, torch.from_numpy(g["prox"]), torch.from_numpy(g["old"]), torch.from_numpy(g["adv"]),
        _kw(g, "eps_clip"), torch.from_numpy(g["mask"]), _kw(g, "eps_clip_higher"),
        _kw(g, "c_clip"), _kw(g, "behav_imp_weight_cap"),
    )
    loss.backward()
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(lp.grad.numpy(), g["g_logp"], rtol=1e-6, atol=1e-9)
    for key, v in stat.items():
        ref = g["stat_" + key]
        if ref.dtype == np.bool_:
            assert v.dtype == torch.bool and np.array_equal(v.numpy(), ref), key
        else:
            assert v.dtype == torch.float32
            np.testing.assert_allclose(v.numpy(), ref, rtol=1e-6, atol=1e-7, err_msg=key)


def test_h2_advnorm_all_levels():
    g = load_golden("h2_advnorm")
    for tag in ("a", "b"):
        x, mask = torch.from_numpy(g[f"x_{tag}"]), torch.from_numpy(g[f"mask_{tag}"])
        for m in ("batch", "group", "none"):
            for s in ("batch", "group", "none"):
                for use_mask in (1, 0):
                    for unbiased in (0, 1):
                        ref = g[f"out_{tag}_{m}_{s}_{use_mask}_{unbiased}"]
                        out = torch_ref.adv_norm(x, mask if use_mask else None, m, s, 4,
                                                 unbiased=bool(unbiased))
                        assert out.shape == ref.shape and out.dtype == torch.float32
                        np.testing.assert_allclose(out.numpy(), ref, rtol=1e-6, atol=1e-6,
                                                   err_msg=f"{tag} {m} {s} {use_mask} {unbiased}")
    x, mask = torch.from_numpy(g["mn_x"]), torch.from_numpy(g["mn_mask"])
    np.testing.assert_allclose(torch_ref.masked_normalization(x, mask).numpy(), g["mn_all"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(torch_ref.masked_normalization(x).numpy(), g["mn_nomask"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(torch_ref.masked_normalization(x, mask, dim=(1,)).numpy(), g["mn_dim1"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(torch_ref.masked_normalization(x, mask, dim=(0,), unbiased=True).numpy(),
                               g["mn_dim0_unbiased"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(torch_ref.masked_normalization(x, mask, high_precision=False).numpy(),
                               g["mn_lowprec"], rtol=1e-5, atol=1e-5)
    # zero input -> zero output (reference test areal/tests/test_adv_norm_config.py:243-254)
    assert np.array_equal(g["mn_zero"], np.zeros((4, 5), np.float32))
    assert torch.equal(torch_ref.masked_normalization(torch.zeros(4, 5), torch.ones(4, 5)), torch.zeros(4, 5))


@pytest.mark.parametrize("name", GAE)
def test_gae_bit_exact_vs_pygae(name):
    g = load_golden(name)
    adv, ret = cpu_ref.gae_py(g["rewards"], g["values"], g["cu_seqlens"], g["bootstrap"],
                              float(g["gamma"]), float(g["lam"]))
    assert np.array_equal(adv, g["adv"]), np.abs(adv - g["adv"]).max()
    assert np.array_equal(ret, g["ret"])
    # the CUDA reference's contraction (FMA) stays within the reference test's tolerance
    # (realhf/tests/cpp_extensions/test_cugae.py:50-51, atol=1e-5 + rtol=1e-5)
    adv_c, ret_c = cpu_ref.gae_cuda_order(g["rewards"], g["values"], g["cu_seqlens"],
                                          g["bootstrap"], float(g["gamma"]), float(g["lam"]))
    if float(g["gamma"]) == 1.0 and float(g["lam"]) == 1.0:
        assert np.array_equal(adv_c, g["adv"]) and np.array_equal(ret_c, g["ret"])
    else:
        np.testing.assert_allclose(adv_c, g["adv"], rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(ret_c, g["ret"], rtol=1e-5, atol=1e-5)


def test_gae_python_loop_agrees_with_c():
    g = load_golden("gae_small_0905")
    a1, r1 = cpu_ref.gae_py_loops(g["rewards"], g["values"], g["cu_seqlens"], g["bootstrap"],
                                  float(g["gamma"]), float(g["lam"]))
    assert np.array_equal(a1, g["adv"]) and np.array_equal(r1, g["ret"])


@pytest.mark.parametrize("tag,dtype", [("bf16_a", np.uint16), ("fp32_b", np.uint32),
                                       ("fp16_c", np.uint16), ("adv", np.uint16)])
def test_intervals_bit_exact(tag, dtype):
    g = load_golden("h4_intervals")
    iv = g[f"{tag}_intervals"]
    src = g[f"{tag}_src_bits"].view(dtype)
    got = cpu_ref.slice_intervals(src, iv)
    assert np.array_equal(got, g[f"{tag}_slice_bits"].view(dtype))
    dst = np.zeros_like(src)
    cpu_ref.set_intervals(got, dst, iv)
    assert np.array_equal(dst, g[f"{tag}_set_bits"].view(dtype))


def test_merge_intervals_vs_reference_cpp():
    g = load_golden("h4_intervals")
    for a, b in (("merge_in", "merge_out"), ("merge_in2", "merge_out2")):
        got = cpu_ref.merge_intervals([tuple(x) for x in g[a].tolist()])
        assert got == [tuple(x) for x in g[b].tolist()]
    assert cpu_ref.merge_intervals([]) == []


def test_padded_gae_vs_reference_loop():
    g = load_golden("gae_padded")
    for k in range(2):
        out = torch_ref.padded_gae(torch.from_numpy(g[f"rewards_{k}"]), torch.from_numpy(g[f"values_{k}"]),
                                   torch.from_numpy(g[f"no_eos_{k}"]), float(g[f"gamma_{k}"]), float(g[f"lam_{k}"]))
        assert np.array_equal(out.numpy(), g[f"adv_{k}"])


@pytest.mark.parametrize("k", range(4))
def test_critic_loss_golden(k):
    g = load_golden(f"critic_{k}")
    value = torch.from_numpy(g["value"]).requires_grad_(True)
    mask = torch.from_numpy(g["mask"]) if bool(g["use_mask"]) else None
    loss, stat = torch_ref.critic_loss(value, torch.from_numpy(g["old"]), torch.from_numpy(g["target"]),
                                       float(g["eps"]), loss_mask=mask,
                                       loss_fn_type="huber" if bool(g["huber"]) else "mse")
    (loss * 1.7).backward()
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-6)
    np.testing.assert_allclose(value.grad.numpy(), g["g_value"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(stat["loss"].numpy(), g["stat_loss"], rtol=1e-6)
    assert np.array_equal(stat["clip_mask"].numpy(), g["stat_clip_mask"])


@pytest.mark.parametrize("k", range(2))
def test_packed_rewards_golden(k):
    g = load_golden(f"rewards_{k}")
    kl, tot = torch_ref.packed_rewards(float(g["kl_ctl"]), float(g["clip"]), torch.from_numpy(g["lp"]),
                                       torch.from_numpy(g["ref"]), torch.from_numpy(g["score"]),
                                       torch.from_numpy(g["cu"]), torch.from_numpy(g["no_eos"]),
                                       bool(g["mask_zero"]))
    assert np.array_equal(kl.numpy(), g["kl"])
    assert np.array_equal(tot.numpy(), g["tot"])


@pytest.mark.parametrize("k", range(2))
def test_sft_loss_golden(k):
    g = load_golden(f"sft_{k}")
    logits, dref = golden_logits(g)
    x = logits.clone().requires_grad_(True)
    loss, stat = torch_ref.packed_sft_loss(x, torch.from_numpy(g["input_ids"]), torch.from_numpy(g["cu"]),
                                           torch.from_numpy(g["loss_mask"]))
    (loss * 2.5).backward()
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-6)
    tol = 1e-6 if logits.dtype == torch.float32 else 8e-3
    np.testing.assert_allclose(x.grad.float().numpy(), dref.float().numpy(), rtol=tol, atol=1e-8)
    np.testing.assert_allclose(stat["ppl"].numpy(), g["stat_ppl"], rtol=1e-5)
    np.testing.assert_allclose(stat["loss"].numpy(), g["stat_loss"], rtol=1e-6, atol=1e-7)
    assert np.array_equal(stat["vocab_min_logits"].numpy(), g["stat_vocab_min_logits"])
    assert np.array_equal(stat["vocab_max_logits"].numpy(), g["stat_vocab_max_logits"])


def test_gae_2d_oracle_reduces_to_1d():
    """The 2-D restatement (gae.cu:62-92) on rows without inner dones equals the pinned 1-D
    CUDA-order oracle (gae.cu:10-28) bit for bit; with dones, segments are independent."""
    rng = np.random.default_rng(0)
    B, H = 5, 40
    r = rng.standard_normal((B, H)).astype(np.float32)
    v = rng.standard_normal((B, H + 1)).astype(np.float32)
    done = np.zeros((B, H + 1), dtype=bool)
    done[:, H] = [True, False, True, False, False]
    cu0 = np.zeros(B + 1, np.int32)
    a2, t2 = cpu_ref.gae_2d_olp(r, v, done, np.zeros(0, np.int32), cu0, np.zeros(0, bool), 0.99, 0.95)
    cu1 = (np.arange(B + 1) * H).astype(np.int32)
    a1, t1 = cpu_ref.gae_cuda_order(r.reshape(-1), v.reshape(-1), cu1, ~done[:, H], 0.99, 0.95)
    assert np.array_equal(a2.reshape(-1), a1) and np.array_equal(t2.reshape(-1), t1)
    # one inner done at t=10 in row 0, truncated: adv[9] = 0, ret[9] = v[9]; first part unaffected by the tail
    done[0, 10] = True
    cu = np.array([0, 1, 1, 1, 1, 1], np.int32)
    a3, t3 = cpu_ref.gae_2d_olp(r, v, done, np.array([10], np.int32), cu, np.array([True]), 0.99, 0.95)
    assert a3[0, 9] == 0.0 and t3[0, 9] == v[0, 9]
    assert np.array_equal(a3[0, 10:], a2[0, 10:]) and np.array_equal(a3[1:], a2[1:])


def test_oracle_properties_on_random_ragged_inputs():
    """Size-independent properties of the oracle itself on seeded random inputs (empty sequences,
    single tokens, every bootstrap pattern; empty and touching intervals): the C statement of the
    reference kernel, the python-loop statement and the float64 statement agree; GAE is linear in
    (rewards, values); slice -> set -> slice is the identity on the covered elements."""
    rng = np.random.default_rng(7)
    for trial in range(20):
        B = int(rng.integers(1, 40))
        lens = rng.integers(0, 70, B)
        if lens.sum() == 0:
            lens[0] = 1
        total = int(lens.sum())
        cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        boot = (rng.random(B) < 0.5)
        gl = [(1.0, 1.0), (0.99, 0.95), (0.9, 0.5)][trial % 3]
        r1, r2 = (rng.standard_normal(total).astype(np.float32) for _ in range(2))
        v1, v2 = (rng.standard_normal(total + B).astype(np.float32) for _ in range(2))
        a_c, t_c = cpu_ref.gae_cuda_order(r1, v1, cu, boot, *gl)
        a_l, t_l = cpu_ref.gae_py_loops(r1, v1, cu, boot.astype(np.float32), *gl)
        a_d, t_d = cpu_ref.gae_f64(r1, v1, cu, boot, *gl)
        np.testing.assert_allclose(a_c, a_l, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(a_c, a_d, rtol=1e-4, atol=1e-4)
        np.testing.assert_allclose(t_c, t_d, rtol=1e-4, atol=1e-4)
        # linearity (exact in real arithmetic; fp32 rounding only)
        a_s, _ = cpu_ref.gae_cuda_order(r1 + r2, v1 + v2, cu, boot, *gl)
        a_2, _ = cpu_ref.gae_cuda_order(r2, v2, cu, boot, *gl)
        np.testing.assert_allclose(a_s, a_c + a_2, rtol=1e-4, atol=1e-3)
        # intervals: disjoint, unsorted, some empty, some touching
        size = 4000
        cuts = np.sort(rng.choice(size, 40, replace=False))
        iv = [(int(cuts[i]), int(cuts[i + 1])) for i in range(0, 38, 2)] + [(int(cuts[5]), int(cuts[5]))]
        iv += [(int(cuts[39]), int(cuts[39]) + 3)] if cuts[39] + 3 <= size else []
        rng.shuffle(iv)
        iv = np.asarray(iv, dtype=np.int64)
        src = rng.integers(0, 65535, size).astype(np.uint16)
        packed = cpu_ref.slice_intervals(src, iv)
        assert packed.size == int((iv[:, 1] - iv[:, 0]).sum())
        dst = np.full(size, 0xABCD, np.uint16)
        cpu_ref.set_intervals(packed, dst, iv)
        covered = np.zeros(size, bool)
        for a, b in iv:
            covered[a:b] = True
        assert np.array_equal(dst[covered], src[covered]) and np.all(dst[~covered] == 0xABCD)
        assert np.array_equal(cpu_ref.slice_intervals(dst, iv), packed)
