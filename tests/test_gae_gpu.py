"""H3 parity on the GPU. Bar: 1e-5 (north_star), as the reference test states it —
allclose(atol=1e-5, rtol=1e-5) (realhf/tests/cpp_extensions/test_cugae.py:50-51)."""

import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden

pytestmark = pytest.mark.gpu

GAE = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "gae_*.npz"))
             if "padded" not in p)


def _run(g_or_tuple, cuda, mode):
    from areal_vllm_b200 import ppo_functional as PF

    r, v, cu, b, gamma, lam = g_or_tuple
    adv, ret = PF.gae_1d_nolp_misalign(
        torch.from_numpy(r).to(cuda), torch.from_numpy(v).to(cuda),
        torch.from_numpy(cu).int().to(cuda), torch.from_numpy(np.asarray(b) != 0).to(cuda),
        gamma, lam, mode=mode)
    return adv.cpu().numpy(), ret.cpu().numpy()


def _make(B, lo, hi, seed, boot="mixed"):
    rng = np.random.default_rng(seed)
    lens = rng.integers(lo, hi + 1, B)
    total = int(lens.sum())
    r = rng.standard_normal(total).astype(np.float32)
    v = rng.standard_normal(total + B).astype(np.float32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    b = {"ones": np.ones(B), "zeros": np.zeros(B), "mixed": rng.random(B) < 0.5}[boot].astype(np.float32)
    return r, v, cu, b


@pytest.mark.parametrize("name", GAE)
@pytest.mark.parametrize("mode", ["auto", "serial", "scan"])
def test_golden(name, mode, cuda):
    g = load_golden(name)
    gamma, lam = float(g["gamma"]), float(g["lam"])
    adv, ret = _run((g["rewards"], g["values"], g["cu_seqlens"], g["bootstrap"], gamma, lam), cuda, mode)
    if mode == "scan" and gamma * lam > 0.99:
        # re-associated fp32 suffix sums with no forgetting: only the sequence-scale bar can hold
        scale = max(1.0, float(np.abs(g["adv"]).max()))
        assert np.abs(adv - g["adv"]).max() <= 1e-5 * scale * 4
        return
    np.testing.assert_allclose(adv, g["adv"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ret, g["ret"], rtol=1e-5, atol=1e-5)
    if mode != "scan" and gamma == 1.0 and lam == 1.0:
        assert np.array_equal(adv, g["adv"]) and np.array_equal(ret, g["ret"])  # same rounding sequence


@pytest.mark.parametrize("max_seqlen", [32, 128, 512])
@pytest.mark.parametrize("bs", [2, 4])
@pytest.mark.parametrize("gamma", [0.9, 1.0])
@pytest.mark.parametrize("lam", [0.5, 1.0])
def test_reference_test_grid(max_seqlen, bs, gamma, lam, cuda):
    """The parameter grid of realhf/tests/cpp_extensions/test_cugae.py:16-51 (bootstrap = ones)."""
    from oracle import cpu_ref

    r, v, cu, b = _make(bs, 1, max_seqlen - 1, 1000 + max_seqlen + bs, boot="ones")
    adv, ret = _run((r, v, cu, b, gamma, lam), cuda, "auto")
    a_ref, r_ref = cpu_ref.gae_py(r, v, cu, b, gamma, lam)
    np.testing.assert_allclose(adv, a_ref, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ret, r_ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("gl", [(0.99, 0.95), (1.0, 1.0)])
@pytest.mark.parametrize("boot", ["ones", "zeros", "mixed"])
def test_cfg1_full(gl, boot, cuda):
    """BASELINE config 1 at full size: 256 packed sequences, lengths 64-2048."""
    from oracle import cpu_ref

    r, v, cu, b = _make(256, 64, 2048, 1, boot)
    for mode in ("auto", "serial"):
        adv, ret = _run((r, v, cu, b, *gl), cuda, mode)
        a_ref, r_ref = cpu_ref.gae_py(r, v, cu, b, *gl)
        np.testing.assert_allclose(adv, a_ref, rtol=1e-5, atol=1e-5, err_msg=mode)
        np.testing.assert_allclose(ret, r_ref, rtol=1e-5, atol=1e-5, err_msg=mode)
    a_cuda, r_cuda = cpu_ref.gae_cuda_order(r, v, cu, b, *gl)
    adv, ret = _run((r, v, cu, b, *gl), cuda, "serial")
    assert np.array_equal(adv, a_cuda) and np.array_equal(ret, r_cuda)  # bit-identical to gae.cu's order


@pytest.mark.parametrize("c", [(0.9, 0.5), (0.99, 0.95), (0.995, 0.99), (1.0, 0.999)])
def test_scan_vs_oracle_long(c, cuda):
    """SCAN on long sequences (look-back across many tiles) and many tiny ones."""
    from oracle import cpu_ref

    for B, lo, hi, seed in [(7, 3000, 16384, 11), (3000, 1, 5, 12), (64, 1, 3000, 13), (1, 70000, 70000, 14)]:
        r, v, cu, b = _make(B, lo, hi, seed)
        adv, ret = _run((r, v, cu, b, *c), cuda, "scan")
        a64, r64 = cpu_ref.gae_f64(r, v, cu, b, *c)
        a_ref, _ = cpu_ref.gae_py(r, v, cu, b, *c)
        # never further from the exact value than a few times the reference's own fp32 error + 1e-5
        ref_err = np.abs(a_ref - a64).max()
        assert np.abs(adv - a64).max() <= 4 * ref_err + 1e-5, (B, np.abs(adv - a64).max(), ref_err)
        if c[0] * c[1] <= 0.99:
            np.testing.assert_allclose(adv, a_ref, rtol=1e-5, atol=1e-5)
            np.testing.assert_allclose(ret, cpu_ref.gae_py(r, v, cu, b, *c)[1], rtol=1e-5, atol=1e-5)


def test_edge_cases(cuda):
    from areal_vllm_b200 import ppo_functional as PF
    from oracle import cpu_ref

    # empty sequences in the middle, single-token sequences, tile-boundary-aligned sequences
    lens = np.array([0, 1, 1024, 0, 0, 2048, 1, 1023, 1, 0, 3000, 0], dtype=np.int64)
    rng = np.random.default_rng(3)
    total, B = int(lens.sum()), len(lens)
    r = rng.standard_normal(total).astype(np.float32)
    v = rng.standard_normal(total + B).astype(np.float32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    b = (rng.random(B) < 0.5).astype(np.float32)
    for mode in ("serial", "scan"):
        adv, ret = _run((r, v, cu, b, 0.97, 0.9), cuda, mode)
        a_ref, r_ref = cpu_ref.gae_py(r, v, cu, b, 0.97, 0.9)
        np.testing.assert_allclose(adv, a_ref, rtol=1e-5, atol=1e-5, err_msg=mode)
        np.testing.assert_allclose(ret, r_ref, rtol=1e-5, atol=1e-5, err_msg=mode)
    # gamma*lam == 0
    adv, ret = _run((r, v, cu, b, 0.0, 0.9), cuda, "scan")
    a_ref, r_ref = cpu_ref.gae_py(r, v, cu, b, 0.0, 0.9)
    np.testing.assert_allclose(adv, a_ref, rtol=1e-5, atol=1e-5)
    # inputs are not modified (pygae zeroes values in place; we must not)
    vt = torch.from_numpy(v).to(cuda)
    vt0 = vt.clone()
    PF.cugae1d_nolp_misalign_func(torch.from_numpy(r).to(cuda), vt, torch.from_numpy(cu).to(cuda),
                                  torch.zeros(B, dtype=torch.bool, device=cuda), 1.0, 1.0)
    assert torch.equal(vt, vt0)
    # argument errors of gae.cu:36-50
    with pytest.raises(RuntimeError):
        PF.gae_1d_nolp_misalign(torch.from_numpy(r), vt, torch.from_numpy(cu).to(cuda),
                                torch.zeros(B, dtype=torch.bool, device=cuda), 1.0, 1.0)
    with pytest.raises(RuntimeError):
        PF.gae_1d_nolp_misalign(torch.from_numpy(r).to(cuda), vt[:-1], torch.from_numpy(cu).to(cuda),
                                torch.zeros(B, dtype=torch.bool, device=cuda), 1.0, 1.0)
    with pytest.raises(RuntimeError):
        PF.gae_1d_nolp_misalign(torch.from_numpy(r).to(cuda), vt, torch.from_numpy(cu).long().to(cuda),
                                torch.zeros(B, dtype=torch.bool, device=cuda), 1.0, 1.0)
    # get_packed_advantages_and_returns casts like the reference (ppo_functional.py:383-389)
    a, _ = PF.get_packed_advantages_and_returns(0.97, 0.9, vt, torch.from_numpy(r).to(cuda),
                                                torch.from_numpy(cu).long().to(cuda),
                                                torch.from_numpy(b).to(cuda))
    np.testing.assert_allclose(a.cpu().numpy(), cpu_ref.gae_py(r, v, cu, b, 0.97, 0.9)[0], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("gl", [(1.0, 1.0), (0.99, 0.95)])
def test_padded_gae_adapter(gl, cuda):
    """actor.padded_gae (the [B, L] GAE of compute_advantages, actor.py:131-148) vs the oracle loop."""
    from areal_vllm_b200.actor import padded_gae
    from oracle import torch_ref

    gen = torch.Generator().manual_seed(8)
    B, L = 37, 301
    lens = torch.randint(5, L + 1, (B,), generator=gen)
    att = torch.arange(L)[None, :] < lens[:, None]
    rewards = torch.randn(B, L, generator=gen) * att
    values = torch.randn(B, L, generator=gen) * att
    no_eos = lens == L
    ref = torch_ref.padded_gae(rewards, values, no_eos, *gl)
    v0 = values.clone().to(cuda)
    out = padded_gae(rewards.to(cuda), v0, no_eos.to(cuda), *gl)
    assert out.shape == (B, L) and out.dtype == torch.float32
    assert torch.equal(v0.cpu(), values)  # inputs untouched
    np.testing.assert_allclose(out.cpu().numpy(), ref.numpy(), rtol=1e-5, atol=1e-5)
    assert float(out[:, -1].abs().max()) == 0.0


def test_padded_gae_golden(cuda):
    from areal_vllm_b200.actor import padded_gae

    g = load_golden("gae_padded")
    for k in range(2):
        t = lambda a: torch.from_numpy(a).to(cuda)
        out = padded_gae(t(g[f"rewards_{k}"]), t(g[f"values_{k}"]), t(g[f"no_eos_{k}"]),
                         float(g[f"gamma_{k}"]), float(g[f"lam_{k}"]))
        np.testing.assert_allclose(out.cpu().numpy(), g[f"adv_{k}"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("mode", ["serial_batch8", "serial_batch32", "serial_slots", "serial_relay", "serial_team"])
@pytest.mark.parametrize("B,lo,hi", [(1, 1, 5), (7, 1, 300), (33, 64, 2048), (1000, 1, 130), (4099, 0, 70),
                                     (5, 250, 1030), (200, 0, 3000)])
@pytest.mark.parametrize("gl", [(1.0, 1.0), (0.99, 0.95)])
def test_batched_serial_kernels_bit_identical(mode, B, lo, hi, gl, cuda):
    """The many-sequence serial kernels (one chain per lane) give the same bits as the
    one-warp-per-sequence kernel and as the C statement of the reference kernel (gae.cu:10-28),
    including empty sequences, ragged groups and every bootstrap pattern."""
    from areal_vllm_b200 import ppo_functional as PF
    from oracle import cpu_ref

    rng = np.random.default_rng(B + hi)
    lens = rng.integers(lo, hi + 1, B)
    total = int(lens.sum())
    if total == 0:
        lens[0] = 3
        total = 3
    r = rng.standard_normal(total).astype(np.float32)
    v = rng.standard_normal(total + B).astype(np.float32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    boot = rng.random(B) < 0.5
    to = lambda a: torch.from_numpy(a).to(cuda)
    a_ref, t_ref = cpu_ref.gae_cuda_order(r, v, cu, boot, *gl)
    a, t = PF.gae_1d_nolp_misalign(to(r), to(v), to(cu), to(boot), *gl, mode=mode)
    assert np.array_equal(a.cpu().numpy(), a_ref) and np.array_equal(t.cpu().numpy(), t_ref)
    a2, t2 = PF.gae_1d_nolp_misalign(to(r), to(v), to(cu), to(boot), *gl, mode="serial_warp")
    assert torch.equal(a, a2) and torch.equal(t, t2)


@pytest.mark.parametrize("B,lo,hi", [(3, 1, 200), (300, 0, 700), (9000, 30, 400)])
def test_slots_kernel_without_ticket_counter(B, lo, hi, cuda):
    """B200RL_GAE_SERIAL_SLOTS called through the C ABI with NO workspace (static round-robin over
    the slots instead of the ticket counter): same bits as the one-warp-per-sequence kernel."""
    from areal_vllm_b200 import _lib
    from areal_vllm_b200 import ppo_functional as PF

    rng = np.random.default_rng(B)
    lens = rng.integers(lo, hi + 1, B)
    lens[0] = max(lens[0], 1)
    total = int(lens.sum())
    to = lambda a: torch.from_numpy(a).to(cuda)
    r, v = to(rng.standard_normal(total).astype(np.float32)), to(rng.standard_normal(total + B).astype(np.float32))
    cu = to(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32))
    boot = to(rng.random(B) < 0.5)
    adv, ret = torch.full_like(r, 7.0), torch.full_like(r, 7.0)
    rc = _lib.lib.b200rl_gae_1d_nolp_misalign(_lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(boot.view(torch.uint8)),
                                              _lib.ptr(adv), _lib.ptr(ret), B, total, 1.0, 1.0, 6, None, 0,
                                              _lib.stream_handle(cuda))
    assert rc == 0, _lib.last_error()
    a2, t2 = PF.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 1.0, mode="serial_warp")
    assert torch.equal(adv, a2) and torch.equal(ret, t2)


def test_serial_auto_dispatch_large_batch(cuda):
    """many short sequences take the sequences-per-warp kernels under mode="serial": same bits."""
    from areal_vllm_b200 import ppo_functional as PF

    rng = np.random.default_rng(1)
    B = 40000
    lens = rng.integers(1, 40, B)
    total = int(lens.sum())
    to = lambda a: torch.from_numpy(a).to(cuda)
    r, v = to(rng.standard_normal(total).astype(np.float32)), to(rng.standard_normal(total + B).astype(np.float32))
    cu = to(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32))
    boot = to(rng.random(B) < 0.5)
    a, t = PF.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 1.0, mode="serial")
    a2, t2 = PF.gae_1d_nolp_misalign(r, v, cu, boot, 1.0, 1.0, mode="serial_warp")
    assert torch.equal(a, a2) and torch.equal(t, t2)


@pytest.mark.parametrize("gl", [(0.99, 0.95), (1.0, 1.0)])
def test_cfg4_full(gl, cuda):
    """BASELINE config 4's GAE shape at full size: 1,048,576 tokens in sequences of 1,024-16,384
    tokens (seed 4). AUTO against the C port of pygae (1e-5), SERIAL bit-identical to the
    reference kernel's order."""
    from oracle import cpu_ref

    rng = np.random.default_rng(4)
    lens = []
    while sum(lens) < 1_048_576:
        lens.append(int(rng.integers(1024, 16385)))
    lens[-1] -= sum(lens) - 1_048_576
    B, total = len(lens), 1_048_576
    r = rng.standard_normal(total).astype(np.float32)
    v = rng.standard_normal(total + B).astype(np.float32)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    b = (rng.random(B) < 0.5).astype(np.float32)
    a_ref, r_ref = cpu_ref.gae_py(r, v, cu, b, *gl)
    adv, ret = _run((r, v, cu, b, *gl), cuda, "auto")
    np.testing.assert_allclose(adv, a_ref, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ret, r_ref, rtol=1e-5, atol=1e-5)
    a_cuda, r_cuda = cpu_ref.gae_cuda_order(r, v, cu, b, *gl)
    adv, ret = _run((r, v, cu, b, *gl), cuda, "serial")
    assert np.array_equal(adv, a_cuda) and np.array_equal(ret, r_cuda)


def test_validate_flag_reproduces_the_reference_check(cuda):
    """gae.cu:48-49 / ppo_functional.py:365: cu_seqlens must start at 0 and end at total. Off by
    default (it costs a host sync); validate=True raises the RuntimeError the reference raises."""
    from areal_vllm_b200 import ppo_functional as PF

    r = torch.randn(10, device=cuda)
    v = torch.randn(12, device=cuda)
    b = torch.ones(2, dtype=torch.bool, device=cuda)
    good = torch.tensor([0, 4, 10], dtype=torch.int32, device=cuda)
    PF.gae_1d_nolp_misalign(r, v, good, b, 0.9, 0.9, validate=True)
    for bad in ([1, 4, 10], [0, 4, 9]):
        with pytest.raises(RuntimeError, match="cu_seqlens must start at 0"):
            PF.gae_1d_nolp_misalign(r, v, torch.tensor(bad, dtype=torch.int32, device=cuda), b, 0.9, 0.9, validate=True)
