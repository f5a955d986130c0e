"""Padded 2-D GAE (ref csrc/cugae/gae.cu:62-206) against the C oracle (bit-exact: same operation
order and FMA contractions) and, when oracle/_ref/ref_cugae.so was built from the reference's own
gae.cu, against the reference kernel itself - both bindings (gae_2d_nolp launches the olp kernel
in the reference, gae.cu:199)."""

import numpy as np
import pytest
import torch

from oracle import cpu_ref

pytestmark = pytest.mark.gpu


def _inputs(B, H, p_done, seed):
    rng = np.random.default_rng(seed)
    r = rng.standard_normal((B, H)).astype(np.float32)
    v = rng.standard_normal((B, H + 1)).astype(np.float32)
    done = np.zeros((B, H + 1), dtype=bool)
    done[:, 1:H] = rng.random((B, H - 1)) < p_done if H > 1 else False
    ys, cu = [], [0]
    for b in range(B):
        y = np.nonzero(done[b, :H])[0]
        ys.append(y)
        cu.append(cu[-1] + len(y))
    done[:, H] = rng.random(B) < 0.5  # the step after the horizon: bootstraps the last segment or not
    y = np.concatenate(ys).astype(np.int32) if ys else np.zeros(0, np.int32)
    trunc = rng.random(len(y)) < 0.4
    return r, v, done, y, np.array(cu, dtype=np.int32), trunc, max((len(a) for a in ys), default=0)


@pytest.mark.parametrize("B,H,p", [(1, 1, 0.0), (3, 7, 0.3), (64, 200, 0.05), (16, 4096, 0.002),
                                   (500, 33, 0.5), (8, 1000, 0.0)])
@pytest.mark.parametrize("gl", [(0.99, 0.95), (1.0, 1.0)])
def test_gae_2d_bit_exact_vs_oracle(B, H, p, gl, cuda):
    from areal_vllm_b200 import ppo_functional as PF

    r, v, done, y, cu, trunc, mx = _inputs(B, H, p, seed=B + H)
    a_ref, t_ref = cpu_ref.gae_2d_olp(r, v, done, y, cu, trunc, *gl)
    to = lambda a: torch.from_numpy(a).to(cuda)
    adv, ret = PF.gae_2d_olp(to(r), to(v), to(done), to(y), to(cu), mx, to(trunc), *gl)
    assert np.array_equal(adv.cpu().numpy(), a_ref)
    assert np.array_equal(ret.cpu().numpy(), t_ref)
    adv2, ret2 = PF.gae_2d_nolp(to(r), to(v), to(done), to(y), to(cu), mx, to(trunc), *gl)
    assert torch.equal(adv2, adv) and torch.equal(ret2, ret)


def test_gae_2d_single_segment_equals_1d(cuda):
    """rows without inner dones are plain sequences: same numbers as the packed 1-D kernel"""
    from areal_vllm_b200 import ppo_functional as PF

    B, H = 12, 300
    r, v, done, y, cu, trunc, mx = _inputs(B, H, 0.0, seed=4)
    to = lambda a: torch.from_numpy(a).to(cuda)
    adv, ret = PF.gae_2d_olp(to(r), to(v), to(done), to(y), to(cu), mx, to(trunc), 0.99, 0.95)
    cu1 = torch.arange(B + 1, dtype=torch.int32, device=cuda) * H
    boot = ~to(done[:, H])
    a1, t1 = PF.gae_1d_nolp_misalign(to(r).reshape(-1), to(v).reshape(-1), cu1, boot, 0.99, 0.95, mode="serial")
    assert torch.equal(adv.reshape(-1), a1) and torch.equal(ret.reshape(-1), t1)


@pytest.mark.parametrize("B,H,p", [(64, 200, 0.05), (7, 2000, 0.01)])
def test_gae_2d_vs_reference_kernel(B, H, p, cuda):
    from areal_vllm_b200 import ppo_functional as PF
    from oracle import ref_loader as R

    m = R.ref_cugae_cuda()
    if m is None:
        pytest.skip("oracle/_ref/ref_cugae.so not built")
    r, v, done, y, cu, trunc, mx = _inputs(B, H, p, seed=9)
    to = lambda a: torch.from_numpy(a).to(cuda)
    args = (to(r), to(v), to(done), to(y), to(cu), mx, to(trunc), 0.99, 0.95)
    a_ref, t_ref = m.gae_2d_olp(*args)
    adv, ret = PF.gae_2d_olp(*args)
    assert torch.equal(adv, a_ref) and torch.equal(ret, t_ref)
    a_ref, t_ref = m.gae_2d_nolp(*args)
    adv, ret = PF.gae_2d_nolp(*args)
    assert torch.equal(adv, a_ref) and torch.equal(ret, t_ref)


def test_gae_2d_errors(cuda):
    from areal_vllm_b200 import ppo_functional as PF

    r, v, done, y, cu, trunc, mx = _inputs(4, 10, 0.3, seed=1)
    to = lambda a: torch.from_numpy(a).to(cuda)
    good = [to(r), to(v), to(done), to(y), to(cu), mx, to(trunc), 0.99, 0.95]
    for idx, bad in ((1, to(v)[:, :-1].contiguous()), (3, to(y).long()), (6, to(trunc).int()),
                     (0, torch.from_numpy(r)), (4, to(cu)[:-1].contiguous())):
        a = list(good)
        a[idx] = bad
        with pytest.raises(RuntimeError):
            PF.gae_2d_olp(*a)
