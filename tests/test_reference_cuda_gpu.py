"""Our kernels against the REAL reference CUDA kernels (csrc/cugae/gae.cu, csrc/interval_op/
interval_op.cu), compiled for sm_100a from the reference tree by oracle/build_ref_cuda.py in the
build container and shipped as oracle/_ref/*.so. Skipped when those binaries are absent."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_gae():
    from oracle import ref_loader as R

    m = R.ref_cugae_cuda()
    if m is None:
        pytest.skip("oracle/_ref/ref_cugae.so not built")
    return m


@pytest.fixture(scope="module")
def ref_iv():
    from oracle import ref_loader as R

    m = R.ref_interval_op_cuda()
    if m is None:
        pytest.skip("oracle/_ref/ref_interval_op_cuda.so not built")
    return m


def _gae_inputs(B, lo, hi, seed, cuda):
    rng = np.random.default_rng(seed)
    lens = rng.integers(lo, hi + 1, B)
    total = int(lens.sum())
    r = torch.from_numpy(rng.standard_normal(total).astype(np.float32)).to(cuda)
    v = torch.from_numpy(rng.standard_normal(total + B).astype(np.float32)).to(cuda)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(cuda)
    b = torch.from_numpy(rng.random(B) < 0.5).to(cuda)
    return r, v, cu, b


@pytest.mark.parametrize("gl", [(1.0, 1.0), (0.99, 0.95), (0.9, 0.5), (1.0, 0.97)])
@pytest.mark.parametrize("shape", [(256, 64, 2048), (5, 1, 9), (40, 3000, 16384)])
def test_gae_vs_reference_kernel(gl, shape, cuda, ref_gae):
    from areal_vllm_b200 import ppo_functional as PF

    r, v, cu, b = _gae_inputs(*shape, seed=31, cuda=cuda)
    a_ref, r_ref = ref_gae.gae_1d_nolp_misalign(r, v, cu, b, gl[0], gl[1])
    a, t = PF.gae_1d_nolp_misalign(r, v, cu, b, gl[0], gl[1], mode="serial")
    assert torch.equal(a, a_ref) and torch.equal(t, r_ref)  # same order, same contractions: same bits
    a, t = PF.gae_1d_nolp_misalign(r, v, cu, b, gl[0], gl[1], mode="auto")
    np.testing.assert_allclose(a.cpu().numpy(), a_ref.cpu().numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(t.cpu().numpy(), r_ref.cpu().numpy(), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("gl", [(1.0, 1.0), (0.99, 0.95)])
@pytest.mark.parametrize("shape,kernel", [((2048, 30, 1500), "relay"), ((6000, 64, 700), "relay"),
                                          ((33000, 1, 60), "slots"), ((20000, 1, 50), "batch8"),
                                          ((120, 1024, 16384), "team"), ((512, 300, 9000), "team")])
def test_gae_large_batches_vs_reference_kernel(gl, shape, kernel, cuda, ref_gae):
    """Batch shapes that SERIAL dispatches to the relay / slots / batched / team kernels (gae.cu's dispatch
    comment), and each of those kernels forced, against the reference kernel: same bits."""
    from areal_vllm_b200 import ppo_functional as PF

    r, v, cu, b = _gae_inputs(*shape, seed=37, cuda=cuda)
    a_ref, r_ref = ref_gae.gae_1d_nolp_misalign(r, v, cu, b, gl[0], gl[1])
    for mode in ("serial", "serial_relay", "serial_slots", "serial_team"):
        a, t = PF.gae_1d_nolp_misalign(r, v, cu, b, gl[0], gl[1], mode=mode)
        assert torch.equal(a, a_ref) and torch.equal(t, r_ref), (mode, kernel)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("n", [1000, 20000])
def test_intervals_vs_reference_kernel(dtype, n, cuda, ref_iv):
    from areal_vllm_b200 import flatten_param as FP

    size = 4_000_000
    s = size // n
    iv = np.array([(i * s, i * s + s // 2) for i in range(n)], dtype=np.int64)
    np.random.default_rng(n).shuffle(iv)
    bits = torch.randint(-32768, 32767, (size * (2 if dtype == torch.float32 else 1),), dtype=torch.int16, device=cuda)
    src = bits.view(dtype)
    ivc = torch.from_numpy(iv).to(cuda)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    suffix = {torch.bfloat16: "bf16", torch.float16: "fp16", torch.float32: "fp32"}[dtype]
    ref_out = getattr(ref_iv, "slice_intervals_" + suffix)(src, ivc, total, s // 2)
    out = FP.slice_intervals(src, intervals_cuda=ivc, output_size=total, max_interval_size=s // 2)
    assert torch.equal(out.view(torch.int16), ref_out.view(torch.int16))
    d_ref = torch.zeros_like(src)
    d_our = torch.zeros_like(src)
    getattr(ref_iv, "set_intervals_" + suffix)(ref_out, d_ref, ivc, s // 2)
    FP.set_intervals(out, d_our, intervals_cuda=ivc, max_interval_size=s // 2)
    assert torch.equal(d_our.view(torch.int16), d_ref.view(torch.int16))
