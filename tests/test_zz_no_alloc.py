"""The allocation-free GAE entry point, called through the C ABI exactly as a host without a
device allocator would call it (a trainer squeezed next to an inference server that owns the rest
of HBM, or a plain C / FFI embedder): inputs and outputs come from the caller, the look-back
scratch is the array built into libb200rl.so.

It must agree with the workspace-taking entry point in SCAN mode (same kernel, same tolerance as
tests/test_gae_gpu.py: rtol = atol = 1e-5) on two calls in a row — the second call sees whatever the
first one left in the shared array, so this also checks that each call clears what it needs.
Collected last (file name) so that a run stopping at the first failure has already run the rest.
"""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(B, lo, hi, seed, dev):
    rng = np.random.default_rng(seed)
    lens = rng.integers(lo, hi + 1, B)
    total = int(lens.sum())
    r = torch.from_numpy(rng.standard_normal(total).astype(np.float32)).to(dev)
    v = torch.from_numpy(rng.standard_normal(total + B).astype(np.float32)).to(dev)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(dev)
    b = torch.from_numpy(rng.random(B) < 0.5).to(dev)
    return r, v, cu, b, B, total


def _with_workspace(r, v, cu, b, B, total, gamma, lam, dev):
    from areal_vllm_b200 import _lib

    adv, ret = torch.empty_like(r), torch.empty_like(r)
    ws = _lib.workspace(_lib.lib.b200rl_gae_workspace_bytes(B, total), dev)
    rc = _lib.lib.b200rl_gae_1d_nolp_misalign(
        _lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), _lib.ptr(adv), _lib.ptr(ret), B, total,
        gamma, lam, _lib.GAE_SCAN, _lib.ptr(ws), ws.numel(), _lib.stream_handle(dev))
    assert rc == 0, _lib.last_error()
    return adv, ret


def _no_alloc(r, v, cu, b, B, total, gamma, lam, dev):
    from areal_vllm_b200 import _lib

    adv, ret = torch.empty_like(r), torch.empty_like(r)  # outputs are the caller's
    rc = _lib.lib.b200rl_gae_1d_nolp_misalign_noalloc(
        _lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), _lib.ptr(adv), _lib.ptr(ret), B, total,
        gamma, lam, _lib.stream_handle(dev))
    assert rc == 0, _lib.last_error()
    return adv, ret


def test_no_alloc_matches_workspace_variant_twice(cuda):
    from areal_vllm_b200 import _lib

    gamma, lam = 0.99, 0.95
    # first call: long sequences (many tiles, look-back chains); second call: a different, smaller
    # batch of short sequences, which reuses the front of the same scratch array
    for B, lo, hi, seed in [(37, 2000, 16384, 21), (900, 1, 700, 22)]:
        r, v, cu, b, B, total = _case(B, lo, hi, seed, cuda)
        a_ws, r_ws = _with_workspace(r, v, cu, b, B, total, gamma, lam, cuda)
        before = torch.cuda.memory_allocated(cuda)
        a_na, r_na = _no_alloc(r, v, cu, b, B, total, gamma, lam, cuda)
        torch.cuda.synchronize()
        # only the two output tensors were allocated by this test around the call
        assert torch.cuda.memory_allocated(cuda) - before <= 2 * (total * 4 + 1024)
        np.testing.assert_allclose(a_na.cpu().numpy(), a_ws.cpu().numpy(), rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(r_na.cpu().numpy(), r_ws.cpu().numpy(), rtol=1e-5, atol=1e-5)
        # and, like the workspace variant, it sits within 1e-5 of the reference's python loop
        from oracle import cpu_ref
        a_ref, r_ref = cpu_ref.gae_py(r.cpu().numpy(), v.cpu().numpy(), cu.cpu().numpy(),
                                      b.cpu().numpy().astype(np.float32), gamma, lam)
        np.testing.assert_allclose(a_na.cpu().numpy(), a_ref, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(r_na.cpu().numpy(), r_ref, rtol=1e-5, atol=1e-5)


def test_no_alloc_rejects_oversize(cuda):
    from areal_vllm_b200 import _lib

    limit = 4_194_304
    rc = _lib.lib.b200rl_gae_1d_nolp_misalign_noalloc(None, None, None, None, None, None, 1, limit + 1,
                                                      0.99, 0.95, _lib.stream_handle(cuda))
    assert rc == -1
    msg = _lib.last_error()
    assert str(limit) in msg and "workspace" in msg
    # exactly at the limit is accepted (one sequence of `limit` tokens)
    r = torch.zeros(limit, device=cuda)
    v = torch.zeros(limit + 1, device=cuda)
    cu = torch.tensor([0, limit], dtype=torch.int32, device=cuda)
    b = torch.ones(1, dtype=torch.bool, device=cuda)
    r[-1] = 1.0
    adv, ret = _no_alloc(r, v, cu, b, 1, limit, 0.5, 1.0, cuda)
    tail = adv[-4:].cpu().numpy()
    np.testing.assert_allclose(tail, [0.125, 0.25, 0.5, 1.0], rtol=1e-6)
    assert float(adv[0]) == 0.0


def test_no_alloc_takes_no_device_memory_and_any_stream(cuda):
    """torch's own counters only see torch's caching allocator; a cudaMalloc inside the library
    would slip past them. So the driver's free-memory figure is compared across warmed-up calls
    (pre-allocated outputs, synchronised on both sides), on the current stream and on a side stream."""
    from areal_vllm_b200 import _lib

    r, v, cu, b, B, total = _case(200, 50, 3000, 23, cuda)
    a_ws, r_ws = _with_workspace(r, v, cu, b, B, total, 0.99, 0.95, cuda)
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    side = torch.cuda.Stream(device=cuda)

    def call(stream_ptr):
        rc = _lib.lib.b200rl_gae_1d_nolp_misalign_noalloc(
            _lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), _lib.ptr(adv), _lib.ptr(ret), B, total,
            0.99, 0.95, stream_ptr)
        assert rc == 0, _lib.last_error()

    call(_lib.stream_handle(cuda))  # warm-up: the one-time symbol lookup happens here
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(cuda)
    for _ in range(3):
        call(_lib.stream_handle(cuda))
    torch.cuda.synchronize()
    side.wait_stream(torch.cuda.current_stream(cuda))
    adv.fill_(0)
    torch.cuda.synchronize()
    import ctypes
    call(ctypes.c_void_p(side.cuda_stream))
    side.synchronize()
    free1, _ = torch.cuda.mem_get_info(cuda)
    assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across allocation-free calls"
    np.testing.assert_allclose(adv.cpu().numpy(), a_ws.cpu().numpy(), rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ret.cpu().numpy(), r_ws.cpu().numpy(), rtol=1e-5, atol=1e-5)
