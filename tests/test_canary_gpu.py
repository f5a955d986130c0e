"""Out-of-bounds WRITE check without compute-sanitizer (closed on this pool): every output of every
kernel is carved out of a larger buffer pre-filled with a sentinel; after the call the guard bytes
on both sides must be untouched. Calls go through the raw C ABI so the test owns all buffers."""

import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096  # bytes on each side


class Guarded:
    def __init__(self, nbytes, dev):
        self.n = nbytes
        self.buf = torch.full((GUARD + nbytes + GUARD + 256,), 0xA5, dtype=torch.uint8, device=dev)
        off = (-self.buf.data_ptr() - GUARD) % 256  # keep the payload 256-B aligned
        self.off = GUARD + off
        self.view = self.buf[self.off:self.off + nbytes]

    def ptr(self):
        return ctypes.c_void_p(self.view.data_ptr())

    def intact(self):
        b = self.buf
        return bool((b[:self.off] == 0xA5).all()) and bool((b[self.off + self.n:] == 0xA5).all())

    def as_(self, dtype):
        return self.view.view(dtype)


def test_h1_and_fused_outputs_stay_in_bounds(cuda):
    from areal_vllm_b200 import _lib

    L, st = _lib.lib, _lib.stream_handle(cuda)
    g = torch.Generator().manual_seed(0)
    for V in (4104, 151936):
        T = 150 if V < 10000 else 20
        x = (torch.randn(T, V, generator=g) * 2).bfloat16().to(cuda)
        y = torch.randint(0, V, (T,), generator=g).to(cuda)
        outs = [Guarded(T * 4, cuda) for _ in range(5)]
        rc = L.b200rl_logprob_entropy_fwd(_lib.ptr(x), 0, V, _lib.ptr(y), T, V, 1.0, *[o.ptr() for o in outs], st)
        assert rc == 0, _lib.last_error()
        glp = torch.randn(T, device=cuda)
        d = Guarded(T * V * 2, cuda)
        rc = L.b200rl_logprob_entropy_bwd(_lib.ptr(x), 0, V, _lib.ptr(y), outs[2].ptr(), outs[1].ptr(), _lib.ptr(glp),
                                          None, T, V, 1.0, d.ptr(), V, st)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert all(o.intact() for o in outs) and d.intact()
        # fused step
        f32 = [Guarded(T * 4, cuda) for _ in range(10)]   # logp, ent, lse, max, min, s_loss, s_iw, s_kl, s_bw, s_bkl
        u8 = [Guarded(T, cuda) for _ in range(3)]
        d2, loss, den = Guarded(T * V * 2, cuda), Guarded(4, cuda), Guarded(4, cuda)
        ws = Guarded(L.b200rl_grpo_fused_workspace_bytes(T), cuda)
        prox, old, adv = (torch.randn(T, device=cuda) for _ in range(3))
        mask = (torch.rand(T, device=cuda) < 0.7).view(torch.uint8)
        rc = L.b200rl_grpo_fused_step(
            _lib.ptr(x), 0, V, _lib.ptr(y), T, V, 1.0, _lib.ptr(prox), _lib.ptr(old), _lib.ptr(adv), _lib.ptr(mask),
            0.8, 1.2, 3.0, 5.0, None, f32[0].ptr(), f32[1].ptr(), f32[2].ptr(), f32[3].ptr(), f32[4].ptr(),
            f32[5].ptr(), f32[6].ptr(), f32[7].ptr(), u8[0].ptr(), u8[1].ptr(), f32[8].ptr(), f32[9].ptr(), u8[2].ptr(),
            d2.ptr(), V, loss.ptr(), den.ptr(), ws.ptr(), ws.n, st)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert all(o.intact() for o in f32 + u8 + [d2, loss, den, ws])
        np.testing.assert_allclose(d2.as_(torch.bfloat16).float().abs().sum().item() > 0, True)


def test_gae_loss_moments_outputs_stay_in_bounds(cuda):
    from areal_vllm_b200 import _lib

    L, st = _lib.lib, _lib.stream_handle(cuda)
    rng = np.random.default_rng(1)
    lens = np.array([0, 1, 1024, 0, 2048, 1, 1023, 3001, 5, 0])
    total, B = int(lens.sum()), len(lens)
    r = torch.randn(total, device=cuda); v = torch.randn(total + B, device=cuda)
    cu = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)).to(cuda)
    b = torch.from_numpy(rng.random(B) < 0.5).to(cuda)
    for mode in (1, 2):
        adv, ret = Guarded(total * 4, cuda), Guarded(total * 4, cuda)
        ws = Guarded(L.b200rl_gae_workspace_bytes(B, total), cuda)
        rc = L.b200rl_gae_1d_nolp_misalign(_lib.ptr(r), _lib.ptr(v), _lib.ptr(cu), _lib.ptr(b), adv.ptr(), ret.ptr(), B, total,
                                           0.99, 0.95, mode, ws.ptr(), ws.n, st)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert adv.intact() and ret.intact() and ws.intact()
    T = 5000
    ins = [torch.randn(T, device=cuda) for _ in range(4)]
    mask = (torch.rand(T, device=cuda) < 0.6).view(torch.uint8)
    f = [Guarded(T * 4, cuda) for _ in range(6)]
    u = [Guarded(T, cuda) for _ in range(3)]
    loss, den = Guarded(4, cuda), Guarded(4, cuda)
    ws = Guarded(L.b200rl_ppo_loss_workspace_bytes(T), cuda)
    rc = L.b200rl_ppo_loss_fwd(*[_lib.ptr(t) for t in ins], _lib.ptr(mask), T, 0.8, 1.2, 3.0, 5.0, None, None, None, None,
                               f[0].ptr(), f[1].ptr(), f[2].ptr(), u[0].ptr(), u[1].ptr(), f[3].ptr(), f[4].ptr(), u[2].ptr(),
                               f[5].ptr(), loss.ptr(), den.ptr(), ws.ptr(), ws.n, st)
    assert rc == 0, _lib.last_error()
    x = torch.randn(18 * 12, device=cuda)
    mom = Guarded(3 * 3 * 8, cuda)
    ws2 = Guarded(L.b200rl_moments_workspace_bytes(216, 72), cuda)
    rc = L.b200rl_masked_moments(_lib.ptr(x), None, 216, 72, None, 1, mom.ptr(), ws2.ptr(), ws2.n, st)
    assert rc == 0, _lib.last_error()
    out = Guarded(216 * 4, cuda)
    rc = L.b200rl_masked_normalize(_lib.ptr(x), None, 216, None, 1, None, 1, 0, out.ptr(), st)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert all(o.intact() for o in f + u + [loss, den, ws, mom, ws2, out])


@pytest.mark.parametrize("es", [1, 2, 4, 8])
def test_interval_outputs_stay_in_bounds(es, cuda):
    from areal_vllm_b200 import _lib

    L, st = _lib.lib, _lib.stream_handle(cuda)
    rng = np.random.default_rng(es)
    iv, pos = [], 0
    for k in range(3000):
        pos += int(rng.integers(0, 23)); ln = int(rng.integers(0, 700)); iv.append((pos, pos + ln)); pos += ln
    iv.append((pos + 3, pos + 3 + 300001)); size = pos + 3 + 300001 + 7
    iv = np.array(iv, dtype=np.int64)[rng.permutation(len(iv))]
    total = int((iv[:, 1] - iv[:, 0]).sum())
    src = torch.randint(0, 255, (size * es,), dtype=torch.uint8, device=cuda)
    ivc = torch.from_numpy(iv).to(cuda)
    packed = Guarded(total * es, cuda)
    ws = Guarded(L.b200rl_interval_workspace_bytes(len(iv), total, es), cuda)
    rc = L.b200rl_slice_intervals(_lib.ptr(src), size, packed.ptr(), total, _lib.ptr(ivc), len(iv), es, ws.ptr(), ws.n, 0, st)
    assert rc == 0, _lib.last_error()
    big = Guarded(size * es, cuda)
    rc = L.b200rl_set_intervals(packed.ptr(), total, big.ptr(), size, _lib.ptr(ivc), len(iv), es, ws.ptr(), ws.n, 1, st)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert packed.intact() and ws.intact() and big.intact()
    # and nothing outside the intervals was written inside the big buffer
    touched = torch.zeros(size * es, dtype=torch.bool, device=cuda)
    for s_, e_ in iv.tolist():
        touched[s_ * es:e_ * es] = True
    assert bool((big.view[~touched] == 0xA5).all())
    assert torch.equal(big.view[touched], src[touched])
