"""H4 parity on the GPU: bit-exact against the reference-generated goldens and the C oracle."""

import numpy as np
import pytest
import torch

from conftest import load_golden

pytestmark = pytest.mark.gpu

DT = {"bf16_a": torch.bfloat16, "fp32_b": torch.float32, "fp16_c": torch.float16, "adv": torch.bfloat16}


def _bits_tensor(bits: np.ndarray, dtype, cuda):
    t = torch.from_numpy(bits.view(np.int16).copy())
    return t.view(dtype).to(cuda)


def _same_bits(t: torch.Tensor, bits: np.ndarray) -> bool:
    return np.array_equal(t.cpu().contiguous().view(torch.int16).numpy().view(np.uint16), bits)


@pytest.mark.parametrize("tag", list(DT))
def test_golden_bit_exact(tag, cuda):
    from areal_vllm_b200 import flatten_param as FP

    g = load_golden("h4_intervals")
    dtype = DT[tag]
    iv = g[f"{tag}_intervals"]
    src = _bits_tensor(g[f"{tag}_src_bits"], dtype, cuda)
    iv_cuda = torch.from_numpy(iv).to(cuda)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    out = FP.slice_intervals(src, [tuple(x) for x in iv.tolist()], intervals_cuda=iv_cuda,
                             output_size=total, max_interval_size=int((iv[:, 1] - iv[:, 0]).max()))
    assert out.dtype == dtype and out.shape == (total,)
    assert _same_bits(out, g[f"{tag}_slice_bits"])
    dst = torch.zeros_like(src)
    r = FP.set_intervals(out, dst, intervals_cuda=iv_cuda)
    assert r is None
    assert _same_bits(dst, g[f"{tag}_set_bits"])
    # the interval list alone (uploaded internally), and the typed pybind-style names
    out2 = FP.slice_intervals(src, [tuple(x) for x in iv.tolist()])
    assert torch.equal(out2.view(torch.int16) if dtype != torch.float32 else out2.view(torch.int32),
                       out.view(torch.int16) if dtype != torch.float32 else out.view(torch.int32))
    fn = {torch.bfloat16: FP.slice_intervals_bf16, torch.float16: FP.slice_intervals_fp16,
          torch.float32: FP.slice_intervals_fp32}[dtype]
    out3 = fn(src, iv_cuda, total, 0)
    assert _same_bits(out3, g[f"{tag}_slice_bits"])


@pytest.mark.parametrize("n_intervals", [1, 100, 500, 1000, 2000, 4000, 10000, 100000])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
def test_reference_test_grid(n_intervals, dtype, cuda):
    """Grid of realhf/tests/cpp_extensions/test_interval_ops.py:39-141 (skipped upstream), with the
    reference's interval generator (:20-31) at 1e7 elements and random BIT PATTERNS (NaN payloads,
    denormals included): slice then set, compared bitwise with the C oracle."""
    from areal_vllm_b200 import flatten_param as FP
    from oracle import cpu_ref

    size = int(1e7)
    s = size // n_intervals
    iv = np.array([(i * s, i * s + s // 2) for i in range(n_intervals)], dtype=np.int64)
    np.random.default_rng(n_intervals).shuffle(iv)
    npdt = np.uint32 if dtype == torch.float32 else np.uint16
    rng = np.random.default_rng(7)
    src_np = rng.integers(0, np.iinfo(npdt).max, size, dtype=npdt, endpoint=True)
    src = torch.from_numpy(src_np.view(np.int32 if npdt == np.uint32 else np.int16)).view(dtype).to(cuda)
    iv_cuda = torch.from_numpy(iv).to(cuda)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    out = FP.slice_intervals(src, intervals_cuda=iv_cuda, output_size=total)
    ref = cpu_ref.slice_intervals(src_np, iv)
    got = out.cpu().view(torch.int32 if npdt == np.uint32 else torch.int16).numpy().view(npdt)
    assert np.array_equal(got, ref)
    dst = torch.zeros(size, dtype=dtype, device=cuda)
    FP.set_intervals(out, dst, intervals_cuda=iv_cuda)
    dst_ref = np.zeros(size, dtype=npdt)
    cpu_ref.set_intervals(ref, dst_ref, iv)
    got = dst.cpu().view(torch.int32 if npdt == np.uint32 else torch.int16).numpy().view(npdt)
    assert np.array_equal(got, dst_ref)
    # plan reuse: the second call with the same intervals tensor skips the planning kernels
    out_again = FP.slice_intervals(src, intervals_cuda=iv_cuda, output_size=total)
    assert torch.equal(out_again.view(torch.uint8), out.view(torch.uint8))


def test_misaligned_and_skewed(cuda):
    """Every (src, dst) 16-byte phase combination, lengths 0..40 elements, plus one very long
    interval and empty intervals; uint8-, 2-, 4- and 8-byte elements through the raw C ABI."""
    import ctypes

    from areal_vllm_b200 import _lib
    from oracle import cpu_ref

    rng = np.random.default_rng(11)
    for es, npdt in ((1, np.uint8), (2, np.uint16), (4, np.uint32), (8, np.uint64)):
        size = 3_000_000
        src_np = rng.integers(0, 255, size * es, dtype=np.uint8).view(npdt)
        iv = []
        pos = 0
        for k in range(4000):
            pos += int(rng.integers(0, 23))
            ln = int(rng.integers(0, 41))
            iv.append((pos, pos + ln))
            pos += ln
        iv.append((pos + 3, pos + 3 + 1_500_001))  # long, odd offset, odd length
        iv.append((5, 5))
        order = rng.permutation(len(iv))
        iv = np.array(iv, dtype=np.int64)[order]
        total = int((iv[:, 1] - iv[:, 0]).sum())
        src = torch.from_numpy(src_np.view(np.uint8)).to(cuda)
        out = torch.empty(total * es, dtype=torch.uint8, device=cuda)
        ivc = torch.from_numpy(iv).to(cuda)
        ws = _lib.workspace(_lib.lib.b200rl_interval_workspace_bytes(len(iv), total, es), cuda)
        st = _lib.stream_handle(cuda)
        rc = _lib.lib.b200rl_slice_intervals(_lib.ptr(src), size, _lib.ptr(out), total, _lib.ptr(ivc), len(iv),
                                             es, _lib.ptr(ws), ws.numel(), 0, st)
        assert rc == 0, _lib.last_error()
        ref = cpu_ref.slice_intervals(src_np, iv)
        assert np.array_equal(out.cpu().numpy().view(npdt), ref), es
        hdr = ws[:16].cpu().numpy().view(np.int64)
        assert hdr[0] == total and hdr[1] == 0
        dst = torch.zeros(size * es, dtype=torch.uint8, device=cuda)
        rc = _lib.lib.b200rl_set_intervals(_lib.ptr(out), total, _lib.ptr(dst), size, _lib.ptr(ivc), len(iv),
                                           es, _lib.ptr(ws), ws.numel(), 1, st)  # reuse the plan
        assert rc == 0, _lib.last_error()
        dst_ref = np.zeros(size, dtype=npdt)
        cpu_ref.set_intervals(ref, dst_ref, iv)
        assert np.array_equal(dst.cpu().numpy().view(npdt), dst_ref), es


def test_round_trip_cfg5_scaled(cuda):
    """BASELINE config 5 shape scaled to what every B200 box can hold next to other tests
    (1/8 of the 7B flat buffer): slice -> set round trip restores exactly the covered elements;
    checksum-of-chunks equality instead of a host copy of the whole buffer."""
    from areal_vllm_b200 import flatten_param as FP

    size = 7_615_616_512 // 8
    n = 100000 // 8
    s = size // n
    iv = np.array([(i * s, i * s + s // 2) for i in range(n)], dtype=np.int64)
    np.random.default_rng(5).shuffle(iv)
    gen = torch.Generator(device=cuda).manual_seed(5)
    src = torch.randint(-32768, 32767, (size,), dtype=torch.int16, device=cuda, generator=gen).view(torch.bfloat16)
    ivc = torch.from_numpy(iv).to(cuda)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    out = FP.slice_intervals(src, intervals_cuda=ivc, output_size=total)
    dst = torch.zeros_like(src)
    FP.set_intervals(out, dst, intervals_cuda=ivc)
    covered = torch.zeros(size, dtype=torch.bool, device=cuda)
    starts = torch.from_numpy(iv[:, 0]).to(cuda)
    idx = (starts[:, None] + torch.arange(s // 2, device=cuda)[None, :]).reshape(-1)
    covered[idx] = True
    a, b = src.view(torch.int16), dst.view(torch.int16)
    assert torch.equal(b[covered], a[covered])
    assert int(b[~covered].abs().max()) == 0
    # first interval's payload sits at the start of the packed buffer
    s0, e0 = iv[0]
    assert torch.equal(out.view(torch.int16)[: e0 - s0], a[s0:e0])


def test_errors(cuda):
    from areal_vllm_b200 import flatten_param as FP

    src = torch.zeros(100, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(RuntimeError):
        FP.slice_intervals(src.cpu(), [(0, 4)])
    with pytest.raises(NotImplementedError):
        FP.slice_intervals(src.double(), [(0, 4)])
    with pytest.raises(RuntimeError):
        FP.set_intervals(src, src.float(), [(0, 4)])
    with pytest.raises(RuntimeError):
        FP.slice_intervals(src, intervals_cuda=torch.zeros(3, 2, dtype=torch.int32, device=cuda), output_size=0)
    assert FP.slice_intervals(src, [], output_size=0).numel() == 0


def test_cfg5_full_size(cuda):
    """BASELINE config 5 at FULL size: 7,615,616,512-element bf16 flat buffer (15.2 GB), 100,000
    shuffled intervals covering half of it. Bit-exactness is checked through the structure of the
    interval list (row k of the [n, s] view contributes its first half), with no index tensors:
    slice order, set round trip, untouched gaps, and the fused move against slice + set."""
    from areal_vllm_b200 import flatten_param as FP

    free, _ = torch.cuda.mem_get_info()
    if free < 70 * 2**30:
        pytest.skip("needs ~60 GB of free HBM")
    size, n = 7_615_616_512, 100000
    s = size // n
    h = s // 2
    perm = np.random.default_rng(5).permutation(n)
    iv = np.stack([perm * s, perm * s + h], 1).astype(np.int64)
    src = torch.empty(size, dtype=torch.bfloat16, device=cuda)
    src.view(torch.int16).random_()  # raw bit patterns, NaN/Inf payloads included
    ivc = torch.from_numpy(iv).to(cuda)
    total = n * h
    rows = src[: n * s].view(torch.int16).view(n, s)
    packed = FP.slice_intervals(src, intervals_cuda=ivc, output_size=total)
    permc = torch.from_numpy(perm).to(cuda)
    for lo in range(0, n, 20000):  # packed row k is the first half of src row perm[k]
        assert torch.equal(packed.view(torch.int16).view(n, h)[lo:lo + 20000], rows[permc[lo:lo + 20000], :h])
    dst = torch.zeros(size, dtype=torch.bfloat16, device=cuda)
    FP.set_intervals(packed, dst, intervals_cuda=ivc)
    drows = dst[: n * s].view(torch.int16).view(n, s)
    assert torch.equal(drows[:, :h], rows[:, :h])
    assert int(drows[:, h:].abs().max()) == 0 and int(dst[n * s:].view(torch.int16).abs().max()) == 0
    del packed
    # fused move into a buffer cut differently: n/4 destination intervals of 4*h elements each
    nd = n // 4
    d_iv = np.stack([np.arange(nd) * (4 * s), np.arange(nd) * (4 * s) + 4 * h], 1).astype(np.int64)
    dst.zero_()
    FP.move_intervals(src, dst, ivc, torch.from_numpy(d_iv).to(cuda), total)
    got = dst[: nd * 4 * s].view(torch.int16).view(nd, 4 * s)
    for lo in range(0, nd, 5000):  # destination interval j holds source rows perm[4j..4j+3] (first halves)
        want = rows[permc[4 * lo: 4 * (lo + 5000)], :h].reshape(-1, 4 * h)
        assert torch.equal(got[lo:lo + 5000, : 4 * h], want)
    assert int(got[:, 4 * h:].abs().max()) == 0


def test_failed_call_leaves_no_plan_behind(cuda):
    """A call that raises after the plan workspace was handed out (here: a strided source) must
    not leave that uninitialised workspace in the plan cache: the corrected call with the SAME
    intervals tensor has to build the plan and give the right bytes (flatten_param)."""
    from areal_vllm_b200 import flatten_param as FP

    FP.clear_plan_cache()
    rng = np.random.default_rng(77)
    size, n = 200_003, 900
    s_ = size // n
    iv = np.array([(i * s_ + (i % 5), i * s_ + (i % 5) + 30 + (i % 11)) for i in range(n)], dtype=np.int64)
    rng.shuffle(iv)
    ivc = torch.from_numpy(iv).to(cuda)
    total = int((iv[:, 1] - iv[:, 0]).sum())
    big = torch.from_numpy(rng.integers(-32768, 32767, 2 * size, dtype=np.int16)).to(cuda).view(torch.bfloat16)
    strided = big[::2]
    with pytest.raises(RuntimeError):
        FP.slice_intervals(strided, intervals_cuda=ivc, output_size=total)
    assert len(FP._PLAN_CACHE) == 0
    good = strided.contiguous()
    out = FP.slice_intervals(good, intervals_cuda=ivc, output_size=total)
    ref = torch.cat([good[a:b] for a, b in iv.tolist()])
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))
    # same for set: strided destination first, then the real one
    dst = torch.zeros(2 * size, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(RuntimeError):
        FP.set_intervals(out, dst[::2], intervals_cuda=ivc)
    d2 = torch.zeros(size, dtype=torch.bfloat16, device=cuda)
    FP.set_intervals(out, d2, intervals_cuda=ivc)
    chk = torch.zeros(size, dtype=torch.bfloat16, device=cuda)
    off = 0
    for a, b in iv.tolist():
        chk[a:b] = out[off:off + b - a]
        off += b - a
    assert torch.equal(d2.view(torch.int16), chk.view(torch.int16))
