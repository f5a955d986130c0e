"""move_intervals: slice -> set fused into one kernel with no packed staging buffer (SURVEY §8f
row 3; the TODO at realhf/impl/model/nn/real_llm_api.py:719-720). Bit-exact against the oracle's
_slice_intervals_py followed by _set_intervals_py (flatten_param.py:65-90). The peer-memory
(NVLink) use of the same kernel is covered by tests/test_move_intervals_dist.py."""

import numpy as np
import pytest
import torch

from oracle import cpu_ref

pytestmark = pytest.mark.gpu


def _random_lists(rng, total, n_src, n_dst, src_size, dst_size, max_gap=9, empties=True):
    """two interval lists over the same `total` elements, cut differently, placed at random
    non-overlapping positions of their buffers, listed in shuffled order"""
    def cuts(n):
        c = np.sort(rng.integers(0, total + 1, n - 1))
        return np.diff(np.concatenate([[0], c, [total]]))

    def place(lens, size):
        slack = size - int(lens.sum())
        assert slack >= 0
        gaps = rng.multinomial(slack, np.ones(len(lens) + 1) / (len(lens) + 1))[:-1] if slack else np.zeros(len(lens), int)
        order = rng.permutation(len(lens))  # memory order != stream order
        start = np.zeros(len(lens), dtype=np.int64)
        pos = 0
        for k in order:
            pos += int(gaps[k])
            start[k] = pos
            pos += int(lens[k])
        return np.stack([start, start + lens], 1).astype(np.int64)

    ls, ld = cuts(n_src), cuts(n_dst)
    if not empties:
        pass
    return place(ls, src_size), place(ld, dst_size)


def _oracle_move(src_np, dst_np, s_iv, d_iv):
    cpu_ref.set_intervals(cpu_ref.slice_intervals(src_np, s_iv), dst_np, d_iv)


@pytest.mark.parametrize("dtype,npdt", [(torch.bfloat16, np.uint16), (torch.float32, np.uint32),
                                        (torch.float16, np.uint16)])
@pytest.mark.parametrize("total,n_src,n_dst", [(1, 1, 1), (4097, 7, 3), (300_000, 1000, 37),
                                              (300_000, 5, 4000), (2_000_003, 3000, 3000)])
def test_move_matches_slice_then_set(dtype, npdt, total, n_src, n_dst, cuda):
    from areal_vllm_b200 import flatten_param as FP

    rng = np.random.default_rng(total + n_src)
    src_size, dst_size = total * 2 + 11, total + total // 2 + 5
    s_iv, d_iv = _random_lists(rng, total, n_src, n_dst, src_size, dst_size)
    src_np = rng.integers(0, np.iinfo(npdt).max, src_size, dtype=npdt)  # raw bit patterns
    dst_np = rng.integers(0, np.iinfo(npdt).max, dst_size, dtype=npdt)
    tview = torch.int16 if npdt == np.uint16 else torch.int32
    src = torch.from_numpy(src_np.view(np.int16 if npdt == np.uint16 else np.int32)).to(cuda).view(dtype)
    dst = torch.from_numpy(dst_np.view(np.int16 if npdt == np.uint16 else np.int32)).to(cuda).view(dtype)
    s_c, d_c = torch.from_numpy(s_iv).to(cuda), torch.from_numpy(d_iv).to(cuda)
    FP.move_intervals(src, dst, s_c, d_c, total)
    _oracle_move(src_np, dst_np, s_iv, d_iv)
    assert np.array_equal(dst.view(tview).cpu().numpy().view(npdt), dst_np)
    # second call reuses the cached plan; total_size inferred
    dst.zero_()
    FP.move_intervals(src, dst, s_c, d_c)
    ref2 = np.zeros(dst_size, dtype=npdt)
    _oracle_move(src_np, ref2, s_iv, d_iv)
    assert np.array_equal(dst.view(tview).cpu().numpy().view(npdt), ref2)
    # and equals the two-kernel path of this package
    dst3 = torch.zeros_like(dst)
    FP.set_intervals(FP.slice_intervals(src, intervals_cuda=s_c, output_size=total), dst3, intervals_cuda=d_c)
    assert torch.equal(dst3.view(tview), dst.view(tview))


def test_move_all_byte_phases_raw_abi(cuda):
    """1-, 2-, 4-, 8-byte elements; tiny pieces at every mutual 16-byte phase plus one long
    misaligned piece (funnel-shift path), through the C ABI."""
    from areal_vllm_b200 import _lib

    rng = np.random.default_rng(3)
    for es, npdt in ((1, np.uint8), (2, np.uint16), (4, np.uint32), (8, np.uint64)):
        lens_s = list(rng.integers(0, 41, 3000)) + [1_200_001]
        total = int(np.sum(lens_s))
        lens_d = list(rng.integers(0, 57, 2000))
        rest = total - int(np.sum(lens_d))
        lens_d += [rest - 1_000_003, 1_000_003]
        def place(lens, gap_hi):
            iv, pos = [], int(rng.integers(0, 9))
            for ln in lens:
                pos += int(rng.integers(0, gap_hi))
                iv.append((pos, pos + int(ln)))
                pos += int(ln)
            return np.array(iv, dtype=np.int64), pos + 7
        s_iv, src_size = place(lens_s, 23)
        d_iv, dst_size = place(lens_d, 5)
        src_np = rng.integers(0, 255, src_size * es, dtype=np.uint8).view(npdt)
        dst_np = np.zeros(dst_size, dtype=npdt)
        src = torch.from_numpy(src_np.view(np.uint8)).to(cuda)
        dst = torch.zeros(dst_size * es, dtype=torch.uint8, device=cuda)
        s_c, d_c = torch.from_numpy(s_iv).to(cuda), torch.from_numpy(d_iv).to(cuda)
        ws = _lib.workspace(_lib.lib.b200rl_move_workspace_bytes(len(s_iv), len(d_iv), total, es), cuda)
        rc = _lib.lib.b200rl_move_intervals(_lib.ptr(src), src_size, _lib.ptr(dst), dst_size, _lib.ptr(s_c),
                                            len(s_iv), _lib.ptr(d_c), len(d_iv), total, es, _lib.ptr(ws),
                                            ws.numel(), 0, _lib.stream_handle(cuda))
        assert rc == 0, _lib.last_error()
        _oracle_move(src_np, dst_np, s_iv, d_iv)
        assert np.array_equal(dst.cpu().numpy().view(npdt), dst_np), es


def test_move_errors(cuda):
    from areal_vllm_b200 import _lib
    from areal_vllm_b200 import flatten_param as FP

    src = torch.zeros(100, device=cuda)
    dst = torch.zeros(100, device=cuda)
    iv = torch.tensor([[0, 10]], device=cuda)
    with pytest.raises(RuntimeError):
        FP.move_intervals(src, dst.half(), iv, iv, 10)
    with pytest.raises(RuntimeError):
        FP.move_intervals(src, dst, iv.int(), iv, 10)
    with pytest.raises(NotImplementedError):
        FP.move_intervals(src.double(), dst.double(), iv, iv, 10)
    FP.move_intervals(src, dst, iv[:0], iv[:0], 0)  # nothing to do
    # workspace too small -> error code, nothing launched
    ws = _lib.workspace(256, cuda)
    rc = _lib.lib.b200rl_move_intervals(_lib.ptr(src), 100, _lib.ptr(dst), 100, _lib.ptr(iv), 1, _lib.ptr(iv), 1,
                                        10, 4, _lib.ptr(ws), 16, 0, _lib.stream_handle(cuda))
    assert rc == _lib.EWORKSPACE if hasattr(_lib, "EWORKSPACE") else rc == -3
    # lists that disagree on the total copy only the common prefix and never write out of range
    src = torch.arange(100, device=cuda, dtype=torch.float32)
    dst = torch.full((100,), -1.0, device=cuda)
    FP.move_intervals(src, dst, torch.tensor([[0, 50]], device=cuda), torch.tensor([[10, 40]], device=cuda), 50)
    assert torch.equal(dst[10:40], src[0:30]) and float(dst[40]) == -1.0 and float(dst[9]) == -1.0
